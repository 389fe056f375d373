#!/usr/bin/env python
"""In-process experiment grid on one workload: (trace variant, fused-tail start, lanes, spp) -> Mrays/s of whole frames
(CUDA events, 3 warm-up + 5 timed frames each) and, per configuration, the per-kernel event shares of one profiled frame.
  python tools/exp_grid.py C3 "variant,tail,overlap,spp;..."   e.g.  "20,-1,1,64;21,-1,1,64;21,2,1,64"
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("smallpt-enoki-optix_b200",):
    sys.path.insert(0, os.path.join(ROOT, p))
import scenes
import wpt

name = sys.argv[1] if len(sys.argv) > 1 else "C3"
grid = [tuple(int(x) for x in g.split(",")) for g in (sys.argv[2] if len(sys.argv) > 2 else "0,0,1,64").split(";") if g]
gen, W, H, spp0, depth = scenes.CONFIGS[name]
sd = gen()
builder = sys.argv[3] if len(sys.argv) > 3 else "sah"
scene = wpt.scene_from(sd, builder=builder)
print(json.dumps({"builder": builder, **{k: scene.stats()[k] for k in ("n_nodes", "max_depth", "build_ms", "sah_cost")}}), flush=True)
cam = wpt.camera_from(sd)
N = W * H
film = torch.zeros(3, N, device="cuda")
renderers = {}
for variant, tail, overlap, spp in grid:
    per_pass = int(max(1, min(spp, 64, (1 << 28) // N)))
    if per_pass not in renderers:
        renderers[per_pass] = wpt.Renderer(scene, W, H, per_pass)
    r = renderers[per_pass]
    r.set_tuning(variant, 0, overlap)
    r.set_tail(tail)
    for _ in range(3):
        film.zero_(); r.render(cam, film, 0, spp, depth, seed=0x5EED); st = r.finish()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    segs = 0
    for _ in range(5):
        film.zero_(); r.render(cam, film, 0, spp, depth, seed=0x5EED); segs += r.finish()["segments"]
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    r.set_tuning(variant, 0, 0)
    r.set_profile(True)
    film.zero_(); r.render(cam, film, 0, spp, depth, seed=0x5EED); p = r.finish()
    r.set_profile(False)
    print(json.dumps({"workload": name, "variant": variant, "tail": tail, "overlap": overlap, "spp": spp,
                      "mrays_s": round(segs / ms / 1e3, 1), "ms_per_frame": round(ms / 5, 3),
                      "serial_ms": round(p["ms_total"], 3), "trace_ms": round(p["ms_trace"], 3), "shade_ms": round(p["ms_shade"], 3),
                      "raygen_ms": round(p["ms_raygen"], 3), "trace_only_mrays_s": round(p["segments"] / p["ms_trace"] / 1e3, 1),
                      "launches": p["kernel_launches"], "film_sum": float(film.sum())}), flush=True)

#!/usr/bin/env python
"""One serial C3 frame (16 spp in one pass, no fused tail) for the DRAM-traffic capture of the trace kernel
(tools/gpu_r2_call7.sh): prints the rays per launch so that dram bytes per ray can be formed from ncu's per-launch bytes."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "smallpt-enoki-optix_b200"))
import scenes, wpt
gen, W, H, spp, depth = scenes.CONFIGS["C3"]
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 16
sd = gen()
scene, cam = wpt.scene_from(sd), wpt.camera_from(sd)
r = wpt.Renderer(scene, W, H, spp)
r.set_tuning(0, 0, 0); r.set_tail(-1)
film = torch.zeros(3, W * H, device="cuda")
for _ in range(2):
    film.zero_(); r.render(cam, film, 0, spp, depth, seed=0x5EED); st = r.finish()
print(json.dumps({"spp": spp, "live": st["live"][:depth], "segments": st["segments"], "retraced": st["retraced"], "pixels": W * H}))

#!/usr/bin/env python
"""Workloads for tools/sanitize.sh (compute-sanitizer racecheck / memcheck): small on purpose — the tools slow kernels
by 10-100x.  (1) a C1-like render through every kernel of the pipeline: depth pre-pass + seeded traces + retry, wavefront
bounces, fused tail, both lanes; (2) the deep-tree spiral scene with the two-row shared stack (local-memory overflow,
trace variant 12) and the default stack; (3) a masked wpt_intersect with per-ray windows; (4) the per-entry flush."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("smallpt-enoki-optix_b200", "tests", "oracle"):
    sys.path.insert(0, os.path.join(ROOT, p))
import scenes
import wpt

which = sys.argv[1] if len(sys.argv) > 1 else "all"
torch.manual_seed(0)
if which in ("all", "render"):
    sd = scenes.cornell(12, colored=True)
    scene, cam = wpt.scene_from(sd), wpt.camera_from(sd)
    for overlap, tail, variant in ((0, 0, 0), (1, 1, 0), (0, -1, 20)):
        r = wpt.Renderer(scene, 48, 40, 4)
        r.set_tuning(variant, 2, overlap)
        r.set_tail(tail)
        f = torch.zeros(3, 48 * 40, device="cuda")
        r.render(cam, f, 0, 8, 5, seed=1)
        st = r.finish()
        print("render", overlap, tail, variant, st["segments"], round(float(f.sum()), -1))
        r.close()
if which in ("all", "deep"):
    # 13 node levels: the two-row shared stack of variant 12 overflows to local memory on most rays
    sd = scenes.spiral(60, 0.7)
    scene = wpt.scene_from(sd)
    cam = wpt.ThinlensCamera((6.0, 0.45, 0.0), (0.0, 0.0, 0.0), (0, 1, 0), 0.0, 1.0, 8.0 / 180.0 * np.pi)   # down the axis of the spiral: every ray passes all 60 spheres, the deep levels included
    assert scene.stats()["max_depth"] > 12
    for variant, tail in ((12, -1), (12, 1), (0, 0)):
        r = wpt.Renderer(scene, 64, 32, 2)
        r.set_tuning(variant, 2, 0)
        r.set_tail(tail)
        f = torch.zeros(3, 64 * 32, device="cuda")
        r.render(cam, f, 0, 2, 4, seed=2)
        st = r.finish()
        print("deep", variant, tail, st["segments"], round(float(f.sum()), -1))
        r.close()
if which in ("all", "masked"):
    sd = scenes.cornell(12)
    scene = wpt.scene_from(sd)
    rng = np.random.default_rng(2)
    n = 5000
    o = rng.uniform(-1, 1, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    ot = [torch.from_numpy(np.ascontiguousarray(o[:, k])).cuda() for k in range(3)]
    dt = [torch.from_numpy(np.ascontiguousarray(d[:, k])).cuda() for k in range(3)]
    m = torch.from_numpy((np.arange(n) % 3 != 0).astype(np.uint8)).cuda()
    tmin = torch.rand(n, device="cuda") * 0.5
    tmax = tmin + 2.0
    h = scene.intersect(ot, dt, active=m, tmin=tmin, tmax=tmax)
    att = scene.hit_attributes(ot, dt, h)
    torch.cuda.synchronize()
    print("masked", int((h[0] >= 0).sum()), float(torch.nan_to_num(att["p"]).sum()))
if which in ("all", "c3"):
    # one full-size bounce mix: C3 at 2 spp (4 M camera rays, every bounce, fused tail on and off)
    gen, W, H, spp, depth = scenes.CONFIGS["C3"]
    sd = gen()
    scene, cam = wpt.scene_from(sd), wpt.camera_from(sd)
    for tail in (-1, 2):
        r = wpt.Renderer(scene, W, H, 2)
        r.set_tail(tail)
        f = torch.zeros(3, W * H, device="cuda")
        r.render(cam, f, 0, 2, depth, seed=3)
        st = r.finish()
        print("c3", tail, st["segments"], round(float(f.sum()), -1))
        r.close()
# self-checking build (WPT_LIB=libwpt_checked.so): violated kernel invariants per kind, all must be zero
import ctypes as C
out = (C.c_uint32 * 8)()
rc = wpt.lib().wpt_debug_check_failures(out, C.c_int(0))
print("checks", "unavailable" if rc != 0 else list(out))

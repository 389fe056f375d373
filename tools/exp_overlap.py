#!/usr/bin/env python
"""does an HBM-bound stream kernel overlap the persistent trace kernel?  (C3 bounce-1-like rays)
usage: [WPT_INTERSECT_CTAS=k] python tools/exp_overlap.py"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("smallpt-enoki-optix_b200", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import numpy as np, torch, scenes, wpt
sd = scenes.CONFIGS["C3"][0]()
scene = wpt.scene_from(sd); cam = wpt.camera_from(sd)
W, H = 1920, 1080
o, d = wpt.raygen(cam, W, H, 0, seed=1)
hit = scene.intersect(o, d)
live = hit[0] >= 0
idx = live.nonzero().squeeze(1)
n = idx.numel()
s0 = torch.rand(n, device="cuda"); s1 = torch.rand(n, device="cuda")
sel = lambda ts: [t[idx].contiguous() for t in ts]
no, nd, w = scene.shade_step(sel(o), sel(d), hit[0][idx].contiguous(), hit[1][idx].contiguous(), hit[2][idx].contiguous(), hit[3][idx].contiguous(), s0, s1)
# 8 copies of the bounce rays -> ~15 M incoherent rays per launch
no = [t.repeat(8) for t in no]; nd = [t.repeat(8) for t in nd]
nr = no[0].numel()
out = scene.intersect(no, nd)
a = torch.empty(1 << 28, dtype=torch.float32, device="cuda"); b = torch.empty_like(a)   # 1 GiB each
st1, st2 = torch.cuda.Stream(), torch.cuda.Stream()
def timed(fn, reps=5):
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / reps
def trace(): scene.intersect(no, nd, out=out)
def copy(): b.copy_(a)
def both(order):
    ev = torch.cuda.Event(); ev.record()
    st1.wait_event(ev); st2.wait_event(ev)
    def t():
        with torch.cuda.stream(st1): scene.intersect(no, nd, out=out, stream=st1)
    def c():
        with torch.cuda.stream(st2):
            for _ in range(K): b.copy_(a)
    (t(), c()) if order == 0 else (c(), t())
    e1 = torch.cuda.Event(); e2 = torch.cuda.Event(); e1.record(st1); e2.record(st2)
    torch.cuda.current_stream().wait_event(e1); torch.cuda.current_stream().wait_event(e2)
for _ in range(3): trace(); copy()
T = timed(trace); Cc = timed(copy)
K = max(1, int(round(0.3 * T / Cc)))     # stream work ~30 % of the trace time
res = {"rays": nr, "trace_ms": T, "copy_ms_each": Cc, "copies": K, "ctas": os.environ.get("WPT_INTERSECT_CTAS", "auto")}
for order in (0, 1):
    for _ in range(2): both(order)
    res[f"both_order{order}_ms"] = timed(lambda: both(order))
res["serial_ms"] = T + K * Cc
print(json.dumps(res))

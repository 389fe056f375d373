"""Experiment (GPU): how much would re-ordering bounce rays buy the trace kernel?
Builds bounce-0/1/2 ray sets of C3 through the per-function entry points, then times wpt_intersect on
(a) the wavefront's natural order, (b) a random shuffle, (c) sorted by direction octant,
(d) sorted by a Morton code of the origin, (e) octant + Morton."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "smallpt-enoki-optix_b200"))
import numpy as np, torch
import scenes, wpt

def morton(o, lo, hi, bits=10):
    q = [((x - l) / (h - l) * ((1 << bits) - 1)).clamp(0, (1 << bits) - 1).long() for x, l, h in zip(o, lo, hi)]
    def spread(v):
        v = (v | (v << 16)) & 0x030000FF
        v = (v | (v << 8)) & 0x0300F00F
        v = (v | (v << 4)) & 0x030C30C3
        v = (v | (v << 2)) & 0x09249249
        return v
    return spread(q[0]) | (spread(q[1]) << 1) | (spread(q[2]) << 2)

def timeit(scene, o, d, reps=5):
    out = scene.intersect(o, d)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        scene.intersect(o, d, out=out)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out

def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "C3"
    gen, W, H, spp, depth = scenes.CONFIGS[wl]
    sd = gen()
    scene = wpt.scene_from(sd, 0)
    cam = wpt.camera_from(sd)
    N = W * H
    os_, ds_ = [], []
    for s in range(4):
        o, d = wpt.raygen(cam, W, H, s, seed=1)
        os_.append(o); ds_.append(d)
    o = [torch.cat([x[k] for x in os_]) for k in range(3)]
    d = [torch.cat([x[k] for x in ds_]) for k in range(3)]
    res = {}
    for bounce in range(4):
        n = o[0].numel()
        ms, hit = timeit(scene, o, d)
        row = {"n": n, "natural": n / ms / 1e3}
        p = torch.randperm(n, device="cuda")
        row["shuffled"] = n / timeit(scene, [x[p] for x in o], [x[p] for x in d])[0] / 1e3
        octant = (d[0] >= 0).long() | ((d[1] >= 0).long() << 1) | ((d[2] >= 0).long() << 2)
        lo = [float(x.min()) for x in o]; hi = [float(x.max()) + 1e-6 for x in o]
        mc = morton(o, lo, hi)
        for name, key in (("octant", octant), ("morton", mc), ("octant_morton", (octant << 30) | mc),
                          ("morton_octant", (mc << 3) | octant),
                          ("morton15_octant", ((mc >> 15) << 3) | octant)):
            p = torch.argsort(key, stable=True)
            row[name] = n / timeit(scene, [x[p] for x in o], [x[p] for x in d])[0] / 1e3
        res[f"bounce{bounce}"] = row
        print(bounce, json.dumps(row), flush=True)
        # next bounce: shade the hits
        tri, t, u, v = hit
        keep = tri >= 0
        idx = keep.nonzero().squeeze(1)
        o2 = [x[idx].contiguous() for x in o]; d2 = [x[idx].contiguous() for x in d]
        g = torch.Generator(device="cuda"); g.manual_seed(bounce)
        s0 = torch.rand(idx.numel(), device="cuda", generator=g); s1 = torch.rand(idx.numel(), device="cuda", generator=g)
        no, nd, w = scene.shade_step(o2, d2, tri[idx].contiguous(), t[idx].contiguous(), u[idx].contiguous(), v[idx].contiguous(), s0, s1)
        o, d = no, nd
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", f"exp_sort_{wl}.json"), "w"), indent=1)

main()

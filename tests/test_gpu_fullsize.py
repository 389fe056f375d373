"""-m gpu: BASELINE.json's full-size configurations.  The CPU oracle cannot render these in seconds, so
the checks are (a) closest hits of 2^20 camera + bounce rays against the reference tree's Embree
(oracle/_ref) on the SAME 1 M / 10 M triangle scene, and (b) size-independent properties of whole frames:
every kernel variant produces the same film, disjoint sample ranges add up, each pixel is a mean of
escape indicators (albedo 1 => 0 <= film <= 1, film * spp sums to the number of escaped paths the
device counted), and the segment counter equals the sum of the per-bounce queue sizes."""
import numpy as np
import pytest

import oracle_py as O
import scenes
import util

pytestmark = pytest.mark.gpu


def _rays_vs_embree(torch, sd, scene, W, H, what, n_max=1 << 20):
    import wpt
    cam = wpt.camera_from(sd)
    om = util.oracle_mesh(sd)
    backend = "embree" if O.ref_available() else "bvh"
    o_t, d_t = wpt.raygen(cam, W, H, 0, seed=5)
    step = max(1, (W * H) // n_max)
    pick = torch.arange(0, W * H, step, device="cuda")
    o_t = [x[pick].contiguous() for x in o_t]
    d_t = [x[pick].contiguous() for x in d_t]
    hit = scene.intersect(o_t, d_t)
    torch.cuda.synchronize()
    o = np.stack([x.cpu().numpy() for x in o_t], 1)
    d = np.stack([x.cpu().numpy() for x in d_t], 1)
    got = tuple(x.cpu().numpy() for x in hit)
    ref = om.intersect(backend, o, d)
    assert util.compare_hits(ref, got, what + "/camera") >= 0.9999
    live = torch.nonzero(hit[0] >= 0).squeeze(1)
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    s0 = torch.rand(live.numel(), device="cuda", generator=g)
    s1 = torch.rand(live.numel(), device="cuda", generator=g)
    sel = lambda ts: [t_[live].contiguous() for t_ in ts]
    no, nd, w = scene.shade_step(sel(o_t), sel(d_t), *[h[live].contiguous() for h in hit], s0, s1)
    hit2 = scene.intersect(no, nd)
    torch.cuda.synchronize()
    o2 = np.stack([x.cpu().numpy() for x in no], 1)
    d2 = np.stack([x.cpu().numpy() for x in nd], 1)
    ref2 = om.intersect(backend, o2, d2)
    assert util.compare_hits(ref2, tuple(x.cpu().numpy() for x in hit2), what + "/bounce") >= 0.9999
    return live.numel()


def test_c3_one_million_triangles_vs_embree():
    import torch
    import wpt
    gen, W, H, spp, depth = scenes.CONFIGS["C3"]
    sd = gen()
    assert sd.n_tris == 1007232
    scene = wpt.scene_from(sd)
    assert _rays_vs_embree(torch, sd, scene, W, H, "C3") > 100000


def test_c3_full_frame_properties():
    import torch
    import wpt
    gen, W, H, spp, depth = scenes.CONFIGS["C3"]
    sd = gen()
    scene = wpt.scene_from(sd)
    cam = wpt.camera_from(sd)
    N = W * H
    spp = 8                                   # the properties do not depend on the sample count
    films, stats = [], []
    for variant, per_pass in ((0, 8), (1, 8), (0, 3)):
        r = wpt.Renderer(scene, W, H, per_pass)
        r.set_tuning(trace_variant=variant)
        f = torch.zeros(3, N, device="cuda")
        r.render(cam, f, 0, spp, depth, seed=21)
        stats.append(r.finish())
        films.append(f)
        r.close()
    # same paths whatever the kernel variant or the pass split: identical segment counts; films equal up to
    # the order of the float atomics (sums of a few 1.0s: exact)
    assert stats[0]["segments"] == stats[1]["segments"] == stats[2]["segments"]
    assert torch.equal(films[0], films[1]) and torch.equal(films[0], films[2])
    st = stats[0]
    live = st["live"][:depth]
    assert live[0] == N * spp and all(a >= b for a, b in zip(live, live[1:])) and sum(live) == st["segments"]
    f = films[0]
    assert float(f.min()) >= 0.0 and float(f.max()) <= spp            # albedo 1: a pixel counts its escaped paths
    assert torch.equal(f[0], f[1]) and torch.equal(f[0], f[2])         # white materials: grey image
    assert torch.equal(f[0], f[0].round())                             # ... of whole numbers before the 1/spp scale
    escaped_by_depth = N * spp - live[-1]                              # paths that left before the last segment
    # every escaped path adds exactly 1; the last segment's escapes are the only unknown
    assert escaped_by_depth <= float(f[0].sum()) <= N * spp
    # disjoint sample ranges add up
    r = wpt.Renderer(scene, W, H, 8)
    a = torch.zeros(3, N, device="cuda"); b = torch.zeros(3, N, device="cuda")
    r.render(cam, a, 0, 5, depth, seed=21); r.finish()
    r.render(cam, b, 5, 3, depth, seed=21); r.finish()
    assert torch.equal(a + b, f)


def test_c4_ten_million_triangles():
    """configs[3] geometry on one GPU: 10 M flattened triangles.  Closest hits vs Embree on a 2^18 ray sample,
    then the per-thread kernel against the persistent warp-queue kernel on a 4K frame."""
    import torch
    import wpt
    gen, W, H, spp, depth = scenes.CONFIGS["C4"]
    sd = gen()
    assert sd.n_tris >= 10_000_000
    scene = wpt.scene_from(sd)
    st = scene.stats()
    assert st["n_triangles"] == sd.n_tris and st["max_depth"] <= 32
    _rays_vs_embree(torch, sd, scene, W, H, "C4", n_max=1 << 18)
    cam = wpt.camera_from(sd)
    films = []
    for variant in (0, 1):
        r = wpt.Renderer(scene, W, H, 1)
        r.set_tuning(trace_variant=variant)
        f = torch.zeros(3, W * H, device="cuda")
        r.render(cam, f, 0, 1, depth, seed=2)
        s = r.finish()
        films.append((f, s["segments"]))
        r.close()
    assert films[0][1] == films[1][1] and torch.equal(films[0][0], films[1][0])

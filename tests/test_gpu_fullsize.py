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
    ref, robust = util.embree_pair(om, o, d)
    assert util.compare_hits(ref, got, what + "/camera", robust=robust) >= 0.9999
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
    ref2, robust2 = util.embree_pair(om, o2, d2)
    assert util.compare_hits(ref2, tuple(x.cpu().numpy() for x in hit2), what + "/bounce", robust=robust2) >= 0.9999
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
    # and the 3-sigma image of the 10 M-triangle scene against the oracle on Embree (1080p, 8 spp: seconds on the host)
    _three_sigma_image(sd, scene, 1920, 1080, 8, depth, "C4")


# ---------------------------------------------------------------- round 2: every BASELINE config at its configured size
def _chain_vs_embree(torch, sd, scene, W, H, what, check_bounces, want=1 << 20, max_spp=24):
    """Hit parity on rays of EVERY listed bounce, at the configured resolution: camera rays of successive global sample
    indices (wpt_raygen) are traced and bounced with the product's own per-function entry points (wpt_intersect /
    wpt_shade_step — the code the render kernels fuse), so the rays entering bounce k have the distribution of the
    renderer's own bounce-k queue; each listed bounce is compared with the reference tree's Embree (fast + robust
    intersector, oracle/_ref) until at least `want` rays of it have been checked."""
    import wpt
    cam = wpt.camera_from(sd)
    om = util.oracle_mesh(sd)
    done = {b: 0 for b in check_bounces}
    g = torch.Generator(device="cuda"); g.manual_seed(7)
    for s in range(max_spp):
        if all(v >= want for v in done.values()):
            break
        o_t, d_t = wpt.raygen(cam, W, H, s, seed=11)
        for b in range(max(check_bounces) + 1):
            hit = scene.intersect(o_t, d_t)
            torch.cuda.synchronize()
            if b in done and done[b] < want:
                o = np.stack([x.cpu().numpy() for x in o_t], 1)
                d = np.stack([x.cpu().numpy() for x in d_t], 1)
                ref, robust = util.embree_pair(om, o, d)
                rate = util.compare_hits(ref, tuple(x.cpu().numpy() for x in hit), f"{what}/bounce {b}", robust=robust)
                assert rate >= 0.9999
                done[b] += o.shape[0]
            live = torch.nonzero(hit[0] >= 0).squeeze(1)
            if live.numel() == 0:
                break
            sel = lambda ts: [t_[live].contiguous() for t_ in ts]
            s0 = torch.rand(live.numel(), device="cuda", generator=g)
            s1 = torch.rand(live.numel(), device="cuda", generator=g)
            o_t, d_t, _ = scene.shade_step(sel(o_t), sel(d_t), *[h[live].contiguous() for h in hit], s0, s1)
    return done


def _three_sigma_image(sd, scene, W, H, spp, depth, what):
    """north_star image bar at a BASELINE config: the device frame (Philox) against the oracle loop on the reference
    tree's Embree with the REFERENCE's generator (per-pixel PCG32, src/main.cpp:376) at EQUAL spp — every pixel, all
    three channels, 3 sigma of the difference of two independent means."""
    import wpt
    cam, ocam = wpt.camera_from(sd), util.oracle_camera(sd)
    r = wpt.Renderer(scene, W, H, min(spp, 16))
    film, st = r.render_host(cam, 0, spp, depth, seed=0x5EED)
    r.close()
    om = util.oracle_mesh(sd)
    ofilm, oseg = om.render("embree" if O.ref_available() else "bvh", ocam, W, H, 0, spp, depth, rng_mode=O.RNG_PCG32)
    ofilm /= spp
    frac, z2 = util.three_sigma_outliers(film, ofilm, spp)
    assert frac < 0.005, f"{what}: {frac:.5f} of the pixel values lie outside 3 sigma (Gaussian: 0.0027)"
    assert z2 < 1.15, f"{what}: mean squared z-score {z2:.3f} (a correct renderer: <= 1 + O(1/spp))"
    assert abs(film.mean() - ofilm.mean()) <= 1e-2 * max(ofilm.mean(), 1e-3), (what, film.mean(), ofilm.mean())
    assert abs(st["segments"] - oseg) <= 2e-3 * oseg, (what, st["segments"], oseg)
    return frac, z2


def test_c2_full_size_hits_and_image():
    """configs[1] as configured: knob(500, 200) = 200 K triangles, 1024 x 1024, 16 spp, depth 8"""
    import torch
    import wpt
    gen, W, H, spp, depth = scenes.CONFIGS["C2"]
    sd = gen()
    assert (W, H, spp, depth) == (1024, 1024, 16, 8) and sd.n_tris >= 199_000
    scene = wpt.scene_from(sd)
    done = _chain_vs_embree(torch, sd, scene, W, H, "C2", check_bounces=(0, 3), want=1 << 20)
    assert done[0] >= 1 << 20 and done[3] >= 1 << 20, done
    _three_sigma_image(sd, scene, W, H, spp, depth, "C2")


def test_c3_three_sigma_image_at_the_headline_config():
    """configs[2]: the 1 007 232-triangle scene at 1920 x 1080, depth 8; 16 of its 64 spp on both sides (equal spp)"""
    import wpt
    gen, W, H, spp, depth = scenes.CONFIGS["C3"]
    sd = gen()
    scene = wpt.scene_from(sd)
    _three_sigma_image(sd, scene, W, H, 16, depth, "C3")


def test_c5_full_size_hits_and_image():
    """configs[4] as configured: interior() = 1.12 M triangles, 1920 x 1080, depth 16 — the incoherent stress case.
    Hit parity on >= 2^20 camera rays and >= 2^20 rays each of bounce 3 and bounce 8; the 3-sigma image at 480 x 270,
    128 spp, depth 16 (the oracle renders that in seconds)."""
    import torch
    import wpt
    gen, W, H, spp, depth = scenes.CONFIGS["C5"]
    sd = gen()
    assert (W, H, spp, depth) == (1920, 1080, 128, 16) and sd.n_tris > 1_000_000
    scene = wpt.scene_from(sd)
    done = _chain_vs_embree(torch, sd, scene, W, H, "C5", check_bounces=(0, 3, 8), want=1 << 20, max_spp=3)
    assert all(v >= 1 << 20 for v in done.values()), done
    _three_sigma_image(sd, scene, 480, 270, spp, depth, "C5")

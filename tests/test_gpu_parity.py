"""-m gpu parity tests proper: the CUDA path through the C ABI (libwpt.so) against the CPU oracle
and, where oracle/_ref is present, against the real Embree 3.6.1 of the reference tree."""
import numpy as np
import pytest

import oracle_py as O
import scenes
import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available()
    return torch


@pytest.fixture(scope="module")
def cornell(torch_cuda):
    import wpt
    sd = scenes.cornell(50)
    return sd, wpt.scene_from(sd), util.oracle_mesh(sd)


def _dev_intersect(torch, scene, o, d, **kw):
    ot = [torch.from_numpy(np.ascontiguousarray(o[:, k])).cuda() for k in range(3)]
    dt = [torch.from_numpy(np.ascontiguousarray(d[:, k])).cuda() for k in range(3)]
    out = scene.intersect(ot, dt, **kw)
    torch.cuda.synchronize()
    return tuple(x.cpu().numpy() for x in out)


def test_minimal_kat(torch_cuda):
    """Embree tutorials/minimal/minimal.cpp:72-213: unit triangle, hit at t=1 prim 0; second ray misses."""
    import wpt
    s = wpt.Scene()
    s.add_triangle_mesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]])
    s.commit(0)
    o = np.array([[0, 0, -1], [1, 1, -1]], np.float32)
    d = np.array([[0, 0, 1], [0, 0, 1]], np.float32)
    hid, t, u, v = _dev_intersect(torch_cuda, s, o, d, tmin=0.0, tmax=float("inf"))
    assert hid.tolist() == [0, -1]
    assert abs(t[0] - 1.0) < 1e-6 and u[0] == 0 and v[0] == 0
    assert np.isnan(t[1])  # t,u,v untouched on miss


def test_triangle_hit_kat(torch_cuda):
    """Embree verify.cpp:2312-2399 TriangleHitTest: 256 rays from (0,0,-1) at the unit triangle:
    u,v,t within 16 ulp of the analytic values."""
    import wpt
    s = wpt.Scene()
    s.add_triangle_mesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]])
    s.commit(0)
    g = np.arange(16, dtype=np.float32)
    U, V = np.meshgrid(g, g, indexing="ij")
    u0 = (U.reshape(-1) / 32.0 + 1 / 64.0).astype(np.float32)
    v0 = (V.reshape(-1) / 32.0 + 1 / 64.0).astype(np.float32)
    o = np.tile(np.array([[0, 0, -1]], np.float32), (256, 1))
    d = np.stack([u0, v0, np.ones_like(u0)], 1).astype(np.float32)
    hid, t, u, v = _dev_intersect(torch_cuda, s, o, d, tmin=0.0, tmax=float("inf"))
    assert (hid == 0).all()
    eps = 16 * np.finfo(np.float32).eps
    assert np.abs(u - u0).max() <= eps and np.abs(v - v0).max() <= eps and np.abs(t - 1.0).max() <= eps


@pytest.mark.parametrize("inside", [True, False])
def test_intersect_vs_oracle_cornell(torch_cuda, cornell, inside):
    sd, scene, om = cornell
    o, d = util.random_rays(sd, 400_000, seed=11 + inside, inside=inside)
    got = _dev_intersect(torch_cuda, scene, o, d)
    ref = om.intersect("bvh", o, d)
    rate = util.compare_hits(ref, got, "cornell/oracle")
    assert rate >= 0.9999
    if O.ref_available():
        fast, robust = util.embree_pair(om, o, d)
        util.compare_hits(fast, got, "cornell/embree", robust=robust)


def test_intersect_mask_semantics(torch_cuda, cornell):
    """wavefront_isect.cu:86-87: size-1 mask broadcasts, masked-out lanes are left untouched."""
    torch = torch_cuda
    sd, scene, om = cornell
    o, d = util.random_rays(sd, 10_000, seed=5)
    full = _dev_intersect(torch, scene, o, d)
    m = (np.arange(10_000) % 3 != 0).astype(np.uint8)
    part = _dev_intersect(torch, scene, o, d, active=torch.from_numpy(m).cuda())
    assert (part[0][m == 0] == -2).all() and np.isnan(part[1][m == 0]).all()
    assert (part[0][m == 1] == full[0][m == 1]).all()
    one = _dev_intersect(torch, scene, o, d, active=torch.ones(1, dtype=torch.uint8, device="cuda"))
    assert (one[0] == full[0]).all()
    zero = _dev_intersect(torch, scene, o, d, active=torch.zeros(1, dtype=torch.uint8, device="cuda"))
    assert (zero[0] == -2).all()
    bmask = _dev_intersect(torch, scene, o, d, active=torch.from_numpy(m.astype(bool)).cuda())
    assert (bmask[0] == part[0]).all()


def test_intersect_host_and_empty(torch_cuda, cornell):
    sd, scene, om = cornell
    o, d = util.random_rays(sd, 5_000, seed=9)
    a = scene.intersect_host(o, d)
    b = _dev_intersect(torch_cuda, scene, o, d)
    assert (a[0] == b[0]).all() and np.array_equal(a[1], b[1], equal_nan=True)
    e = scene.intersect_host(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32))
    assert e[0].size == 0


def test_intersect_tmin_tmax(torch_cuda, cornell):
    """tmin exclusive, tmax inclusive-ish window: results must equal the oracle's under the same window."""
    sd, scene, om = cornell
    o, d = util.random_rays(sd, 100_000, seed=21)
    for tmin, tmax in ((0.0, 0.5), (0.3, 1.0), (1e-3, 1e20)):
        got = _dev_intersect(torch_cuda, scene, o, d, tmin=tmin, tmax=tmax)
        ref = om.intersect("bvh", o, d, tmin=tmin, tmax=tmax)
        util.compare_hits(ref, got, f"window {tmin},{tmax}")


def test_intersect_knob_camera_and_bounce_rays(torch_cuda):
    """C2-like scene: coherent camera rays and incoherent second-bounce rays vs oracle (+Embree)."""
    import wpt
    torch = torch_cuda
    sd = scenes.knob(250, 100)
    scene, om = wpt.scene_from(sd), util.oracle_mesh(sd)
    cam = wpt.camera_from(sd)
    W = H = 512
    o_t, d_t = wpt.raygen(cam, W, H, 0, seed=7)
    hit = scene.intersect(o_t, d_t)
    torch.cuda.synchronize()
    o = np.stack([x.cpu().numpy() for x in o_t], 1)
    d = np.stack([x.cpu().numpy() for x in d_t], 1)
    got = tuple(x.cpu().numpy() for x in hit)
    ref, robust = util.embree_pair(om, o, d)
    assert util.compare_hits(ref, got, "knob/camera", robust=robust) >= 0.9999
    # bounce rays from the hit points
    live = got[0] >= 0
    n = int(live.sum())
    s0 = torch.rand(n, device="cuda")
    s1 = torch.rand(n, device="cuda")
    idx = torch.from_numpy(np.nonzero(live)[0]).cuda()
    sel = lambda ts: [t_[idx].contiguous() for t_ in ts]
    no, nd, w = scene.shade_step(sel(o_t), sel(d_t), hit[0][idx].contiguous(), hit[1][idx].contiguous(),
                                 hit[2][idx].contiguous(), hit[3][idx].contiguous(), s0, s1)
    hit2 = scene.intersect(no, nd)
    torch.cuda.synchronize()
    o2 = np.stack([x.cpu().numpy() for x in no], 1)
    d2 = np.stack([x.cpu().numpy() for x in nd], 1)
    got2 = tuple(x.cpu().numpy() for x in hit2)
    ref2, robust2 = util.embree_pair(om, o2, d2)
    assert util.compare_hits(ref2, got2, "knob/bounce", robust=robust2) >= 0.9999


def test_philox_kat(torch_cuda):
    """Random123 known-answer vectors for philox4x32-10 + device == oracle on a sweep."""
    import ctypes as C
    import wpt
    out = wpt.philox(1, 0, 0, 0).cpu().numpy().view(np.uint32)
    assert [hex(x) for x in out[0]] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    L = O.lib()
    dev = wpt.philox(1000, 0xDEAD, 7, 0x0123456789ABCDEF).cpu().numpy().view(np.uint32)
    for i in (0, 1, 17, 999):
        ctr = (C.c_uint32 * 4)(i, 0xDEAD, 7, 0)
        key = (C.c_uint32 * 2)(0x89ABCDEF, 0x01234567)
        o = (C.c_uint32 * 4)()
        L.orc_philox4x32_10(ctr, key, o)
        assert list(o) == dev[i].tolist()


def test_raygen_vs_oracle(torch_cuda):
    """camera rays (src/pinhole.h:27-56) device vs oracle with the same Philox numbers: <= 1e-6 abs."""
    import ctypes as C
    import wpt
    sd = scenes.cornell(8)
    cam = wpt.camera_from(sd)
    ocam = util.oracle_camera(sd)
    W, H = 64, 48
    o_t, d_t = wpt.raygen(cam, W, H, 3, seed=99)
    d = np.stack([x.cpu().numpy() for x in d_t], 1)
    o = np.stack([x.cpu().numpy() for x in o_t], 1)
    L = O.lib()
    for pix in (0, 1, 63, 64, 1000, W * H - 1):
        ctr = (C.c_uint32 * 4)(pix, 3, 0, 0)
        key = (C.c_uint32 * 2)(99, 0)
        r = (C.c_uint32 * 4)()
        L.orc_philox4x32_10(ctr, key, r)
        xi = (C.c_float * 4)(*[L.orc_u32_to_f32(C.c_uint32(x)) for x in r])
        po, pd = (C.c_float * 3)(), (C.c_float * 3)()
        L.orc_camera_sample_pos(C.byref(ocam), xi, po)
        xi2 = (C.c_float * 2)(xi[2], xi[3])
        L.orc_camera_sample_dir(C.byref(ocam), C.c_int(pix % W), C.c_int(pix // W), po, C.c_int(W), C.c_int(H), xi2, pd)
        assert np.abs(np.array(po) - o[pix]).max() <= 1e-6
        assert np.abs(np.array(pd) - d[pix]).max() <= 2e-6


def test_shade_step_vs_oracle(torch_cuda, cornell):
    """one bounce (optix_backend.h:463-486 + main.cpp:414-424) device vs oracle on the same hits/uniforms"""
    torch = torch_cuda
    sd, scene, om = cornell
    o, d = util.random_rays(sd, 2000, seed=3)
    hid, t, u, v = om.intersect("bvh", o, d)
    live = np.nonzero(hid >= 0)[0][:500]
    rng = np.random.default_rng(4)
    s = rng.random((live.size, 2)).astype(np.float32)
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    no, nd, w = scene.shade_step([cu(o[live, k]) for k in range(3)], [cu(d[live, k]) for k in range(3)],
                                 cu(hid[live]), cu(t[live]), cu(u[live]), cu(v[live]), cu(s[:, 0]), cu(s[:, 1]))
    no = np.stack([x.cpu().numpy() for x in no], 1)
    nd = np.stack([x.cpu().numpy() for x in nd], 1)
    w = np.stack([x.cpu().numpy() for x in w], 1)
    for k, i in enumerate(live):
        ro, rd, rw = om.shade_step(o[i], d[i], hid[i], t[i], u[i], v[i], s[k])
        assert np.abs(ro - no[k]).max() <= 1e-6 * max(1.0, np.abs(ro).max())
        assert np.abs(rd - nd[k]).max() <= 2e-6
        assert np.array_equal(rw, w[k])


def _render_both(sd, W, H, spp, depth, seed, rng_mode, spp_per_pass=2):
    import wpt
    scene, om = wpt.scene_from(sd), util.oracle_mesh(sd)
    cam, ocam = wpt.camera_from(sd), util.oracle_camera(sd)
    r = wpt.Renderer(scene, W, H, spp_per_pass)
    film, st = r.render_host(cam, 0, spp, depth, seed=seed)
    ofilm, oseg = om.render("embree" if O.ref_available() else "bvh", ocam, W, H, 0, spp, depth, rng_mode=rng_mode, seed=seed)
    return film, st, ofilm / spp, oseg


def test_render_matches_oracle_same_rng(torch_cuda):
    """C1 (configs[0]): 256x256, 4 spp, depth 5.  With the oracle driven by the same Philox counters the
    two images must be the same paths: identical except where a ulp-level difference flipped a hit."""
    sd = scenes.cornell(50, colored=True)
    film, st, ofilm, oseg = _render_both(sd, 256, 256, 4, 5, seed=1234, rng_mode=O.RNG_PHILOX)
    diff = np.abs(film - ofilm).max(0)
    assert (diff > 1e-5).mean() < 0.01, f"{(diff > 1e-5).mean()} of pixels differ"
    assert abs(film.mean() - ofilm.mean()) < 1e-3 * max(ofilm.mean(), 1e-6)
    assert abs(st["segments"] - oseg) <= 1e-3 * oseg


def test_render_three_sigma_vs_reference_rng(torch_cuda):
    """north_star image bar: 3 sigma per pixel at equal spp against the oracle running the REFERENCE's
    generator (per-pixel PCG32, main.cpp:376)."""
    sd = scenes.cornell(50)
    spp = 64
    film, st, ofilm, oseg = _render_both(sd, 96, 96, spp, 5, seed=77, rng_mode=O.RNG_PCG32, spp_per_pass=8)
    frac, z2 = util.three_sigma_outliers(film, ofilm, spp)           # all three channels
    assert frac < 0.005 and z2 < 1.15, f"{frac} of pixel values outside 3 sigma, mean z^2 {z2}"
    assert abs(film.mean() - ofilm.mean()) < 0.01 * ofilm.mean()
    assert abs(st["segments"] - oseg) < 0.01 * oseg


def test_render_spp_sharding_is_additive(torch_cuda):
    """disjoint spp ranges with the global-index keyed RNG sum to the single-range render."""
    import torch
    import wpt
    sd = scenes.cornell(16)
    scene = wpt.scene_from(sd)
    cam = wpt.camera_from(sd)
    W = H = 64
    r = wpt.Renderer(scene, W, H, 4)
    whole = torch.zeros(3, W * H, device="cuda")
    r.render(cam, whole, 0, 8, 4, seed=5)
    s_all = r.finish()
    parts = torch.zeros(3, W * H, device="cuda")
    r.render(cam, parts, 0, 3, 4, seed=5); s1 = r.finish()
    r.render(cam, parts, 3, 5, 4, seed=5); s2 = r.finish()
    assert torch.equal(whole, parts)          # albedo 1: sums of exact 0/1 contributions
    assert s_all["segments"] == s1["segments"] + s2["segments"]
    # spp_per_pass must not change the result either
    r2 = wpt.Renderer(scene, W, H, 1)
    one = torch.zeros(3, W * H, device="cuda")
    r2.render(cam, one, 0, 8, 4, seed=5); r2.finish()
    assert torch.equal(whole, one)


def test_trace_variants_agree(torch_cuda):
    import torch
    import wpt
    sd = scenes.cornell(16)
    scene = wpt.scene_from(sd)
    cam = wpt.camera_from(sd)
    W = H = 64
    films = []
    for variant in (0, 1, 2, 12, 20, 21, 22, 23):   # 12: two shared stack rows, local-memory overflow exercised; 20: per-entry flush
        r = wpt.Renderer(scene, W, H, 2)
        r.set_tuning(trace_variant=variant)
        f = torch.zeros(3, W * H, device="cuda")
        r.render(cam, f, 0, 4, 5, seed=3); r.finish()
        films.append(f)
    assert all(torch.equal(films[0], f) for f in films[1:])


def test_sealed_room_is_black(torch_cuda):
    """C5 semantics: no emitter — light only enters through escaping rays.  A sealed
    interior is black except for the reference's own leaks (the bounce is sampled about the
    interpolated SHADING normal with no offset, main.cpp:414-424, so on a coarsely tessellated bumpy
    wall a direction can dip under the geometric surface): the device image must equal the oracle's
    path for path, the sealed room must be >= 10x darker than the one with the skylight."""
    import wpt
    means = []
    for sky in (False, True):
        sd = scenes.interior(33, 2, (16, 32), skylight=sky)
        scene = wpt.scene_from(sd)
        r = wpt.Renderer(scene, 64, 64, 4)
        film, st = r.render_host(wpt.camera_from(sd), 0, 8, 6, seed=1)
        ofilm, oseg = util.oracle_mesh(sd).render("bvh", util.oracle_camera(sd), 64, 64, 0, 8, 6,
                                                  rng_mode=O.RNG_PHILOX, seed=1)
        assert (np.abs(film - ofilm / 8).max(0) > 1e-5).mean() < 0.01
        assert abs(st["segments"] - oseg) <= 2e-3 * oseg
        means.append(float(film.mean()))
    assert means[1] > 0 and means[0] < 0.1 * means[1], means


def test_pfm_roundtrip(tmp_path, torch_cuda):
    import wpt
    W, H = 5, 3
    rng = np.random.default_rng(0)
    r, g, b = (rng.random(W * H).astype(np.float32) for _ in range(3))
    p = tmp_path / "x.pfm"
    wpt.Fimage.save_pfm(r, g, b, W, H, p)
    raw = p.read_bytes()
    head = b"PF\n5 3\n-1\n"
    assert raw.startswith(head)
    data = np.frombuffer(raw[len(head):], np.float32).reshape(H, W, 3)
    assert np.array_equal(data[::-1, :, 0].reshape(-1), r)      # rows bottom-up (fimage.h:49)
    assert np.array_equal(data[::-1, :, 2].reshape(-1), b)


def test_errors(torch_cuda):
    import wpt
    s = wpt.Scene()
    with pytest.raises(wpt.WptError):
        s.commit(0)
    with pytest.raises(wpt.WptError):
        s.add_triangle_mesh(np.zeros((0, 3)), np.zeros((0, 3)))
    s.add_triangle_mesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 5]])
    with pytest.raises(wpt.WptError):
        s.commit(0)           # index out of range is caught by the builder


# ---------------------------------------------------------------- row f4: thin lens with lens_radius > 0
def _lens_cameras(sd, radius, focal):
    import wpt
    c = sd.camera
    cam = wpt.ThinlensCamera(c["look_from"], c["look_at"], c["up"], radius, focal, c["fov_y"])
    ocam = util.oracle_camera(sd, lens_radius=radius, focal_dist=focal)
    return cam, ocam


def test_thin_lens_raygen_vs_oracle(torch_cuda):
    """src/pinhole.h:27-32 with lens_radius > 0: Shirley disk (src/mapping.h:15-27) scaled by the radius,
    placed in the camera's local x-z plane (sic), then sample_dir through the focal point (pinhole.h:34-56).
    Every pixel of one sample against the oracle with the same Philox numbers."""
    import ctypes as C
    import wpt
    sd = scenes.cornell(8)
    cam, ocam = _lens_cameras(sd, 0.35, 2.5)
    W, H = 40, 24
    o_t, d_t = wpt.raygen(cam, W, H, 5, seed=4242)
    o = np.stack([x.cpu().numpy() for x in o_t], 1)
    d = np.stack([x.cpu().numpy() for x in d_t], 1)
    L = O.lib()
    on_lens = 0
    for pix in range(W * H):
        ctr = (C.c_uint32 * 4)(pix, 5, 0, 0)
        key = (C.c_uint32 * 2)(4242, 0)
        r = (C.c_uint32 * 4)()
        L.orc_philox4x32_10(ctr, key, r)
        xi = (C.c_float * 4)(*[L.orc_u32_to_f32(C.c_uint32(x)) for x in r])
        po, pd = (C.c_float * 3)(), (C.c_float * 3)()
        L.orc_camera_sample_pos(C.byref(ocam), xi, po)
        xi2 = (C.c_float * 2)(xi[2], xi[3])
        L.orc_camera_sample_dir(C.byref(ocam), C.c_int(pix % W), C.c_int(pix // W), po, C.c_int(W), C.c_int(H), xi2, pd)
        assert np.abs(np.array(po) - o[pix]).max() <= 2e-6, pix        # sincos.approx vs libm: stated tolerance
        assert np.abs(np.array(pd) - d[pix]).max() <= 4e-6, pix
        on_lens += np.linalg.norm(o[pix] - np.float32(sd.camera["look_from"])) <= 0.35 * (1 + 1e-5)
    assert on_lens == W * H
    assert np.ptp(o[:, 0]) > 0.3                                       # the origins really spread over the lens


def test_thin_lens_render_matches_oracle(torch_cuda):
    """whole loop with depth of field: same-RNG image equal to the oracle's except at hit flips"""
    import wpt
    sd = scenes.cornell(24, colored=True)
    cam, ocam = _lens_cameras(sd, 0.2, 3.0)
    W = H = 96
    scene = wpt.scene_from(sd)
    film, st = wpt.Renderer(scene, W, H, 2).render_host(cam, 0, 4, 5, seed=9)
    om = util.oracle_mesh(sd)
    ofilm, oseg = om.render("bvh", ocam, W, H, 0, 4, 5, rng_mode=O.RNG_PHILOX, seed=9)
    ofilm /= 4
    diff = np.abs(film - ofilm).max(0)
    assert (diff > 1e-5).mean() < 0.02, f"{(diff > 1e-5).mean()} of pixels differ"
    assert abs(st["segments"] - oseg) <= 2e-3 * oseg
    # and it is not the pinhole image
    pin = wpt.Renderer(scene, W, H, 2).render_host(wpt.camera_from(sd), 0, 4, 5, seed=9)[0]
    assert (np.abs(pin - film).max(0) > 1e-3).mean() > 0.2


def test_overlap_lanes_render_the_same_frame(torch_cuda):
    """wpt_renderer_set_tuning third knob: two lanes on two streams (every other pass each).  Same counters,
    same paths: segment counts equal, film equal (sums of the same terms; atomics order only)."""
    import torch
    import wpt
    sd = scenes.cornell(16, colored=True)
    scene = wpt.scene_from(sd)
    cam = wpt.camera_from(sd)
    W = H = 96
    out = []
    for overlap, per_pass in ((0, 4), (1, 4), (1, 3), (1, 16)):
        r = wpt.Renderer(scene, W, H, per_pass)
        r.set_tuning(0, 0, overlap)
        f = torch.zeros(3, W * H, device="cuda")
        for _ in range(2):                                   # twice: the lanes' buffers are reused
            f.zero_()
            r.render(cam, f, 3, 11, 5, seed=8)
            st = r.finish()
        out.append((f.clone(), st))
        r.close()
    for f, st in out[1:]:
        assert st["segments"] == out[0][1]["segments"] and st["live"] == out[0][1]["live"]
        assert torch.allclose(f, out[0][0], rtol=0, atol=1e-5)
    assert out[1][1]["passes"] == 6 and out[0][1]["passes"] == 3


def test_non_finite_rays_miss_and_terminate(torch_cuda, cornell):
    """Embree's NaNTest / InfTest (tutorials/verify/verify.cpp:3051,3124) at this boundary: rays with a NaN or
    infinite component come back as misses, the persistent kernel terminates, finite rays in the same launch
    are untouched."""
    from test_host_builder import _poison
    sd, scene, om = cornell
    o, d = util.random_rays(sd, 200_000, seed=31)
    o2, d2, bad = _poison(o, d, seed=4)
    base = _dev_intersect(torch_cuda, scene, o, d)
    got = _dev_intersect(torch_cuda, scene, o2, d2)
    assert bad.sum() > 10_000 and (got[0][bad] == -1).all()
    for a, b in zip(base, got):
        assert np.array_equal(a[~bad], b[~bad], equal_nan=True)


def test_intersect_concurrent_streams(torch_cuda, cornell):
    """one committed scene serves wpt_intersect calls on different streams at the same time
    (read-only BVH, one fetch cursor per call).  Results equal the serial ones."""
    torch = torch_cuda
    sd, scene, om = cornell
    sets = [util.random_rays(sd, 300_000, seed=40 + k) for k in range(4)]
    dev = []
    for o, d in sets:
        dev.append(([torch.from_numpy(np.ascontiguousarray(o[:, k])).cuda() for k in range(3)],
                    [torch.from_numpy(np.ascontiguousarray(d[:, k])).cuda() for k in range(3)]))
    serial = [tuple(x.clone() for x in scene.intersect(o, d)) for o, d in dev]
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream() for _ in dev]
    for _ in range(3):
        outs = []
        for (o, d), st in zip(dev, streams):
            with torch.cuda.stream(st):
                outs.append(scene.intersect(o, d, stream=st))
        torch.cuda.synchronize()
        for a, b in zip(serial, outs):
            # ids may differ at exact t-ties only; t itself is the same minimum either way
            assert (a[0] == b[0]).float().mean().item() >= 0.9999
            assert torch.equal(a[0] >= 0, b[0] >= 0)
            hit = a[0] >= 0
            assert torch.equal(a[1][hit], b[1][hit])


def test_deep_trees_all_stack_variants(torch_cuda):
    """the launcher sizes the shared-memory stack from the tree depth (10 rows up to 11 levels, 14 up to 15, local
    memory overflow beyond): scenes.spiral gives 13 and 17 levels; hits must equal the oracle's either way"""
    import wpt
    from test_host_builder import _spiral_rays
    for n, q, min_depth in ((60, 0.7, 12), (80, 0.75, 16)):
        sd = scenes.spiral(n, q)
        scene, om = wpt.scene_from(sd), util.oracle_mesh(sd)
        assert scene.stats()["max_depth"] > min_depth
        o, d = _spiral_rays(300_000, seed=n)
        got = _dev_intersect(torch_cuda, scene, o, d)
        ref = om.intersect("bvh", o, d)
        assert (ref[0] >= 0).mean() > 0.2
        assert util.compare_hits(ref, got, f"spiral {n}") >= 0.9999


def test_render_is_cuda_graph_capturable(torch_cuda):
    """wpt_render only enqueues kernels, memsets and event fork/joins, so a whole frame (film zero -> render -> 1/spp)
    can be recorded into a CUDA graph and replayed: same film, same counters, in both the serial and the two-lane mode"""
    import torch
    import wpt
    sd = scenes.cornell(16, colored=True)
    scene = wpt.scene_from(sd)
    cam = wpt.camera_from(sd)
    W = H = 96
    for overlap in (0, 1):
        r = wpt.Renderer(scene, W, H, 4)
        r.set_tuning(0, 0, overlap)
        ref = torch.zeros(3, W * H, device="cuda")
        r.render(cam, ref, 0, 8, 5, seed=5)
        st_ref = r.finish()
        wpt.film_scale(ref, 1.0 / 8)
        film = torch.zeros(3, W * H, device="cuda")
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                film.zero_()
                r.render(cam, film, 0, 8, 5, seed=5, stream=side)
                wpt.film_scale(film, 1.0 / 8, stream=side)
        torch.cuda.current_stream().wait_stream(side)
        for _ in range(3):
            film.fill_(7.0)
            g.replay()
            st = r.finish()
            assert st["segments"] == st_ref["segments"] and st["live"] == st_ref["live"]
            assert torch.allclose(film, ref, rtol=0, atol=1e-5)
        # a plain render afterwards works again and is timed
        r.render(cam, film, 0, 8, 5, seed=5)
        assert r.finish()["ms_total"] > 0
        r.close()


def test_depth_seeding_is_exact(torch_cuda):
    """wpt_renderer_set_depth_seeding: camera rays start with a per-pixel tmax seed from a one-ray-per-pixel
    pre-pass; rays that find nothing inside it are traced again.  Same paths as the unseeded render: equal counters,
    equal film up to fp32 atomic order — on a scene full of silhouettes (retries happen) and in both lane modes;
    thin-lens cameras and short renders are not seeded at all."""
    import torch
    import wpt
    for sd, W, H in ((scenes.cornell(16, colored=True), 128, 96), (scenes.knob(120, 50), 160, 128)):
        scene = wpt.scene_from(sd)
        cam = wpt.camera_from(sd)
        out = {}
        for seeding in (0, 1, 9500):   # 9500: seeds 5 % too short, most camera rays take the second trace
            for overlap in (0, 1):
                r = wpt.Renderer(scene, W, H, 8)
                r.set_tuning(0, 0, overlap)
                r.set_depth_seeding(seeding)
                f = torch.zeros(3, W * H, device="cuda")
                r.render(cam, f, 3, 16, 5, seed=21)
                out[(seeding, overlap)] = (f, r.finish())
                r.close()
        f0, st0 = out[(0, 0)]
        assert st0["retraced"] == 0
        for key, (f, st) in out.items():
            assert st["segments"] == st0["segments"] and st["live"] == st0["live"], key
            assert torch.allclose(f, f0, rtol=0, atol=2e-5), key
            if key[0] == 1:
                assert st["retraced"] < 0.05 * st["live"][0], (key, st["retraced"])   # a few silhouette rays at most
            if key[0] == 9500:
                assert st["retraced"] > 0.3 * st["live"][0], (key, st["retraced"])
        # not seeded: fewer than 8 samples, or a thin lens
        r = wpt.Renderer(scene, W, H, 4)
        r.render(cam, torch.zeros(3, W * H, device="cuda"), 0, 4, 3, seed=1)
        assert r.finish()["kernel_launches"] == 1 + 2 * 2 + 1      # raygen, two wavefront bounces, the fused tail
        c = sd.camera
        lens = wpt.ThinlensCamera(c["look_from"], c["look_at"], c["up"], 0.05, 3.0, c["fov_y"])
        r.render(lens, torch.zeros(3, W * H, device="cuda"), 0, 8, 2, seed=1)
        st = r.finish()
        assert st["kernel_launches"] == 2 * (1 + 2 * 2) and st["retraced"] == 0      # depth 2: no tail
        r.close()

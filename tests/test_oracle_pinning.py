"""Pins the CPU oracle: known-answer vectors, the committed golden fixtures generated from the real
Embree 3.6.1 of the reference tree, and — where oracle/_ref travelled — Embree itself, live."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_py as O
import scenes
import util

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_pcg32_canonical_sequence():
    """pcg32-demo (pcg-random.org, seed 42 stream 54) — the generator enoki/random.h:62-77 restates"""
    L = O.lib()
    p = O.OrcPcg()
    L.orc_pcg32_seed(C.byref(p), C.c_uint64(42), C.c_uint64(54))
    got = [L.orc_pcg32_next_u32(C.byref(p)) for _ in range(6)]
    assert got == [0xa15c02b7, 0x7b47f409, 0xba1d3330, 0x83d2f293, 0xbfa4784b, 0xcbed606e]


def test_pcg32_float_range_and_formula():
    L = O.lib()
    for u, want in ((0, 0.0), (0xFFFFFFFF, 1.0 - 2.0 ** -23), (0x80000000, 0.5)):
        assert L.orc_u32_to_f32(C.c_uint32(u)) == np.float32(want)


@pytest.mark.parametrize("ctr,key,want", [
    ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
     (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
])
def test_philox_random123_kat(ctr, key, want):
    """Random123 kat_vectors, philox4x32 10 rounds"""
    o = (C.c_uint32 * 4)()
    O.lib().orc_philox4x32_10((C.c_uint32 * 4)(*ctr), (C.c_uint32 * 2)(*key), o)
    assert tuple(o) == want


@pytest.mark.parametrize("backend", ["brute", "bvh"])
def test_embree_minimal_kat(backend):
    """dependencies/embree/tutorials/minimal/minimal.cpp:72-213"""
    m = O.Mesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]])
    o = np.array([[0, 0, -1], [1, 1, -1]], np.float32)
    d = np.array([[0, 0, 1], [0, 0, 1]], np.float32)
    hid, t, u, v = m.intersect(backend, o, d, tmin=0.0, tmax=np.inf)
    assert hid.tolist() == [0, -1] and t[0] == 1.0


@pytest.mark.parametrize("backend", ["brute", "bvh"])
def test_embree_triangle_hit_kat(backend):
    """verify.cpp:2312-2399 TriangleHitTest: u,v,t within 16 ulp"""
    m = O.Mesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]])
    g = np.arange(16, dtype=np.float32)
    U, V = np.meshgrid(g, g, indexing="ij")
    u0 = (U.reshape(-1) / 32 + 1 / 64).astype(np.float32)
    v0 = (V.reshape(-1) / 32 + 1 / 64).astype(np.float32)
    o = np.tile(np.array([[0, 0, -1]], np.float32), (256, 1))
    d = np.stack([u0, v0, np.ones_like(u0)], 1)
    hid, t, u, v = m.intersect(backend, o, d, tmin=0.0, tmax=np.inf)
    eps = 16 * np.finfo(np.float32).eps
    assert (hid == 0).all()
    assert np.abs(u - u0).max() <= eps and np.abs(v - v0).max() <= eps and np.abs(t - 1).max() <= eps


@pytest.mark.parametrize("name,gen", [("cornell8", lambda: scenes.cornell(8)), ("knob40", lambda: scenes.knob(40, 16))])
def test_oracle_matches_golden_embree_hits(name, gen):
    """committed vectors produced by Embree 3.6.1 (tests/golden/make_golden.py)"""
    g = np.load(os.path.join(GOLD, f"hits_{name}.npz"))
    om = util.oracle_mesh(gen())
    for backend in ("bvh", "brute") if name == "cornell8" else ("bvh",):
        got = om.intersect(backend, g["o"], g["d"])
        rate = util.compare_hits((g["id"], g["t"], g["u"], g["v"]), got, f"{name}/{backend}")
        assert rate == 1.0


def test_oracle_image_matches_golden():
    """the restated loop, reference RNG, on the oracle's own intersector == the same loop on Embree"""
    sd = scenes.cornell(8, colored=True)
    om, cam = util.oracle_mesh(sd), util.oracle_camera(sd)
    for mode, fn, seed in ((O.RNG_PCG32, "image_cornell8_pcg32.npz", 0), (O.RNG_PHILOX, "image_cornell8_philox.npz", 2024)):
        g = np.load(os.path.join(GOLD, fn))
        film, seg = om.render("bvh", cam, 32, 32, 0, 16, 5, rng_mode=mode, seed=seed)
        assert seg == int(g["segments"])
        assert np.array_equal(film, g["film"])


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not present")
def test_oracle_vs_live_embree():
    sd = scenes.knob(120, 48)
    om = util.oracle_mesh(sd)
    o, d = util.random_rays(sd, 200_000, seed=1)
    ref = om.intersect("embree", o, d)
    got = om.intersect("bvh", o, d)
    assert util.compare_hits(ref, got, "oracle/embree") >= 0.99999
    # robust (Pluecker) Embree disagrees only at grazing hits
    rob = om.intersect("embree_robust", o, d)
    assert (rob[0] == ref[0]).mean() > 0.9995


def test_mask_semantics_oracle():
    sd = scenes.cornell(8)
    om = util.oracle_mesh(sd)
    o, d = util.random_rays(sd, 1000, seed=2)
    m = (np.arange(1000) % 2).astype(np.uint8)
    hid, t, u, v = om.intersect("bvh", o, d, mask=m)
    assert (hid[m == 0] == -2).all() and np.isnan(t[m == 0]).all()
    hid1, *_ = om.intersect("bvh", o, d, mask=np.ones(1, np.uint8))
    hid0, *_ = om.intersect("bvh", o, d, mask=np.zeros(1, np.uint8))
    assert (hid1 != -2).all() and (hid0 == -2).all()


def test_camera_center_ray_and_frame():
    """pinhole.h: the centre of the film looks along look_at - look_from; frame is orthonormal"""
    cam = O.camera((0, 3.03, 5), (0, 0.03, 0), fov_y=40 / 180 * np.pi)
    x, y, z = np.array(cam.x), np.array(cam.y), np.array(cam.z)
    assert abs(x @ y) < 1e-6 and abs(y @ z) < 1e-6 and abs(np.linalg.norm(z) - 1) < 1e-6
    L = O.lib()
    po, pd = (C.c_float * 3)(), (C.c_float * 3)()
    L.orc_camera_sample_pos(C.byref(cam), (C.c_float * 2)(0.3, 0.7), po)
    assert np.allclose(np.array(po), [0, 3.03, 5])           # lens radius 0 -> pinhole (main.cpp:383)
    W = H = 64
    L.orc_camera_sample_dir(C.byref(cam), C.c_int(31), C.c_int(31), po, C.c_int(W), C.c_int(H), (C.c_float * 2)(1.0, 1.0), pd)
    want = np.array([0, -3, -5.0]) / np.linalg.norm([0, -3, -5.0])
    assert np.allclose(np.array(pd), want, atol=1e-6)
    # +x pixel goes towards -film x (film_xy = (0.5 - ndc) * size, pinhole.h:46)
    L.orc_camera_sample_dir(C.byref(cam), C.c_int(0), C.c_int(31), po, C.c_int(W), C.c_int(H), (C.c_float * 2)(0.0, 1.0), pd)
    assert np.array(pd) @ x > 0


def test_frame_and_warp_properties():
    L = O.lib()
    rng = np.random.default_rng(0)
    for _ in range(200):
        n = rng.normal(size=3); n /= np.linalg.norm(n)
        m = (C.c_float * 9)()
        L.orc_frame_from_normal((C.c_float * 3)(*n), m)
        M = np.array(m).reshape(3, 3)
        assert np.allclose(M.T @ M, np.eye(3), atol=2e-6)          # orthonormal
        assert np.allclose(M[:, 1], n, atol=1e-7)                  # normal is local Y (coordframe.h:23)
        s = rng.random(2).astype(np.float32)
        out = (C.c_float * 3)()
        L.orc_cosine_hemisphere((C.c_float * 2)(*s), out)
        v = np.array(out)
        assert abs(np.linalg.norm(v) - 1) < 1e-6 and v[1] >= 0 and abs(v[1] - np.sqrt(s[0])) < 1e-7
        w = (C.c_float * 3)()
        L.orc_frame_to_world(m, out, w)
        back = (C.c_float * 3)()
        L.orc_frame_to_local(m, w, back)
        assert np.allclose(np.array(back), v, atol=2e-6)
    d = (C.c_float * 2)()
    L.orc_disk_from_square((C.c_float * 2)(1.0, 0.5), d)
    assert np.allclose(np.array(d), [1, 0], atol=1e-7)


def test_render_reference_semantics():
    """with every albedo 1 the pixel value is P(path escapes within depth segments);
    nothing to hit -> exactly 1; sealed box -> exactly 0; depth 1 -> primary visibility of the sky."""
    far = O.Mesh([[100, 100, 100], [101, 100, 100], [100, 101, 100]], [[0, 1, 2]],
                 nrm=[[0, 0, 1]], nrm_idx=[[0, 0, 0]])
    cam = O.camera((0, 0, 5), (0, 0, 0))
    film, seg = far.render("bvh", cam, 8, 8, 0, 4, 3)
    assert (film == 4.0).all() and seg == 8 * 8 * 4
    sd = scenes.interior(9, 1, (8, 8), skylight=False)
    om = util.oracle_mesh(sd)
    film, seg = om.render("bvh", util.oracle_camera(sd), 16, 16, 0, 2, 4)
    assert film.max() == 0.0 and seg == 16 * 16 * 2 * 4


def test_spp_ranges_compose_pcg_and_philox():
    sd = scenes.cornell(8)
    om, cam = util.oracle_mesh(sd), util.oracle_camera(sd)
    for mode in (O.RNG_PCG32, O.RNG_PHILOX):
        whole, s = om.render("bvh", cam, 24, 24, 0, 6, 4, rng_mode=mode, seed=9)
        a, s1 = om.render("bvh", cam, 24, 24, 0, 2, 4, rng_mode=mode, seed=9)
        b, s2 = om.render("bvh", cam, 24, 24, 2, 4, 4, rng_mode=mode, seed=9)
        assert np.array_equal(whole, a + b) and s == s1 + s2


def test_texture_eval_restatement():
    """ImageTexture::eval (src/main.cpp:62-76) as restated: a 1x1 texture is the constant the reference builds for every
    material (weights sum to 1 within 2 ulp); texel centres reproduce the texel; the midpoint of four texels is their
    mean; coordinates clamp (texel_fetch_wrap :56-60); a NON-square image is indexed by its row length (the :52 bug)."""
    import ctypes as C
    L = O.lib()

    def ev(img, u, v):
        img = np.ascontiguousarray(img, np.float32)
        t = O.OrcTexture(img.ctypes.data, img.shape[1], img.shape[0])
        out = (C.c_float * 3)()
        L.orc_texture_eval(C.byref(t), C.c_float(u), C.c_float(v), out)
        return np.array(out, np.float32)

    rng = np.random.default_rng(0)
    one = np.array([[[0.25, 0.5, 0.75]]], np.float32)
    for u, v in rng.random((50, 2)):
        assert np.abs(ev(one, u, v) - one[0, 0]).max() <= 2 * np.finfo(np.float32).eps
    img = rng.random((5, 8, 3)).astype(np.float32)          # h = 5, w = 8: non-square
    for y in range(5):
        for x in range(8):
            assert np.allclose(ev(img, (x + 0.5) / 8, (y + 0.5) / 5), img[y, x], atol=1e-6)
    assert np.allclose(ev(img, 3.0 / 8, 2.0 / 5), img[1:3, 2:4].reshape(-1, 3).mean(0), atol=1e-6)
    assert np.allclose(ev(img, -3.0, 9.0), img[4, 0], atol=1e-6) and np.allclose(ev(img, 2.0, -1.0), img[0, 7], atol=1e-6)

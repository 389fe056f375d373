"""-m gpu: round-2 rows — per-ray tmin/tmax STREAMS at the intersect boundary (the form the reference passes,
src/ray.h:20-26, src/optixdata.h:20-21), the fused tail kernel and the balanced warp-queue flush against the wavefront
pipeline, deterministic t-ties, the derived hit attributes of OptixBackend::intersect (src/accel/optix_backend.h:463-486),
wpt_film_reduce on the C ABI, pinned host films, re-commit after a device-side refit."""
import ctypes as C

import numpy as np
import pytest

import oracle_py as O
import scenes
import util

pytestmark = pytest.mark.gpu


def _cu(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _streams(torch, o, d):
    return [_cu(torch, o[:, k]) for k in range(3)], [_cu(torch, d[:, k]) for k in range(3)]


def test_intersect_per_ray_tmin_tmax_streams():
    """Ray3C carries tmin / tmax as per-ray ARRAYS (src/ray.h:20-26) and struct Params passes their device pointers
    (src/optixdata.h:20-21).  Random windows per ray, through wpt_intersect, against the oracle called ray by ray with
    the same windows; then mixed forms (stream + scalar)."""
    import torch
    import wpt
    sd = scenes.cornell(50)
    scene, om = wpt.scene_from(sd), util.oracle_mesh(sd)
    n = 120_000
    o, d = util.random_rays(sd, n, seed=61)
    rng = np.random.default_rng(62)
    tmin = rng.uniform(0.0, 0.8, n).astype(np.float32)
    tmax = (tmin + rng.uniform(0.05, 3.0, n)).astype(np.float32)
    tmax[rng.random(n) < 0.2] = np.float32(1e20)
    ot, dt = _streams(torch, o, d)
    got = tuple(x.cpu().numpy() for x in scene.intersect(ot, dt, tmin=_cu(torch, tmin), tmax=_cu(torch, tmax)))
    ref = om.intersect("bvh", o, d, tmin=tmin, tmax=tmax)
    assert (got[0] >= 0).mean() > 0.2 and (got[0] < 0).mean() > 0.2       # the windows really cut hits away
    util.compare_hits(ref, got, "per-ray windows / oracle")
    if O.ref_available():
        fast = om.intersect("embree", o, d, tmin=tmin, tmax=tmax)
        robust = om.intersect("embree_robust", o, d, tmin=tmin, tmax=tmax)
        util.compare_hits(fast, got, "per-ray windows / embree", robust=robust)
    hit = got[0] >= 0
    assert (got[1][hit] > tmin[hit]).all() and (got[1][hit] <= tmax[hit]).all()
    # a constant stream equals the scalar form; a stream for one bound and a scalar for the other is allowed
    full = lambda v: torch.full((n,), v, dtype=torch.float32, device="cuda")
    a = scene.intersect(ot, dt, tmin=full(0.25), tmax=full(1.5))
    b = scene.intersect(ot, dt, tmin=0.25, tmax=1.5)
    c = scene.intersect(ot, dt, tmin=full(0.25), tmax=1.5)
    torch.cuda.synchronize()
    for x, y, z in zip(a, b, c):
        assert torch.equal(x, y, ) or torch.equal(torch.nan_to_num(x), torch.nan_to_num(y))
        assert torch.equal(torch.nan_to_num(x), torch.nan_to_num(z))
    # a negative tmin is clamped to 0 (include/wpt.h): same hits as tmin = 0
    e = scene.intersect(ot, dt, tmin=full(-5.0), tmax=1e20)
    g = scene.intersect(ot, dt, tmin=0.0, tmax=1e20)
    torch.cuda.synchronize()
    assert torch.equal(e[0], g[0])


def test_fused_tail_equals_the_wavefront_pipeline():
    """wpt_renderer_set_tail: bounces >= k of a pass in ONE fused trace+shade launch.  Same Philox counters, same
    paths: per-bounce ray counts identical, film equal up to fp32 atomic order, for every k, in both lane modes, with
    and without depth seeding, on coloured albedos (throughput products travel through the kernel's path state)."""
    import torch
    import wpt
    for sd, W, H, depth in ((scenes.cornell(16, colored=True), 96, 80, 6), (scenes.knob(120, 50), 128, 96, 8),
                            (scenes.interior(33, 2, (16, 32)), 64, 64, 12)):
        scene = wpt.scene_from(sd)
        cam = wpt.camera_from(sd)
        ref = None
        for tail in (-1, 1, 2, 3, depth - 1, depth, depth + 5):
            for overlap in (0, 1):
                r = wpt.Renderer(scene, W, H, 4)
                r.set_tuning(0, 0, overlap)
                r.set_tail(tail)
                f = torch.zeros(3, W * H, device="cuda")
                r.render(cam, f, 2, 9, depth, seed=33)
                st = r.finish()
                r.close()
                if ref is None:
                    ref = (f, st)
                    assert st["kernel_launches"] >= 3 * (1 + 2 * depth)      # the pure wavefront pipeline
                    continue
                assert st["live"] == ref[1]["live"] and st["segments"] == ref[1]["segments"], (sd.name, tail, overlap)
                assert torch.allclose(f, ref[0], rtol=0, atol=2e-5), (sd.name, tail, overlap)
                if 1 <= tail < depth and not overlap:
                    assert st["kernel_launches"] < ref[1]["kernel_launches"]
    # a tail that starts at bounce 1 of a depth-2 render; depth 1 has no tail at all
    sd = scenes.cornell(8)
    scene, cam = wpt.scene_from(sd), wpt.camera_from(sd)
    for depth in (1, 2):
        out = []
        for tail in (-1, 1):
            r = wpt.Renderer(scene, 48, 48, 2)
            r.set_tail(tail)
            f = torch.zeros(3, 48 * 48, device="cuda")
            r.render(cam, f, 0, 4, depth, seed=1)
            out.append((f, r.finish()))
        assert torch.equal(out[0][0], out[1][0]) and out[0][1]["live"] == out[1][1]["live"]


def test_fused_tail_under_cuda_graph_and_profile():
    import torch
    import wpt
    sd = scenes.cornell(16, colored=True)
    scene, cam = wpt.scene_from(sd), wpt.camera_from(sd)
    W = H = 64
    r = wpt.Renderer(scene, W, H, 4)
    r.set_tail(1)
    ref = torch.zeros(3, W * H, device="cuda")
    r.render(cam, ref, 0, 8, 6, seed=3)
    st_ref = r.finish()
    r.set_profile(True)
    f = torch.zeros(3, W * H, device="cuda")
    r.render(cam, f, 0, 8, 6, seed=3)
    st = r.finish()
    r.set_profile(False)
    assert st["live"] == st_ref["live"] and st["ms_trace"] > 0 and torch.allclose(f, ref, rtol=0, atol=2e-5)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    film = torch.zeros(3, W * H, device="cuda")
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            film.zero_()
            r.render(cam, film, 0, 8, 6, seed=3, stream=side)
    torch.cuda.current_stream().wait_stream(side)
    for _ in range(2):
        film.fill_(3.0)
        g.replay()
        st = r.finish()
        assert st["live"] == st_ref["live"] and torch.allclose(film, ref, rtol=0, atol=2e-5)


def test_ties_resolve_to_the_lowest_primitive():
    """Equal-t hits (coincident triangles, shared edges) are folded with an integer min on (t bits, primitive): the
    lowest primitive index wins whatever the order of the tests, (u, v) come from that same triangle, and two runs of
    the same rays are bit-identical (round 1 took the last writer)."""
    import torch
    import wpt
    rng = np.random.default_rng(5)
    tri = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    # 37 coincident copies of one triangle, each with its own vertices, in a shuffled spatial order among 500 others
    far = rng.uniform(2, 3, (500, 1, 3)).astype(np.float32) + rng.normal(0, 0.05, (500, 3, 3)).astype(np.float32)
    pos = np.concatenate([np.tile(tri, (37, 1)), far.reshape(-1, 3)])
    idx = np.arange(len(pos), dtype=np.int32).reshape(-1, 3)
    perm = rng.permutation(len(idx))
    idx = idx[perm]
    s = wpt.Scene()
    s.add_triangle_mesh(pos, idx)
    s.commit(0)
    n = 50_000
    o = np.stack([rng.uniform(0.05, 0.45, n), rng.uniform(0.05, 0.45, n), np.full(n, -1.0)], 1).astype(np.float32)
    d = np.tile(np.array([[0, 0, 1]], np.float32), (n, 1))
    ot, dt = _streams(torch, o, d)
    a = s.intersect(ot, dt)
    b = s.intersect(ot, dt)
    torch.cuda.synchronize()
    copies = np.nonzero(perm < 37)[0]                   # primitive ids (positions in idx) of the coincident copies
    assert (a[0].cpu().numpy() == copies.min()).all()
    for x, y in zip(a, b):
        assert torch.equal(x, y)
    assert torch.allclose(a[2], ot[0]) and torch.allclose(a[3], ot[1])      # u, v of THAT triangle (= x, y here)
    # shared edges of a closed mesh: run-to-run identical ids on grazing-rich random rays
    sd = scenes.cornell(50)
    scene = wpt.scene_from(sd)
    o, d = util.random_rays(sd, 500_000, seed=3)
    ot, dt = _streams(torch, o, d)
    runs = [tuple(x.clone() for x in scene.intersect(ot, dt)) for _ in range(3)]
    torch.cuda.synchronize()
    for r in runs[1:]:
        assert torch.equal(r[0], runs[0][0])
        hit = r[0] >= 0
        assert torch.equal(r[1][hit], runs[0][1][hit]) and torch.equal(r[2][hit], runs[0][2][hit])


def test_hit_attributes_vs_oracle():
    """position, geometry normal, interpolated shading normal, texcoord and material index for every hit lane against
    the oracle's restatement of src/accel/optix_backend.h:463-486; missed lanes untouched; scenes without normals /
    without texcoords."""
    import torch
    import wpt
    sd = scenes.cornell(24, colored=True)
    rng = np.random.default_rng(8)
    nuv = 97
    uvs = rng.random((nuv, 2)).astype(np.float32)
    uvi = rng.integers(0, nuv, (sd.n_tris, 3)).astype(np.int32)
    uvi[::11] = -1                                             # faces without texcoords
    s = wpt.Scene()
    s.add_triangle_mesh(sd.pos, sd.idx, sd.nrm, sd.nidx, sd.mat, sd.albedo, texcoords=uvs, texcoord_triplets=uvi)
    s.commit(0)
    om = util.oracle_mesh(sd)
    n = 40_000
    o, d = util.random_rays(sd, n, seed=12)
    ot, dt = _streams(torch, o, d)
    hit = s.intersect(ot, dt)
    att = s.hit_attributes(ot, dt, hit)
    torch.cuda.synchronize()
    hid, t, u, v = (x.cpu().numpy() for x in hit)
    got = {k: x.cpu().numpy() for k, x in att.items()}
    miss = hid < 0
    assert miss.any() and np.isnan(got["p"][:, miss]).all() and (got["mat"][miss] == -1).all()
    L = O.lib()
    uvs_c, uvi_c = np.ascontiguousarray(uvs), np.ascontiguousarray(uvi)
    f3 = lambda: (C.c_float * 3)()
    checked = 0
    for i in np.nonzero(~miss)[0][:3000]:
        p, gn, sn, uv, mat = f3(), f3(), f3(), (C.c_float * 2)(), C.c_int32()
        L.orc_hit_attributes(C.byref(om.c), uvs_c.ctypes.data_as(C.c_void_p), uvi_c.ctypes.data_as(C.c_void_p),
                             (C.c_float * 3)(*o[i]), (C.c_float * 3)(*d[i]), C.c_int32(int(hid[i])), C.c_float(t[i]),
                             C.c_float(u[i]), C.c_float(v[i]), p, gn, sn, uv, C.byref(mat))
        assert np.array_equal(np.array(p, np.float32), got["p"][:, i])                     # same two roundings
        assert np.abs(np.array(gn) - got["gn"][:, i]).max() <= 2e-6                        # rsqrt.approx vs 1/sqrt
        assert np.array_equal(np.array(sn, np.float32), got["sn"][:, i])
        assert np.array_equal(np.array(uv, np.float32), got["uv"][:, i])
        assert mat.value == got["mat"][i] == sd.mat[hid[i]]
        checked += 1
    assert checked == 3000
    assert np.abs(np.linalg.norm(got["gn"][:, ~miss], axis=0) - 1).max() < 1e-5
    # partial outputs; no normals, no texcoords
    only = s.hit_attributes(ot, dt, hit, want=("gn",))
    assert set(only) == {"gn"} and torch.equal(torch.nan_to_num(only["gn"]), torch.nan_to_num(att["gn"]))
    bare = wpt.Scene()
    bare.add_triangle_mesh(sd.pos, sd.idx)
    bare.commit(0)
    hb = bare.intersect(ot, dt)
    ab = bare.hit_attributes(ot, dt, hb)
    torch.cuda.synchronize()
    hitb = hb[0] >= 0
    assert torch.allclose(ab["sn"][:, hitb], ab["gn"][:, hitb], atol=2e-6)                 # geometric-normal fallback
    assert float(ab["uv"][:, hitb].abs().max()) == 0.0 and int(ab["mat"][hitb].max()) == 0


def test_film_reduce_on_the_c_abi():
    """wpt_film_reduce: NCCL bound at run time inside libwpt.so.  A one-rank communicator is the identity (runs on the
    single-GPU test box); with >= 2 GPUs in one process (wpt_nccl_comm_create_all + group calls) two spp shards summed
    onto device 0 equal the unsharded film — what `wurst --gpus N` does."""
    import torch
    import wpt
    L = wpt.lib()
    sd = scenes.cornell(16, colored=True)
    W = H = 64
    N = W * H
    ndev = torch.cuda.device_count()
    comms = (C.c_void_p * max(ndev, 1))()
    devs = (C.c_int * max(ndev, 1))(*range(max(ndev, 1)))
    wpt._check(L.wpt_nccl_comm_create_all(C.c_int(1), devs, comms))
    scene, cam = wpt.scene_from(sd), wpt.camera_from(sd)
    r = wpt.Renderer(scene, W, H, 4)
    whole = torch.zeros(3, N, device="cuda:0")
    r.render(cam, whole, 0, 8, 5, seed=5); r.finish()
    keep = whole.clone()
    base = whole.data_ptr()
    wpt._check(L.wpt_film_reduce(comms[0], C.c_void_p(base), C.c_void_p(base + 4 * N), C.c_void_p(base + 8 * N),
                                 C.c_uint32(N), C.c_int(0), wpt._stream_handle(None)))
    # non-contiguous planes take the grouped three-call form
    planes = [whole[k].clone() for k in (2, 0, 1)]
    wpt._check(L.wpt_film_reduce(comms[0], *[C.c_void_p(p.data_ptr()) for p in planes], C.c_uint32(N), C.c_int(0),
                                 wpt._stream_handle(None)))
    torch.cuda.synchronize()
    assert torch.equal(whole, keep) and torch.equal(planes[0], keep[2])
    L.wpt_nccl_comm_destroy(comms[0])
    if ndev < 2:
        return
    wpt._check(L.wpt_nccl_comm_create_all(C.c_int(2), devs, comms))
    parts = []
    for dev, (b, c) in enumerate(((0, 3), (3, 5))):
        torch.cuda.set_device(dev)
        sc = wpt.scene_from(sd, device=dev)
        rr = wpt.Renderer(sc, W, H, 4)
        f = torch.zeros(3, N, device=f"cuda:{dev}")
        rr.render(cam, f, b, c, 5, seed=5); rr.finish()
        parts.append((f, sc, rr))
    wpt._check(L.wpt_nccl_group_start())
    for dev, (f, _, _) in enumerate(parts):
        torch.cuda.set_device(dev)
        p = f.data_ptr()
        wpt._check(L.wpt_film_reduce(comms[dev], C.c_void_p(p), C.c_void_p(p + 4 * N), C.c_void_p(p + 8 * N),
                                     C.c_uint32(N), C.c_int(0), wpt._stream_handle(None)))
    wpt._check(L.wpt_nccl_group_end())
    for dev in range(2):
        torch.cuda.synchronize(dev)
    torch.cuda.set_device(0)
    assert torch.allclose(parts[0][0], keep, rtol=0, atol=2e-5)
    for dev in range(2):
        L.wpt_nccl_comm_destroy(comms[dev])


def test_render_host_into_pinned_and_pageable_films():
    import wpt
    sd = scenes.cornell(16, colored=True)
    scene, cam = wpt.scene_from(sd), wpt.camera_from(sd)
    W = H = 80
    r = wpt.Renderer(scene, W, H, 4)
    pin = wpt.PinnedArray((3, W * H))
    a, st = r.render_host(cam, 0, 8, 5, seed=2, out=pin.array)
    b, _ = r.render_host(cam, 0, 8, 5, seed=2, out=np.empty((3, W * H), np.float32))
    c, _ = r.render_host(cam, 0, 8, 5, seed=2)
    assert a is pin.array and np.allclose(a, b, atol=1e-6) and np.allclose(a, c, atol=1e-6) and a.mean() > 0.01
    assert st["segments"] > 0


def test_recommit_after_device_refit_uses_the_moved_vertices():
    """wpt_scene_refit_device left the host copy stale, so a later commit silently rebuilt the OLD mesh.
    Now the scene keeps the device-side positions and reads them back at the next commit."""
    import torch
    import wpt
    sd = scenes.cornell(16)
    scene = wpt.scene_from(sd)
    moved = (sd.pos * np.float32(1.25) + np.float32(0.1)).astype(np.float32)
    scene.refit(torch.from_numpy(moved.reshape(-1)).cuda())
    torch.cuda.synchronize()
    scene.commit(0)                                    # rebuild: must see `moved`
    fresh = wpt.Scene()
    fresh.add_triangle_mesh(moved, sd.idx, sd.nrm, sd.nidx, sd.mat, sd.albedo)
    fresh.commit(0)
    o, d = util.random_rays(sd, 100_000, seed=4)
    o = (o * np.float32(1.25) + np.float32(0.1)).astype(np.float32)
    ot, dt = _streams(torch, o, d)
    a, b = scene.intersect(ot, dt), fresh.intersect(ot, dt)
    torch.cuda.synchronize()
    assert (a[0] >= 0).float().mean() > 0.5
    for x, y in zip(a, b):
        assert torch.equal(torch.nan_to_num(x), torch.nan_to_num(y))
    assert scene.stats()["n_nodes"] == fresh.stats()["n_nodes"]


def test_image_textures_bilinear_eval_vs_oracle():
    """row f3: per-material ImageTexture with the bilinear eval of src/main.cpp:62-76 (index bug of :52 fixed).
    (1) wpt_shade_step weights == the oracle's texture_eval at the interpolated texcoord, bit for bit, on a non-square
    image; (2) a whole render with textured and constant materials == the oracle's render with the same Philox numbers;
    (3) a 1x1 texture is the constant albedo."""
    import torch
    import wpt
    sd = scenes.cornell(20, colored=True)
    rng = np.random.default_rng(17)
    nuv = 211
    uvs = rng.random((nuv, 2)).astype(np.float32) * 1.4 - 0.2          # some outside [0,1]: clamping
    uvi = rng.integers(0, nuv, (sd.n_tris, 3)).astype(np.int32)
    uvi[::13] = -1
    n_mat = len(sd.albedo)
    tex = {0: rng.random((5, 8, 3)).astype(np.float32), n_mat - 1: rng.random((16, 16, 3)).astype(np.float32)}
    s = wpt.Scene()
    s.add_triangle_mesh(sd.pos, sd.idx, sd.nrm, sd.nidx, sd.mat, sd.albedo, texcoords=uvs, texcoord_triplets=uvi)
    for m, img in tex.items():
        s.set_texture(m, img)
    s.commit(0)
    om = util.oracle_mesh(sd)
    om.set_textures(uvs, uvi, tex)
    o, d = util.random_rays(sd, 30_000, seed=5)
    hid, t, u, v = om.intersect("bvh", o, d)
    live = np.nonzero(hid >= 0)[0][:4000]
    sq = rng.random((live.size, 2)).astype(np.float32)
    cu = lambda a: _cu(torch, a)
    no, nd, w = s.shade_step([cu(o[live, k]) for k in range(3)], [cu(d[live, k]) for k in range(3)],
                             cu(hid[live]), cu(t[live]), cu(u[live]), cu(v[live]), cu(sq[:, 0]), cu(sq[:, 1]))
    w = np.stack([x.cpu().numpy() for x in w], 1)
    textured = 0
    for k, i in enumerate(live[:1500]):
        _, _, rw = om.shade_step(o[i], d[i], hid[i], t[i], u[i], v[i], sq[k])
        assert np.array_equal(rw, w[k]), (k, rw, w[k])
        textured += sd.mat[hid[i]] in tex
    assert textured > 100
    W = H = 96
    film, st = wpt.Renderer(s, W, H, 4).render_host(wpt.camera_from(sd), 0, 8, 5, seed=4)
    assert st["tail_from"] == -1                               # textured scenes run the wavefront pair
    ofilm, oseg = om.render("bvh", util.oracle_camera(sd), W, H, 0, 8, 5, rng_mode=O.RNG_PHILOX, seed=4)
    ofilm /= 8
    assert (np.abs(film - ofilm).max(0) > 1e-5).mean() < 0.01 and abs(st["segments"] - oseg) <= 2e-3 * oseg
    # not the untextured image
    plain = wpt.scene_from(sd)
    pfilm, _ = wpt.Renderer(plain, W, H, 4).render_host(wpt.camera_from(sd), 0, 8, 5, seed=4)
    assert (np.abs(pfilm - film).max(0) > 1e-3).mean() > 0.05
    # 1x1 == constant albedo
    s1 = wpt.Scene()
    s1.add_triangle_mesh(sd.pos, sd.idx, sd.nrm, sd.nidx, sd.mat, sd.albedo)
    s1.set_texture(0, np.array([[sd.albedo[0]]], np.float32))
    s1.commit(0)
    f1, _ = wpt.Renderer(s1, W, H, 4).render_host(wpt.camera_from(sd), 0, 8, 5, seed=4)
    assert np.allclose(f1, pfilm, atol=2e-6)


def test_commit_cached_shares_one_host_build(tmp_path):
    """wpt_scene_commit_cached: the first commit builds and writes the tree file, the second loads it (same hits, same
    stats); another mesh or a damaged file is rebuilt, never trusted."""
    import torch
    import wpt
    sd = scenes.knob(100, 40)
    path = tmp_path / "tree.bin"
    a = wpt.scene_from(sd, cache_path=path)
    assert a.from_cache is False and path.exists()
    b = wpt.scene_from(sd, cache_path=path)
    assert b.from_cache is True
    sa, sb = a.stats(), b.stats()
    assert all(sa[k] == sb[k] for k in ("n_nodes", "max_depth", "sah_cost", "n_leaf_slots"))
    o, d = util.random_rays(sd, 100_000, seed=6)
    ot, dt = _streams(torch, o, d)
    for x, y in zip(a.intersect(ot, dt), b.intersect(ot, dt)):
        assert torch.equal(torch.nan_to_num(x), torch.nan_to_num(y))
    r = wpt.Renderer(b, 64, 64, 2)
    film, st = r.render_host(wpt.camera_from(sd), 0, 4, 4, seed=1)
    assert st["segments"] > 0                                   # refit levels etc. came along: the renderer works
    b.refit(sd.pos)                                             # and so does the refit (level lists from the file)
    other = scenes.knob(100, 41)
    c = wpt.scene_from(other, cache_path=path)                  # different mesh: hash mismatch -> rebuilt and rewritten
    assert c.from_cache is False
    raw = bytearray(path.read_bytes()); raw[8] ^= 0xFF
    path.write_bytes(bytes(raw))                                # damaged header
    e = wpt.scene_from(other, cache_path=path)
    assert e.from_cache is False and wpt.scene_from(other, cache_path=path).from_cache is True


def _tree_words_hash(chunks):
    """csrc/api.cu tree_payload_hash: FNV-1a over 64-bit words with an extra xor-shift, the trailing half word bytewise"""
    M = (1 << 64) - 1
    h = 0xcbf29ce484222325
    for raw in chunks:
        k = len(raw) // 8 * 8
        for w in np.frombuffer(raw[:k], "<u8").tolist():
            h ^= w
            h = (h * 0x100000001b3) & M
            h ^= h >> 29
        for byte in raw[k:]:
            h ^= byte
            h = (h * 0x100000001b3) & M
    return h


def test_commit_cached_checks_what_it_loads(tmp_path):
    """a tree file is outside input.  max_depth picks the traversal stack and `base` is followed without a
    range check by the kernels, so a stale, damaged or forged file must be refused at load (and the tree built), even
    when its checksum has been made to fit."""
    import struct
    import torch
    import wpt
    sd = scenes.knob(40, 16)
    path = tmp_path / "tree.bin"
    ref = wpt.scene_from(sd, cache_path=path)
    good = path.read_bytes()
    assert wpt.scene_from(sd, cache_path=path).from_cache is True
    assert (path.stat().st_mode & 0o077) == 0                    # written 0600
    HDR = 88                                                     # TreeFileHeader
    (magic, mhash, version, n_tris, n_nodes, max_depth, n_blocks, n_lnodes, n_lbegin, sah, bms, tag,
     phash) = struct.unpack("<QQIIIIQQQddQQ", good[:HDR])
    assert n_lbegin == max_depth + 1 and n_lnodes == n_nodes and len(good) == HDR + 64 * n_blocks + 4 * (n_lnodes + n_lbegin)
    blocks = good[HDR:HDR + 64 * n_blocks]
    lnodes = good[HDR + 64 * n_blocks:HDR + 64 * n_blocks + 4 * n_lnodes]
    lbegin = good[HDR + 64 * n_blocks + 4 * n_lnodes:]
    assert _tree_words_hash([blocks, lnodes, lbegin]) == phash   # this test's restatement of the checksum is right

    def header(**kw):
        f = dict(magic=magic, mhash=mhash, version=version, n_tris=n_tris, n_nodes=n_nodes, max_depth=max_depth,
                 n_blocks=n_blocks, n_lnodes=n_lnodes, n_lbegin=n_lbegin, sah=sah, bms=bms, tag=tag, phash=phash)
        f.update(kw)
        return struct.pack("<QQIIIIQQQddQQ", *f.values())

    o, d = util.random_rays(sd, 20_000, seed=8)
    ot, dt = _streams(torch, o, d)
    want = ref.intersect(ot, dt)

    def refused(raw, what):
        path.write_bytes(raw)
        s = wpt.scene_from(sd, cache_path=path)
        assert s.from_cache is False, what
        assert s.stats()["max_depth"] == max_depth, what
        for x, y in zip(want, s.intersect(ot, dt)):
            assert torch.equal(torch.nan_to_num(x), torch.nan_to_num(y)), what
        now = path.read_bytes()                                  # and the file has been replaced by a sound one
        assert now[:64] == good[:64] and now[72:] == good[72:], what     # (bytes 64..72: the build time)

    refused(header(max_depth=40) + good[HDR:], "stack depth beyond the traversal stack")
    refused(header(max_depth=max_depth - 1) + good[HDR:], "depth that disagrees with the level table")
    refused(header(tag=tag ^ (1 << 32)) + good[HDR:], "another builder revision")
    refused(good[:-4], "truncated")
    refused(good + b"\0" * 64, "trailing bytes")
    flipped = bytearray(good); flipped[HDR + 64 * (n_blocks // 2) + 17] ^= 0x40
    refused(bytes(flipped), "payload changed, checksum not")
    # forged with a fitting checksum: the root's `base` points past the buffer
    forged = bytearray(blocks); forged[12:16] = struct.pack("<I", n_blocks)
    refused(header(phash=_tree_words_hash([bytes(forged), lnodes, lbegin])) + bytes(forged) + lnodes + lbegin,
            "child run outside the buffer")
    # ... a leaf triangle whose face index is outside the mesh (last block of a breadth-first tree is a triangle)
    forged = bytearray(blocks); forged[64 * (n_blocks - 1) + 48:64 * (n_blocks - 1) + 52] = struct.pack("<i", n_tris)
    refused(header(phash=_tree_words_hash([bytes(forged), lnodes, lbegin])) + bytes(forged) + lnodes + lbegin,
            "face index outside the mesh")
    # ... a level table that does not end at the node count
    lb = np.frombuffer(lbegin, "<u4").copy(); lb[-1] += 1
    refused(header(phash=_tree_words_hash([blocks, lnodes, lb.tobytes()])) + blocks + lnodes + lb.tobytes(),
            "level table does not end at the node count")


def test_depth_shard_refuses_more_ranks_than_pixels():
    """a rank count beyond W*H would leave ranks whose pre-pass slice starts outside the image"""
    import ctypes as C
    import wpt
    sd = scenes.cornell(4)
    r = wpt.Renderer(wpt.scene_from(sd), 3, 2, 1)
    rc = wpt.lib().wpt_renderer_set_depth_shard(r._h, C.c_void_p(1), C.c_int(0), C.c_int(7))
    assert rc == 1 and b"more ranks than pixels" in wpt.lib().wpt_last_error()
    assert wpt.lib().wpt_renderer_set_depth_shard(r._h, None, C.c_int(0), C.c_int(7)) == 0      # NULL comm: switched off
    r.close()


def test_pinned_film_outlives_the_objects_that_made_it():
    """arrays over wpt_host_alloc memory own it — dropping the PinnedArray or the Renderer does not free the
    film under a live view; render_host(copy=False) hands out the pinned film itself, the default a private copy."""
    import gc
    import wpt
    sd = scenes.cornell(8)
    scene = wpt.scene_from(sd)
    cam = wpt.camera_from(sd)
    pin = wpt.PinnedArray((3, 1000))
    row = pin.array[1]
    row[:] = 7.0
    del pin
    gc.collect()
    junk = [wpt.PinnedArray((3, 1000)) for _ in range(8)]       # would reuse the block had it been freed
    for j in junk:
        j.array[:] = -1.0
    assert (row == 7.0).all()
    r = wpt.Renderer(scene, 32, 32, 2)
    a, _ = r.render_host(cam, 0, 4, 4, seed=3)
    live, _ = r.render_host(cam, 0, 4, 4, seed=3, copy=False)
    assert np.array_equal(a, live) and not np.shares_memory(a, live)
    again, _ = r.render_host(cam, 4, 4, 4, seed=3, copy=False)
    assert np.shares_memory(again, live) and not np.array_equal(a, live)      # the next frame overwrote the live film
    keep = live[0]
    want = keep.copy()
    r.close()
    del r, live, again
    gc.collect()
    assert np.array_equal(keep, want)


def test_render_resumes_from_a_film_checkpoint(tmp_path):
    """SURVEY §5 checkpoint/resume: samples 0..5, save, a NEW renderer loads and renders 5..8 on top == samples 0..8 in
    one go — bit for bit here, because every contribution of an all-white scene is exactly 1.0"""
    import torch
    import wpt
    sd = scenes.cornell(16)
    scene = wpt.scene_from(sd)
    cam = wpt.camera_from(sd)
    W, H, depth, seed = 96, 64, 5, 77
    whole = torch.zeros(3, W * H, device="cuda")
    r = wpt.Renderer(scene, W, H, 4)
    r.render(cam, whole, 0, 8, depth, seed=seed)
    r.finish()
    part = torch.zeros(3, W * H, device="cuda")
    r.render(cam, part, 0, 5, depth, seed=seed)
    r.finish()
    r.close()
    wpt.film_save(tmp_path / "f.ckpt", part.cpu().numpy(), W, H, 5, seed)
    del part
    w, h, done, sd_seed = wpt.film_info(tmp_path / "f.ckpt")
    host, done2, _ = wpt.film_load(tmp_path / "f.ckpt", w, h)
    assert (w, h, done, sd_seed, done2) == (W, H, 5, seed, 5)
    cont = torch.from_numpy(host).cuda()
    r2 = wpt.Renderer(wpt.scene_from(sd), W, H, 4)
    r2.render(cam, cont, done, 8 - done, depth, seed=sd_seed)
    r2.finish()
    assert torch.equal(cont, whole) and float(whole.sum()) > 0


def test_reduce_scatter_entry_points_on_a_one_rank_communicator(tmp_path):
    """wpt_film_shard_layout / wpt_film_reduce_scatter / wpt_film_shard_finish / wpt_host_register on ONE GPU (a
    communicator of one rank: the collective is the identity): symbol binding, slice arithmetic, the scale kernel on a
    slice and the DMA into a page-locked file mapping.  The N-rank sum is tests/test_gpu_dist.py's business."""
    import ctypes as C
    import torch
    import wpt
    L = wpt.lib()
    # slice arithmetic, no device needed: 3 * 91 = 273 floats over 4 ranks -> 69, 69, 69, 66; padded to 276
    got = []
    for g in range(4):
        pad, lo, cnt = C.c_uint64(), C.c_uint64(), C.c_uint64()
        assert L.wpt_film_shard_layout(C.c_uint32(91), C.c_int(g), C.c_int(4), C.byref(pad), C.byref(lo), C.byref(cnt)) == 0
        got.append((pad.value, lo.value, cnt.value))
    assert got == [(276, 0, 69), (276, 69, 69), (276, 138, 69), (276, 207, 66)]
    assert L.wpt_film_shard_layout(C.c_uint32(91), C.c_int(4), C.c_int(4), None, None, None) == 1
    idb = (C.c_char * 128)()
    wpt._check(L.wpt_nccl_unique_id(idb))
    comm = C.c_void_p()
    wpt._check(L.wpt_nccl_comm_create(C.c_char_p(idb.raw), C.c_int(1), C.c_int(0), C.c_int(0), C.byref(comm)))
    n = 91
    film = torch.arange(3 * n, dtype=torch.float32, device="cuda") + 1.0
    want = (film * 0.25).cpu().numpy()
    import shutil
    import tempfile
    import wpt_dist
    box = tempfile.mkdtemp(prefix="wpt_test_", dir="/dev/shm")     # a tmpfs: file mappings of a disk cannot be page-locked
    shared = wpt_dist.SharedHostFilm(box, 3 * n, 0, lambda: None)
    host = shared.array
    try:
        wpt._check(L.wpt_film_reduce_scatter(comm, C.c_void_p(film.data_ptr()), C.c_uint32(n), C.c_int(0), C.c_int(1), None))
        wpt._check(L.wpt_film_shard_finish(C.c_void_p(film.data_ptr()), C.c_uint32(n), C.c_int(0), C.c_int(1), C.c_float(0.25),
                                           C.c_void_p(host.ctypes.data), None))
        torch.cuda.synchronize()
        assert np.array_equal(np.asarray(host), want) and np.array_equal(film.cpu().numpy(), want)
    finally:
        shared.close()
        shutil.rmtree(box, ignore_errors=True)
        L.wpt_nccl_comm_destroy(comm)

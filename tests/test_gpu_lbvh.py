"""-m gpu: row f4 — the device-side builder (wpt_scene_commit_lbvh: Morton codes, radix sort, Karras radix tree,
collapse to the 7-wide block layout, geometry by the refit kernels).  The tree differs from the host SAH builder's, the
closest hits must not: parity against the oracle / Embree, identical renders, degenerate inputs, refit afterwards."""
import numpy as np
import pytest

import oracle_py as O
import scenes
import util

pytestmark = pytest.mark.gpu


def _dev(torch, scene, o, d, **kw):
    ot = [torch.from_numpy(np.ascontiguousarray(o[:, k])).cuda() for k in range(3)]
    dt = [torch.from_numpy(np.ascontiguousarray(d[:, k])).cuda() for k in range(3)]
    out = scene.intersect(ot, dt, **kw)
    torch.cuda.synchronize()
    return tuple(x.cpu().numpy() for x in out)


@pytest.mark.parametrize("name", ["cornell", "knob", "terrain", "interior", "spiral"])
def test_lbvh_hits_equal_the_oracle_and_the_sah_tree(name):
    import torch
    import wpt
    sd = {"cornell": lambda: scenes.cornell(30), "knob": lambda: scenes.knob(150, 60),
          "terrain": lambda: scenes.terrain_spheres(81, 6, 10), "interior": lambda: scenes.interior(49, 3, (16, 32)),
          "spiral": lambda: scenes.spiral(60, 0.7)}[name]()
    lb, sah, om = wpt.scene_from(sd, builder="lbvh"), wpt.scene_from(sd), util.oracle_mesh(sd)
    st = lb.stats()
    assert st["n_triangles"] == sd.n_tris and 1 <= st["max_depth"] <= 31 and st["n_nodes"] < sd.n_tris
    for inside in (True, False):
        o, d = util.random_rays(sd, 200_000, seed=3 + inside, inside=inside)
        got = _dev(torch, lb, o, d)
        ref, robust = util.embree_pair(om, o, d)
        assert util.compare_hits(ref, got, f"lbvh/{name}", robust=robust) >= 0.9999
        # the two trees give the SAME closest hit, tie rule included (lowest primitive): bit-identical records
        other = _dev(torch, sah, o, d)
        for a, b in zip(got, other):
            assert np.array_equal(a, b, equal_nan=True)


def test_lbvh_render_equals_the_sah_render():
    import torch
    import wpt
    sd = scenes.knob(120, 50)
    cam = wpt.camera_from(sd)
    W, H = 160, 128
    films = []
    for builder in ("sah", "lbvh"):
        scene = wpt.scene_from(sd, builder=builder)
        r = wpt.Renderer(scene, W, H, 8)
        f = torch.zeros(3, W * H, device="cuda")
        r.render(cam, f, 0, 16, 6, seed=12)
        films.append((f, r.finish()))
    assert films[0][1]["live"] == films[1][1]["live"]
    assert torch.allclose(films[0][0], films[1][0], rtol=0, atol=2e-5)


def test_lbvh_degenerate_inputs_and_refit():
    import torch
    import wpt
    rng = np.random.default_rng(0)

    def build(pos, idx):
        s = wpt.Scene()
        s.add_triangle_mesh(np.asarray(pos, np.float32), np.asarray(idx, np.int32))
        s.commit(0, "lbvh")
        return s

    tri = [[0, 0, 0], [1, 0, 0], [0, 1, 0]]
    o = np.array([[0.2, 0.2, -1], [2, 2, -1]], np.float32)
    d = np.array([[0, 0, 1], [0, 0, 1]], np.float32)
    for n in (1, 2, 3, 4, 37):                      # coincident copies: equal Morton codes, ties by position
        pos = np.tile(np.array(tri, np.float32), (n, 1))
        idx = np.arange(3 * n).reshape(-1, 3)
        s = build(pos, idx)
        hid, t, u, v = _dev(torch, s, o, d, tmin=0.0)
        assert hid.tolist() == [0, -1] and abs(t[0] - 1) < 1e-6          # lowest primitive of the coincident stack
    # a flat axis-aligned grid (zero extent on one axis) and a huge + tiny mix, against the oracle
    g = np.stack(np.meshgrid(np.arange(12.0), np.arange(12.0), indexing="ij"), -1).reshape(-1, 2)
    flat_pos = np.concatenate([g[:, :1], np.zeros((len(g), 1)), g[:, 1:]], 1)
    ii, jj = np.meshgrid(np.arange(11), np.arange(11), indexing="ij")
    a = (ii * 12 + jj).reshape(-1)
    flat_idx = np.concatenate([np.stack([a, a + 12, a + 13], 1), np.stack([a, a + 13, a + 1], 1)])
    c = rng.uniform(-1, 1, (300, 1, 3))
    small = (c + rng.normal(0, 1e-4, (300, 3, 3))).reshape(-1, 3)
    big = np.array([[-50, -2, -50], [50, -2, -50], [0, -2, 80]])
    mix_pos = np.concatenate([small, big])
    mix_idx = np.concatenate([np.arange(900).reshape(-1, 3), [[900, 901, 902]]])
    for pos, idx in ((flat_pos, flat_idx), (mix_pos, mix_idx)):
        sd = scenes.SceneData("t", np.asarray(pos, np.float32), np.asarray(idx, np.int32), np.zeros((1, 3)),
                              np.zeros_like(np.asarray(idx, np.int32)), np.zeros(len(idx)), [[1, 1, 1]], {})
        s = build(pos, idx)
        om = util.oracle_mesh(sd)
        oo, dd = util.random_rays(sd, 50_000, seed=2, inside=False)
        assert util.compare_hits(om.intersect("bvh", oo, dd), _dev(torch, s, oo, dd), "lbvh/degenerate") >= 0.9999
    # refit of an LBVH-built tree (same kernels, same level lists) after the vertices moved
    sd = scenes.cornell(16)
    s = wpt.scene_from(sd, builder="lbvh")
    moved = (sd.pos * np.float32(1.1) + np.float32(0.05)).astype(np.float32)
    s.refit(moved)
    sd2 = scenes.SceneData("m", moved, sd.idx, sd.nrm, sd.nidx, sd.mat, sd.albedo, sd.camera)
    oo, dd = util.random_rays(sd2, 100_000, seed=8)
    assert util.compare_hits(util.oracle_mesh(sd2).intersect("bvh", oo, dd), _dev(torch, s, oo, dd), "lbvh/refit") >= 0.9999


def test_lbvh_full_size_build_time_and_tree():
    """C3's 1 M triangles: built on the device in milliseconds; hits on 2^18 camera rays against Embree"""
    import torch
    import wpt
    gen, W, H, spp, depth = scenes.CONFIGS["C3"]
    sd = gen()
    s = wpt.scene_from(sd, builder="lbvh")      # first build of the process pays CUB's / the context's one-off costs
    s2 = wpt.scene_from(sd, builder="lbvh")
    st = s2.stats()
    assert st["build_ms"] < 200.0, st["build_ms"]          # host SAH: ~300-900 ms
    assert st["max_depth"] <= 31
    cam = wpt.camera_from(sd)
    o_t, d_t = wpt.raygen(cam, W, H, 0, seed=5)
    pick = torch.arange(0, W * H, 8, device="cuda")
    o_t = [x[pick].contiguous() for x in o_t]; d_t = [x[pick].contiguous() for x in d_t]
    hit = s2.intersect(o_t, d_t)
    torch.cuda.synchronize()
    o = np.stack([x.cpu().numpy() for x in o_t], 1); d = np.stack([x.cpu().numpy() for x in d_t], 1)
    ref, robust = util.embree_pair(util.oracle_mesh(sd), o, d)
    assert util.compare_hits(ref, tuple(x.cpu().numpy() for x in hit), "lbvh/C3", robust=robust) >= 0.9999
    print("LBVH C3: build_ms", st["build_ms"], "nodes", st["n_nodes"], "depth", st["max_depth"])

"""Row f2 of SURVEY §8: OBJ/MTL ingestion (wpt_obj_load) against the reference's own loader chain —
its vendored tinyobjloader flattened like load_meshes (src/main.cpp:143-223).  CPU tests compare with
the committed golden arrays (tests/golden/obj/*.npz, made by tests/golden/make_obj_golden.py) and,
where oracle/_ref/libtinyobj_ref.so exists, with the real thing on random polygon soups and on the
loader's own model zoo; the GPU test pushes a loaded file through the hot path."""
import glob
import os

import numpy as np
import pytest

import oracle_py as O
import wpt

HERE = os.path.dirname(os.path.abspath(__file__))
OBJ = os.path.join(HERE, "golden", "obj")
KEYS = ("positions", "normals", "texcoords", "position_triplets", "normal_triplets", "texcoord_triplets", "material_id")


def _bits(a):
    return a.view(np.uint32) if a.dtype == np.float32 else a


def _same(mine, ref, what):
    ref = dict(ref)
    # documented deviation: a normal index past the end of the `vn` list — which the
    # reference would gather out of bounds — is loaded as "no normal" (-1)
    nt = np.array(ref["normal_triplets"])
    nt[nt >= len(ref["normals"])] = -1
    ref["normal_triplets"] = nt
    for k in KEYS:
        assert mine[k].shape == ref[k].shape, (what, k, mine[k].shape, ref[k].shape)
        assert np.array_equal(_bits(mine[k]), _bits(ref[k])), (what, k)     # bit-exact, floats included
    assert np.array_equal(_bits(mine["albedo"][1:]), _bits(np.asarray(ref["kd"], np.float32))), what


@pytest.mark.parametrize("name", ["box_quads", "polys"])
def test_matches_golden(name):
    m = wpt.load_meshes(os.path.join(OBJ, name + ".obj"), use_kd=True)
    g = np.load(os.path.join(OBJ, name + ".npz"))
    _same(m, g, name)
    assert np.all(m["albedo"][0] == 1.0)                       # default material, src/main.cpp:234


def test_reference_semantics():
    m = wpt.load_meshes(os.path.join(OBJ, "box_quads.obj"))
    assert np.all(m["albedo"] == 1.0)                          # every material white as shipped, src/main.cpp:244
    assert m["material_names"] == ["", "white", "red", "green", "white"]
    # 5 quads -> 10 tris, then pyramid: unknown usemtl -> default material 0; duplicate name -> first wins
    assert m["material_id"].tolist() == [1, 1, 1, 1, 1, 1, 2, 2, 3, 3, 0, 0, 1, 1, 1, 1]
    assert m["position_triplets"][6].tolist() == [0, 4, 7]     # negative indices are relative to the count so far
    p = wpt.load_meshes(os.path.join(OBJ, "polys.obj"))
    assert p["normals"].shape == (0, 3) and np.all(p["normal_triplets"] == -1)
    assert p["positions"][13].tolist() == [10.0, -2.5, 2.0]    # 1e1 -.25E+1 +2


def test_errors(tmp_path):
    with pytest.raises(wpt.WptError) as e:
        wpt.load_meshes(tmp_path / "nope.obj")
    assert e.value.code == 5
    bad = tmp_path / "zero.obj"
    bad.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 0 1 2\n")   # index 0 is not allowed (tiny_obj_loader.h:752)
    with pytest.raises(wpt.WptError) as e:
        wpt.load_meshes(bad)
    assert e.value.code == 1
    oob = tmp_path / "oob.obj"
    oob.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 9\n")    # the reference would gather out of bounds
    with pytest.raises(wpt.WptError):
        wpt.load_meshes(oob)
    empty = tmp_path / "empty.obj"
    empty.write_text("# nothing\nv 0 0 0\n")
    with pytest.raises(wpt.WptError):
        wpt.load_meshes(empty)


def _random_obj(path, rng, n_poly):
    """planar convex and star-shaped polygons with 3..9 corners, random orientation, mixed corner forms"""
    lines, nv = ["mtllib r.mtl"], 0
    (path.parent / "r.mtl").write_text("".join(f"newmtl m{i}\nKd {rng.random():.6f} {rng.random():.5g} {rng.random():.3e}\n"
                                               for i in range(4)))
    lines += ["vn 0 1 0", "vt 0.5 0.25"]
    for i in range(n_poly):
        k = int(rng.integers(3, 10))
        a, b = np.linalg.qr(rng.normal(size=(3, 3)))[0][:2]
        c = rng.uniform(-5, 5, 3)
        ang = np.sort(rng.uniform(0, 2 * np.pi, k))
        rad = rng.uniform(0.3, 1.0, k) if i % 2 else np.full(k, 1.0)
        for t, r in zip(ang, rad):
            p = c + r * (np.cos(t) * a + np.sin(t) * b)
            fmt = ("{:.6f}", "{:.9g}", "{:.4e}")[i % 3]
            lines.append("v " + " ".join(fmt.format(x) for x in p))
        if i % 5 == 0:
            lines.append(f"usemtl m{int(rng.integers(0, 5))}")
        if i % 7 == 0:
            lines.append(f"g grp{i}")
        form = i % 4
        idx = [nv + 1 + j for j in range(k)] if i % 2 else [-(k - j) for j in range(k)]
        lines.append("f " + " ".join((f"{j}", f"{j}/1", f"{j}//1", f"{j}/1/1")[form] for j in idx))
        nv += k
    path.write_text("\n".join(lines) + "\n")


@pytest.mark.skipif(not O.ref_obj_available(), reason="oracle/_ref/libtinyobj_ref.so not built")
def test_random_polygon_soup_vs_tinyobjloader(tmp_path):
    rng = np.random.default_rng(7)
    for trial in range(3):
        p = tmp_path / f"soup{trial}.obj"
        _random_obj(p, rng, 400)
        _same(wpt.load_meshes(p, use_kd=True), O.ref_load_obj(p), p.name)


ZOO = "/root/reference/dependencies/tinyobjloader/models"


@pytest.mark.skipif(not (O.ref_obj_available() and os.path.isdir(ZOO)), reason="reference tree absent")
def test_model_zoo_vs_tinyobjloader():
    """every model of the loader's own test-suite that it accepts and that holds faces"""
    n = 0
    for p in sorted(glob.glob(os.path.join(ZOO, "*.obj"))):
        try:
            ref = O.ref_load_obj(p)
        except RuntimeError:
            continue
        if ref["material_id"].size == 0 or ref["position_triplets"].min() < 0 or \
                ref["position_triplets"].max() >= len(ref["positions"]):
            continue
        _same(wpt.load_meshes(p, use_kd=True), ref, os.path.basename(p))
        n += 1
    assert n >= 20


@pytest.mark.gpu
def test_loaded_obj_through_the_hot_path():
    import torch
    path = os.path.join(OBJ, "box_quads.obj")
    m = wpt.load_meshes(path, use_kd=True)
    scene = wpt.scene_from_obj(path, 0, use_kd=True)
    om = O.Mesh(m["positions"], m["position_triplets"], m["normals"], m["normal_triplets"], m["material_id"], m["albedo"])
    rng = np.random.default_rng(3)
    o = rng.uniform(-0.9, 0.9, (50000, 3)).astype(np.float32) + np.float32([0, 1, 0])
    d = rng.normal(size=(50000, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    got = scene.intersect_host(o, d)
    ref = om.intersect("bvh", o, d)
    assert (got[0] == ref[0]).mean() >= 0.9999
    hit = (ref[0] >= 0) & (got[0] == ref[0])
    assert np.allclose(got[1][hit], ref[1][hit], rtol=1e-5, atol=0)
    # same-RNG render: image equal to the oracle's but for t-tie pixels
    W = H = 48
    cam = wpt.ThinlensCamera((0, 1, 3.2), (0, 1, 0), (0, 1, 0), 0.0, 1.0, 0.8)
    r = wpt.Renderer(scene, W, H, 2)
    film, st = r.render_host(cam, 0, 4, 5, seed=11)
    ocam = O.camera((0, 1, 3.2), (0, 1, 0), (0, 1, 0), fov_y=0.8)
    ofilm, oseg = om.render("bvh", ocam, W, H, 0, 4, 5, rng_mode=O.RNG_PHILOX, seed=11)
    ofilm /= 4
    assert (np.abs(film - ofilm).max(0) > 1e-5).mean() < 0.02
    assert film[0].max() > 0 and not np.allclose(film[0], film[1])      # Kd colours reach the film

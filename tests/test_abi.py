"""The C-ABI library loads without a GPU and exports every symbol include/wpt.h declares; the host-only
entry points behave; compute entry points fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "wpt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(wpt_[a-z0-9_]+)\s*\(", src)))


def test_exports_match_header():
    import wpt
    L = wpt.lib()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"libwpt.so does not export {n}"
    assert L.wpt_version() == 200


def test_host_only_entry_points(tmp_path):
    import wpt
    cam = wpt.ThinlensCamera.reference_default()
    z = np.array(cam.c.z)
    assert np.allclose(z, np.array([0, -3, -5.0]) / np.linalg.norm([0, -3, -5.0]), atol=1e-6)
    with pytest.raises(wpt.WptError):
        wpt.ThinlensCamera((0, 0, 0), (0, 0, 0), (0, 1, 0), 0, 1, 0.7)      # degenerate
    r = np.arange(6, dtype=np.float32)
    wpt.Fimage.save_pfm(r, r + 10, r + 20, 3, 2, tmp_path / "a.pfm")
    raw = (tmp_path / "a.pfm").read_bytes()
    assert raw.startswith(b"PF\n3 2\n-1\n")
    px = np.frombuffer(raw[len(b"PF\n3 2\n-1\n"):], np.float32).reshape(2, 3, 3)
    assert px[0, 0].tolist() == [3.0, 13.0, 23.0]       # first row written is the LAST image row (fimage.h:49)
    with pytest.raises(wpt.WptError):
        wpt.Fimage.save_pfm(r, r, r, 3, 2, "/nonexistent-dir/x.pfm")


def test_scene_validation_and_no_cpu_fallback():
    import torch
    import wpt
    s = wpt.Scene()
    with pytest.raises(wpt.WptError):
        s.add_triangle_mesh(np.zeros((3, 3)), np.zeros((1, 3)), material_id=[4])     # material out of range
    s = wpt.Scene()
    s.add_triangle_mesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]])
    with pytest.raises(wpt.WptError):
        s.stats()                                     # not committed
    if not torch.cuda.is_available():
        with pytest.raises(wpt.WptError) as e:
            s.commit(0)
        assert "no CUDA device" in str(e.value) and e.value.code == 2


def test_product_does_not_reach_the_oracle():
    """the product path must not import/link anything under oracle/"""
    pkg = os.path.join(ROOT, "smallpt-enoki-optix_b200")
    for dp, _, fs in os.walk(pkg):
        if "build" in dp:
            continue
        for f in fs:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh", "Makefile")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle_py" not in txt and "wpt_oracle" not in txt and "embree_ref" not in txt, f
    import subprocess
    out = subprocess.run(["ldd", os.path.join(pkg, "libwpt.so")], capture_output=True, text=True).stdout
    assert "oracle" not in out and "embree" not in out


def test_film_checkpoint_roundtrip_and_damage(tmp_path):
    """wpt_film_save / wpt_film_info / wpt_film_load are host-only: round trip bit for bit, atomic replace, and a
    truncated or altered file, or one of another size, is refused."""
    import wpt
    rng = np.random.default_rng(3)
    W, H = 7, 5
    film = rng.uniform(0, 9, (3, W * H)).astype(np.float32)
    p = tmp_path / "frame.ckpt"
    wpt.film_save(p, film, W, H, 37, seed=0xDEADBEEFCAFE)
    assert wpt.film_info(p) == (W, H, 37, 0xDEADBEEFCAFE)
    back, spp, seed = wpt.film_load(p, W, H)
    assert spp == 37 and seed == 0xDEADBEEFCAFE and np.array_equal(back.view(np.uint32), film.view(np.uint32))
    wpt.film_save(p, film * 2, W, H, 74)                      # replaces the file as a whole
    assert wpt.film_info(p)[2] == 74 and not list(tmp_path.glob("*.tmp*"))
    with pytest.raises(wpt.WptError) as e:
        wpt.film_load(p, H, W)
    assert e.value.code == 1 and "7x5" in str(e.value)
    raw = p.read_bytes()
    for bad in (raw[:-5], raw[:40] + bytes([raw[40] ^ 1]) + raw[41:], b"PF\n7 5\n-1\n" + raw[11:]):
        p.write_bytes(bad)
        with pytest.raises(wpt.WptError) as e:
            wpt.film_load(p, W, H)
        assert e.value.code == 5
    with pytest.raises(wpt.WptError):
        wpt.film_info(tmp_path / "absent.ckpt")

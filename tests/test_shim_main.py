"""include/wpt_shim.hpp + examples/wurst_main.cpp: the reference's main() (src/main.cpp:355-445) on the
C ABI.  CPU: it builds against the public headers only and fails loudly without a device.  GPU: the PFM
it writes is the image the Python mirror renders for the same inputs (same library, same path)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "smallpt-enoki-optix_b200")
WURST = os.path.join(PKG, "wurst")
OBJ = os.path.join(ROOT, "tests", "golden", "obj", "box_quads.obj")


def _read_pfm(path):
    raw = open(path, "rb").read()
    parts = raw.split(b"\n", 3)
    assert parts[0] == b"PF" and parts[2] == b"-1"
    w, h = (int(x) for x in parts[1].split())
    px = np.frombuffer(parts[3], np.float32).reshape(h, w, 3)[::-1]      # rows are stored bottom-up (fimage.h:49)
    return np.ascontiguousarray(px.transpose(2, 0, 1)).reshape(3, -1)


def test_shim_compiles_from_public_headers_only(tmp_path):
    src = tmp_path / "t.cpp"
    src.write_text('#include "wpt_shim.hpp"\nint main(){ wpt_shim::Real3 a(1,2,3); wpt_shim::StopWatch s; s.reset();'
                   ' try { wpt_shim::Scene sc; sc.add_triangle_mesh("/nonexistent.obj"); } catch (const std::runtime_error&)'
                   ' { return 0; } return 1; }\n')
    exe = tmp_path / "t"
    subprocess.run(["g++", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src), "-L", PKG, "-lwpt",
                    f"-Wl,-rpath,{PKG}"], check=True)
    assert subprocess.run([str(exe)]).returncode == 0           # the C code became a std::runtime_error


def test_main_fails_loudly_without_a_device(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("has a device")
    assert os.path.exists(WURST), "make -C smallpt-enoki-optix_b200 builds it"
    r = subprocess.run([WURST, OBJ, str(tmp_path / "o.pfm"), "16", "16", "1", "2"], capture_output=True, text=True)
    assert r.returncode == 1 and "no CUDA device" in r.stderr and not (tmp_path / "o.pfm").exists()


@pytest.mark.gpu
def test_main_writes_the_same_image_as_the_python_mirror(tmp_path):
    import wpt
    out = tmp_path / "wurst.pfm"
    r = subprocess.run([WURST, OBJ, str(out), "64", "48", "8", "3", "--kd"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert r.stdout.strip().endswith("done") and "writing image" in r.stdout
    img = _read_pfm(out)
    scene = wpt.scene_from_obj(OBJ, 0, use_kd=True)
    film, st = wpt.Renderer(scene, 64, 48, 4).render_host(wpt.ThinlensCamera.reference_default(), 0, 8, 3, seed=0)
    # same kernels, same counters; film accumulation order (float atomics) is the only freedom
    assert np.allclose(img, film, rtol=0, atol=2e-6)
    assert img.max() > 0

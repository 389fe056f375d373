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


def test_main_checks_its_command_line(tmp_path):
    """flags may stand anywhere and are not read as positional values; sizes and the GPU count are checked
    before anything is loaded (exit code 2, nothing written)."""
    out = tmp_path / "o.pfm"
    for argv, needle in (([OBJ, str(out), "--gpus"], "needs a count"),
                         ([OBJ, str(out), "0", "16", "1", "2"], "positive"),
                         ([OBJ, str(out), "16", "16", "1", "2", "7"], "positional"),
                         ([OBJ, str(out), "--frobnicate"], "usage"),
                         ([OBJ, str(out), "--gpus", "4096", "16", "16", "1", "2"], "CUDA device"),
                         ([OBJ, str(out), "--gpus", "0"], "CUDA device")):
        r = subprocess.run([WURST] + argv, capture_output=True, text=True)
        assert r.returncode == 2 and needle in r.stderr, (argv, r.returncode, r.stderr)
    assert not out.exists()


@pytest.mark.gpu
def test_main_takes_flags_before_the_positionals(tmp_path):
    """`wurst scene.obj out.pfm --gpus 1 --kd 32 24 4 2` used to read width = atoi("--gpus")"""
    a, b = tmp_path / "a.pfm", tmp_path / "b.pfm"
    for out, argv in ((a, [OBJ, str(a), "32", "24", "4", "2", "--kd", "--gpus", "1"]),
                      (b, ["--kd", OBJ, str(b), "--gpus", "1", "32", "24", "4", "2"])):
        r = subprocess.run([WURST] + argv, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    ia, ib = _read_pfm(a), _read_pfm(b)
    assert ia.shape == ib.shape == (3, 32 * 24) and np.allclose(ia, ib, rtol=0, atol=2e-6)


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


def test_shim_frame3_matches_the_oracle():
    """Frame3 of the shim (src/coordframe.h:5-54, host scalar form) against the oracle's restatement: same matrix,
    to_world / to_local bit for bit"""
    import ctypes as C
    import oracle_py as O
    src = os.path.join(ROOT, "tests", "_frame3_probe.cpp")
    exe = os.path.join(PKG, "build", "frame3_probe")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    open(src, "w").write('#include <cstdio>\n#include <cstdlib>\n#include "wpt_shim.hpp"\nint main(int c, char **v){ using namespace wpt_shim;'
                         ' Real3 n(atof(v[1]),atof(v[2]),atof(v[3])), x(atof(v[4]),atof(v[5]),atof(v[6])); Frame3 f(n);'
                         ' Real3 w=f.to_world(x), l=f.to_local(x); for(int k=0;k<9;k++) printf("%a ",f.m_world_from_local[k]);'
                         ' printf("%a %a %a %a %a %a\\n",w.x,w.y,w.z,l.x,l.y,l.z); return 0; }\n')
    try:
        subprocess.run(["g++", "-std=c++17", "-O1", "-ffp-contract=off", "-I", os.path.join(ROOT, "include"), "-o", exe, src,
                        "-L", PKG, "-lwpt", f"-Wl,-rpath,{PKG}"], check=True)
    finally:
        os.remove(src)
    L = O.lib()
    rng = np.random.default_rng(3)
    for _ in range(20):
        n = rng.normal(size=3).astype(np.float32); n /= np.linalg.norm(n)
        x = rng.normal(size=3).astype(np.float32)
        out = subprocess.run([exe] + [float(a).hex() for a in n] + [float(a).hex() for a in x], capture_output=True, text=True).stdout.split()
        got = np.array([float.fromhex(t) for t in out], np.float32)
        m = (C.c_float * 9)(); w = (C.c_float * 3)(); l = (C.c_float * 3)()
        L.orc_frame_from_normal((C.c_float * 3)(*n), m)
        L.orc_frame_to_world(m, (C.c_float * 3)(*x), w)
        L.orc_frame_to_local(m, (C.c_float * 3)(*x), l)
        assert np.array_equal(got[:9], np.array(m, np.float32))
        assert np.array_equal(got[9:12], np.array(w, np.float32)) and np.array_equal(got[12:], np.array(l, np.float32))


@pytest.mark.gpu
def test_main_sharded_over_two_gpus_writes_the_same_image(tmp_path):
    """wurst --gpus 2: ShardedPathLoop (spp shards + ONE wpt_film_reduce, no Python) == the single-GPU image"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run under gpurun --gpus 2)")
    a, b = tmp_path / "one.pfm", tmp_path / "two.pfm"
    for out, extra in ((a, []), (b, ["--gpus", "2"])):
        r = subprocess.run([WURST, OBJ, str(out), "64", "48", "8", "3", "--kd"] + extra, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
    assert np.allclose(_read_pfm(a), _read_pfm(b), rtol=0, atol=2e-6)


@pytest.mark.gpu
def test_shim_scene_intersect_returns_the_interaction(tmp_path):
    """Scene::intersect of the shim fills an Interaction (src/main.cpp:255-284, 320-340): hit record, position,
    normals, texcoord and material per lane — checked against the Python mirror on the same rays"""
    src = tmp_path / "t.cpp"
    src.write_text('''#include <cstdio>
#include "wpt_shim.hpp"
using namespace wpt_shim;
int main(int, char **v) {
    Scene sc; sc.add_triangle_mesh(v[1], true); sc.commit();
    const uint32_t n = 4096;
    std::vector<float> o[3], d[3];
    for (uint32_t i = 0; i < n; i++) { o[0].push_back(-0.9f + 1.8f * (i % 64) / 63.f); o[1].push_back(0.1f + 1.8f * (i / 64) / 63.f); o[2].push_back(4.f);
                                       d[0].push_back(0.f); d[1].push_back(0.f); d[2].push_back(-1.f); }
    RealC ox = RealC::copy(o[0].data(), n), oy = RealC::copy(o[1].data(), n), oz = RealC::copy(o[2].data(), n);
    RealC dx = RealC::copy(d[0].data(), n), dy = RealC::copy(d[1].data(), n), dz = RealC::copy(d[2].data(), n);
    Ray3C ray{ox.data(), oy.data(), oz.data(), dx.data(), dy.data(), dz.data(), n};
    Interaction it(n);
    sc.intersect(ray, nullptr, 0, it);
    auto id = it.m_hit.m_tri_id.to_host(); auto pz = it.m_hit.m_position[2].to_host(); auto gz = it.m_hit.m_geometry_normal[2].to_host();
    auto gx = it.m_hit.m_geometry_normal[0].to_host(); auto gy = it.m_hit.m_geometry_normal[1].to_host();
    auto mat = it.m_scatter.to_host(); auto t = it.m_hit.m_t.to_host();
    int hits = 0; double err = 0;
    for (uint32_t i = 0; i < n; i++) if (id[i] >= 0) { hits++; err += std::fabs(pz[i] - (4.f - t[i])) + (mat[i] < 0) + (std::fabs(std::sqrt(gx[i] * gx[i] + gy[i] * gy[i] + gz[i] * gz[i]) - 1.f) > 1e-5f); }
    std::printf("%d %g\\n", hits, err);
    return 0; }
''')
    exe = tmp_path / "t"
    subprocess.run(["g++", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src), "-L", PKG, "-lwpt",
                    f"-Wl,-rpath,{PKG}"], check=True)
    r = subprocess.run([str(exe), OBJ], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    hits, err = r.stdout.split()
    assert int(hits) > 300 and float(err) < 1e-3

#!/usr/bin/env python
"""CPU-side traversal statistics of the product's BVH (test-only host library): node visits and
triangle tests per ray for camera rays and for diffuse bounce rays of a workload."""
import ctypes as C, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("smallpt-enoki-optix_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import scenes, util, oracle_py as O

def stats(H, sd, o, d):
    n = o.shape[0]
    cols = [np.ascontiguousarray(o[:, k]) for k in range(3)] + [np.ascontiguousarray(d[:, k]) for k in range(3)]
    hid = np.full(n, -2, np.int32); t = np.full(n, np.nan, np.float32); u = t.copy(); v = t.copy()
    cnt = (C.c_double * 4)()
    rc = H.wpt_hosttest_trace_stats(sd.pos.ctypes.data_as(C.c_void_p), C.c_uint32(sd.pos.shape[0]),
        sd.idx.ctypes.data_as(C.c_void_p), C.c_uint32(sd.n_tris), *[c.ctypes.data_as(C.c_void_p) for c in cols],
        C.c_float(1e-3), C.c_float(1e20), C.c_uint32(n), hid.ctypes.data_as(C.c_void_p), t.ctypes.data_as(C.c_void_p),
        u.ctypes.data_as(C.c_void_p), v.ctypes.data_as(C.c_void_p), cnt)
    assert rc == 0
    return cnt[0] / n, cnt[1] / n, hid, t

def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C3"
    H = C.CDLL(os.path.join(ROOT, "smallpt-enoki-optix_b200", "libwpt_hosttest.so"))
    gen, W, Hh, spp, depth = scenes.CONFIGS[name]
    sd = gen()
    rc, st = util.hosttest_build(H, sd)
    print(name, sd.n_tris, "tris:", {k: (round(v, 3) if isinstance(v, float) else v) for k, v in st.items()})
    # camera rays on a coarse pixel grid
    cam = util.oracle_camera(sd)
    L = O.lib()
    rng = np.random.default_rng(1)
    n = 200_000
    px = rng.integers(0, W, n); py = rng.integers(0, Hh, n)
    o = np.zeros((n, 3), np.float32); d = np.zeros((n, 3), np.float32)
    po, pd = (C.c_float * 3)(), (C.c_float * 3)()
    xi = (C.c_float * 4)(0.5, 0.5, 0.5, 0.5); xi2 = (C.c_float * 2)(0.5, 0.5)
    for i in range(n):
        L.orc_camera_sample_pos(C.byref(cam), xi, po)
        L.orc_camera_sample_dir(C.byref(cam), C.c_int(int(px[i])), C.c_int(int(py[i])), po, C.c_int(W), C.c_int(Hh), xi2, pd)
        o[i] = po[:]; d[i] = pd[:]
    nv, tt, hid, t = stats(H, sd, o, d)
    print(f"  camera rays : {nv:6.2f} nodes/ray {tt:6.2f} tris/ray  hit rate {np.mean(hid>=0):.3f}")
    # bounce rays: from hit points, cosine-ish random directions in the upper hemisphere of +y-ish normal
    hit = hid >= 0
    p = o[hit] + t[hit, None] * d[hit]
    nd = rng.normal(size=p.shape).astype(np.float32); nd /= np.linalg.norm(nd, axis=1, keepdims=True)
    nd[:, 1] = np.abs(nd[:, 1])
    nv, tt, hid2, t2 = stats(H, sd, p.astype(np.float32), nd)
    print(f"  bounce rays : {nv:6.2f} nodes/ray {tt:6.2f} tris/ray  hit rate {np.mean(hid2>=0):.3f}")

if __name__ == "__main__":
    main()

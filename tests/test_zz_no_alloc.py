"""wpt_intersect_host_static, the closest-hit entry point that allocates nothing (include/wpt.h).

Who needs it: a host that has to keep cudaMalloc out of its bounce loop — a plugin living inside an application whose
pool allocator already owns the whole GPU (a raw cudaMalloc then fails although the pool has room), a frame loop that
cannot take the device-wide wait of cudaMalloc / cudaFree, or a thread that works while another one records a CUDA
graph.  For them the staging buffers wpt_intersect_host allocates per call sit in a fixed area inside libwpt.so, big
enough for WPT_STATIC_MAX_RAYS rays.

What is checked, through the C ABI (ctypes on libwpt.so) and on two calls in a row: the new entry point returns the
same bits as wpt_intersect_host for the same host buffers.  The area is shared by all calls and keeps the last call's
contents, so the second call is the smaller one, masked, with per-ray intervals: anything left over from the first call
would show up in its untouched lanes.  Batches beyond the limit must be refused without touching the outputs.

This file is named to be collected last: `pytest -x` stops at the first failure and a failure here must not hide the
parity tests.  It passes alone in a fresh process (the first call on the GPU does the address look-up) and after the
rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

import scenes
import util

pytestmark = pytest.mark.gpu

WPT_STATIC_MAX_RAYS = 1 << 18      # include/wpt.h
WPT_E_LIMIT = 4


@pytest.fixture(scope="module")
def cornell():
    import torch
    import wpt
    assert torch.cuda.is_available()
    sd = scenes.cornell(50)
    return sd, wpt.scene_from(sd)


def _closest_hits(entry, scene, o, d, mask=None, tmin=None, tmax=None):
    """One call of `entry` (either host-buffer entry point) on host arrays.  Returns (status, tri_id, t, u, v); the
    result arrays start as -2 / NaN so that lanes the library leaves alone can be told apart."""
    import wpt
    n = o.shape[0]
    cols = [np.ascontiguousarray(o[:, k]) for k in range(3)] + [np.ascontiguousarray(d[:, k]) for k in range(3)]
    rs = wpt.RayStreams(*[c.ctypes.data for c in cols], None, None, 1e-3, 1e20)
    if tmin is not None:
        rs.tmin, rs.tmax = tmin.ctypes.data, tmax.ctypes.data
    hid = np.full(n, -2, np.int32)
    t, u, v = (np.full(n, np.nan, np.float32) for _ in range(3))
    hs = wpt.HitStreams(hid.ctypes.data, t.ctypes.data, u.ctypes.data, v.ctypes.data)
    mptr, msz = (None, 0) if mask is None else (C.c_void_p(mask.ctypes.data), mask.size)
    L = wpt.lib()
    args = [scene._h, C.byref(rs), mptr, C.c_uint32(msz), C.c_uint32(n), C.byref(hs)]
    if entry == "static":
        rc = L.wpt_intersect_host_static(*args, C.c_void_p(None))
    else:
        rc = L.wpt_intersect_host(*args)
    return rc, hid, t, u, v


def _same_bits(a, b, what):
    assert a[0] == 0 and b[0] == 0, (what, a[0], b[0])
    assert np.array_equal(a[1], b[1]), f"{what}: tri_id differs on {(a[1] != b[1]).sum()} rays"
    for k, name in ((2, "t"), (3, "u"), (4, "v")):
        assert np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32)), f"{what}: {name} differs"


def test_static_entry_point_equals_intersect_host_on_two_calls_in_a_row(cornell):
    sd, scene = cornell

    # call 1: a full staging area, every ray live
    o1, d1 = util.random_rays(sd, WPT_STATIC_MAX_RAYS, seed=31)
    got1 = _closest_hits("static", scene, o1, d1)
    ref1 = _closest_hits("host", scene, o1, d1)
    _same_bits(got1, ref1, "call 1")
    assert (got1[1] >= 0).mean() > 0.5                  # a closed box: most rays hit something

    # call 2, straight after: fewer rays, a third masked out, per-ray intervals that cut some hits away
    n2 = 70_001
    o2, d2 = util.random_rays(sd, n2, seed=32)
    rng = np.random.default_rng(33)
    mask = (np.arange(n2) % 3 != 0).astype(np.uint8)
    tmin = rng.uniform(0.0, 0.2, n2).astype(np.float32)
    tmax = rng.uniform(0.3, 3.0, n2).astype(np.float32)
    got2 = _closest_hits("static", scene, o2, d2, mask, tmin, tmax)
    ref2 = _closest_hits("host", scene, o2, d2, mask, tmin, tmax)
    _same_bits(got2, ref2, "call 2")
    off = mask == 0
    assert (got2[1][off] == -2).all() and np.isnan(got2[2][off]).all()      # masked lanes: the caller's values, not call 1's
    assert (got2[1][~off] >= -1).all() and (got2[1][~off] == -1).any() and (got2[1][~off] >= 0).any()
    miss = got2[1] == -1
    assert np.isnan(got2[2][miss]).all()                                    # t, u, v are written on hits only

    # a broadcast mask of one zero byte switches every lane off
    none = _closest_hits("static", scene, o2, d2, np.zeros(1, np.uint8))
    assert none[0] == 0 and (none[1] == -2).all()


def test_static_entry_point_refuses_batches_beyond_its_staging_area(cornell):
    import wpt
    sd, scene = cornell
    n = WPT_STATIC_MAX_RAYS + 1
    o, d = util.random_rays(sd, n, seed=34)
    rc, hid, t, u, v = _closest_hits("static", scene, o, d)
    assert rc == WPT_E_LIMIT
    msg = wpt.lib().wpt_last_error().decode()
    assert str(WPT_STATIC_MAX_RAYS) in msg and "WPT_STATIC_MAX_RAYS" in msg
    assert (hid == -2).all() and np.isnan(t).all()       # nothing was done
    empty = _closest_hits("static", scene, o[:0], d[:0])
    assert empty[0] == 0

"""shared helpers for the parity tests"""
import ctypes as C

import numpy as np

import oracle_py as O


def oracle_mesh(sd):
    return O.Mesh(sd.pos, sd.idx, sd.nrm, sd.nidx, sd.mat, sd.albedo)


def oracle_camera(sd, **kw):
    c = sd.camera
    return O.camera(c["look_from"], c["look_at"], c["up"], fov_y=c["fov_y"], **kw)


def random_rays(sd, n, seed, inside=True):
    rng = np.random.default_rng(seed)
    lo, hi = sd.pos.min(0), sd.pos.max(0)
    if not inside:
        c, e = 0.5 * (lo + hi), 0.5 * (hi - lo)
        lo, hi = c - 1.5 * e, c + 1.5 * e
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o, d


def hosttest_build(H, sd):
    st = (C.c_double * 8)()
    rc = H.wpt_hosttest_build(sd.pos.ctypes.data_as(C.c_void_p), C.c_uint32(sd.pos.shape[0]),
                              sd.idx.ctypes.data_as(C.c_void_p), C.c_uint32(sd.n_tris), st)
    return rc, dict(n_nodes=st[0], n_tris=st[1], max_depth=st[2], sah=st[3], build_ms=st[4], problems=st[5], children=st[6], leaves=st[7])


def hosttest_trace(H, sd, o, d, tmin=1e-3, tmax=1e20):
    n = o.shape[0]
    cols = [np.ascontiguousarray(o[:, k]) for k in range(3)] + [np.ascontiguousarray(d[:, k]) for k in range(3)]
    hid = np.full(n, -2, np.int32)
    t = np.full(n, np.nan, np.float32)
    u = t.copy()
    v = t.copy()
    rc = H.wpt_hosttest_trace(sd.pos.ctypes.data_as(C.c_void_p), C.c_uint32(sd.pos.shape[0]),
                              sd.idx.ctypes.data_as(C.c_void_p), C.c_uint32(sd.n_tris),
                              *[c.ctypes.data_as(C.c_void_p) for c in cols], C.c_float(tmin), C.c_float(tmax),
                              C.c_uint32(n), hid.ctypes.data_as(C.c_void_p), t.ctypes.data_as(C.c_void_p),
                              u.ctypes.data_as(C.c_void_p), v.ctypes.data_as(C.c_void_p))
    assert rc == 0
    return hid, t, u, v


def compare_hits(ref, got, what="", robust=None, tie_budget=1e-4, nontie_budget=1e-6):
    """north_star bar, as TWO assertions: (1) primitive ids equal except at documented t-ties — both hit, t within
    1e-5 relative, different primitive (shared edges / coincident triangles) — for at most `tie_budget` of the rays;
    (2) every OTHER disagreement (hit<->miss flips, different primitive at a different distance) is an edge-grazing
    ray on which the reference's own two intersectors disagree as well — `robust` = the same rays through Embree's
    ROBUST (Pluecker) intersector (oracle_py "embree_robust") siding with `got` — or it counts against
    `nontie_budget` (1e-6 of the rays; at least one ray is allowed on small sets).  t within 1e-5 relative and u, v
    within 2e-4 wherever the ids agree.  Returns the id agreement rate."""
    rid, rt, ru, rv = ref
    gid, gt, gu, gv = got
    n = rid.size
    same = rid == gid
    rate = same.mean()
    hit = same & (rid >= 0)
    if hit.any():
        rel = np.abs(rt[hit] - gt[hit]) / np.maximum(np.abs(rt[hit]), 1e-30)
        assert rel.max() <= 1e-5, f"{what}: t mismatch {rel.max()}"
        assert np.abs(ru[hit] - gu[hit]).max() <= 2e-4 and np.abs(rv[hit] - gv[hit]).max() <= 2e-4, f"{what}: uv mismatch"
    bad = ~same
    if bad.any():
        both = bad & (rid >= 0) & (gid >= 0)
        tie = np.zeros_like(bad)
        tie[both] = np.abs(rt[both] - gt[both]) <= 1e-5 * np.abs(rt[both])
        other = bad & ~tie
        grazing = np.zeros_like(bad)
        if robust is not None:
            grazing = other & (robust[0] == gid)       # the reference's robust intersector returns OUR answer
        unexplained = other & ~grazing
        assert tie.sum() + grazing.sum() <= tie_budget * n, \
            f"{what}: {tie.sum()} t-ties + {grazing.sum()} grazing rays of {n} exceed the {tie_budget:g} budget"
        assert unexplained.sum() <= max(1.0, nontie_budget * n), \
            f"{what}: {unexplained.sum()} of {n} rays disagree without being a t-tie or an edge-grazing ray " \
            f"(first: ray {np.nonzero(unexplained)[0][0]}, ref id {rid[unexplained][0]} t {rt[unexplained][0]}, " \
            f"got id {gid[unexplained][0]} t {gt[unexplained][0]})"
    return rate


def embree_pair(om, o, d, **kw):
    """(Embree fast Moeller-Trumbore hits, Embree ROBUST hits) of the same rays, or (oracle BVH hits, None) where
    oracle/_ref is absent"""
    if O.ref_available():
        return om.intersect("embree", o, d, **kw), om.intersect("embree_robust", o, d, **kw)
    return om.intersect("bvh", o, d, **kw), None


def three_sigma_outliers(a, b, spp):
    """a, b: per-pixel means of [0,1] contributions at equal spp (any shape).  Returns (fraction of values whose
    difference exceeds 3 sigma of the difference of two independent means, mean squared z-score).  sigma^2 =
    2 p (1 - p) / spp with the pooled mean p: exact for 0/1 contributions (albedo 1: a path adds 1 or nothing), an
    upper bound for any contribution in [0, 1] (Bhatia-Davis), so the outlier fraction of a correct renderer is at
    most the Gaussian 0.27 % and the mean z^2 at most 1."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    p = 0.5 * (a + b)
    var = np.clip(p * (1.0 - p), 0.0, None) / spp
    sigma = np.sqrt(2.0 * var)
    diff = np.abs(a - b)
    out = diff > 3.0 * sigma + 1e-6
    live = sigma > 0
    z2 = float(np.mean((diff[live] / sigma[live]) ** 2)) if live.any() else 0.0
    return float(out.mean()), z2

"""-m gpu: wpt_scene_refit — vertices move, topology stays.  After a refit the closest
hits must be those of the DEFORMED mesh: checked against the oracle (and Embree where oracle/_ref is
present) built on the deformed vertices, and against a scene committed from scratch on them."""
import numpy as np
import pytest

import oracle_py as O
import scenes
import util

pytestmark = pytest.mark.gpu


def _deform(pos, k):
    p = pos.astype(np.float64)
    q = p.copy()
    q[:, 0] = p[:, 0] * (1.0 + 0.15 * k) + 0.2 * k * np.sin(1.3 * p[:, 1] + k)
    q[:, 1] = p[:, 1] + 0.25 * k * np.sin(0.9 * p[:, 0] + 0.7 * p[:, 2])
    q[:, 2] = p[:, 2] * (1.0 - 0.1 * k) + 0.1 * k * np.cos(1.1 * p[:, 0])
    return q.astype(np.float32)


def _rays(sd, pos, n, seed):
    rng = np.random.default_rng(seed)
    lo, hi = pos.min(0), pos.max(0)
    o = rng.uniform(lo - 0.2 * (hi - lo), hi + 0.2 * (hi - lo), (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o, d


@pytest.mark.parametrize("maker", [lambda: scenes.cornell(24), lambda: scenes.knob(120, 60)])
def test_refit_hits_match_the_deformed_mesh(maker):
    import torch
    import wpt
    sd = maker()
    scene = wpt.scene_from(sd)
    backend = "embree" if O.ref_available() else "bvh"
    for k in (1, 2):                                            # refit twice: boxes must not accumulate error
        pos = _deform(sd.pos, k)
        if k == 1:
            scene.refit(pos)                                    # host path
        else:
            scene.refit(torch.from_numpy(pos.reshape(-1)).cuda())   # device path
            torch.cuda.synchronize()
        o, d = _rays(sd, pos, 200_000, k)
        got = scene.intersect_host(o, d)
        om = O.Mesh(pos, sd.idx, sd.nrm, sd.nidx, sd.mat, sd.albedo)
        ref = om.intersect(backend, o, d)
        assert util.compare_hits(ref, got, f"refit k={k}") >= 0.9999
        assert (ref[0] >= 0).mean() > 0.05
        # ... and equal to a from-scratch commit of the deformed mesh (ids may differ at t-ties only)
        fresh = wpt.Scene()
        fresh.add_triangle_mesh(pos, sd.idx, sd.nrm, sd.nidx, sd.mat, sd.albedo)
        fresh.commit(0)
        assert util.compare_hits(fresh.intersect_host(o, d), got, f"refit vs rebuild k={k}") >= 0.9999


def test_refit_back_to_rest_pose_restores_the_image():
    """geometry-normal faces (no vertex normals) follow the deformation; refitting back gives the original film"""
    import wpt
    sd = scenes.cornell(12, colored=True)
    scene = wpt.Scene()
    scene.add_triangle_mesh(sd.pos, sd.idx, None, None, sd.mat, sd.albedo)      # geometric normals
    scene.commit(0)
    cam = wpt.camera_from(sd)
    r = wpt.Renderer(scene, 64, 64, 2)
    film0, st0 = r.render_host(cam, 0, 4, 4, seed=5)
    scene.refit(_deform(sd.pos, 1))
    film1, st1 = r.render_host(cam, 0, 4, 4, seed=5)
    om = O.Mesh(_deform(sd.pos, 1), sd.idx, None, np.full_like(sd.idx, -1), sd.mat, sd.albedo)
    c = sd.camera
    ofilm, oseg = om.render("bvh", O.camera(c["look_from"], c["look_at"], c["up"], fov_y=c["fov_y"]), 64, 64, 0, 4, 4,
                            rng_mode=O.RNG_PHILOX, seed=5)
    assert (np.abs(film1 - ofilm / 4).max(0) > 1e-5).mean() < 0.02
    assert (np.abs(film1 - film0).max(0) > 1e-3).mean() > 0.05
    scene.refit(sd.pos)
    film2, st2 = r.render_host(cam, 0, 4, 4, seed=5)
    assert np.array_equal(film0, film2) and st0["segments"] == st2["segments"]


def test_refit_errors():
    import wpt
    sd = scenes.cornell(8)
    s = wpt.Scene()
    s.add_triangle_mesh(sd.pos, sd.idx, sd.nrm, sd.nidx)
    with pytest.raises(wpt.WptError):
        s.refit(sd.pos)                                         # not committed
    s.commit(0)
    with pytest.raises(wpt.WptError):
        s.refit(sd.pos[:-1])                                    # vertex count differs
    bad = sd.pos.copy(); bad[0, 0] = np.nan
    with pytest.raises(wpt.WptError):
        s.refit(bad)


def test_device_refit_flags_non_finite_vertices():
    """wpt_scene_refit_device cannot validate device-side positions on the host; the refit kernels raise a
    flag word instead (wpt_scene_refit_status), and the next commit (which reads the positions back) refuses them."""
    import torch
    import wpt
    sd = scenes.cornell(8)
    s = wpt.scene_from(sd)
    good = torch.from_numpy(sd.pos.reshape(-1).copy()).cuda()
    s.refit(good)
    assert s.refit_status() is False
    for bad_value in (float("nan"), float("inf")):
        bad = good.clone()
        bad[3 * 17 + 1] = bad_value
        s.refit(bad)
        assert s.refit_status() is True
        with pytest.raises(wpt.WptError):
            s.commit(0)                               # reads the device-side positions back: non-finite -> refused
        s.refit(good)
        assert s.refit_status() is False

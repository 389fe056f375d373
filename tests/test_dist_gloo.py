"""world_size-2 gloo test of the N>1 path's host logic: disjoint spp ranges keyed on the GLOBAL sample
index + one reduce of the partial films == the single-rank film.  The per-rank renderer here
is the CPU oracle in Philox mode (the device renderer is exercised by the -m gpu tests); the sharding and
reduce code (wpt_dist) is the product's."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import wpt_dist


def test_spp_ranges():
    for world in (1, 2, 3, 8):
        for total in (0, 1, 7, 64, 256):
            got = [wpt_dist.spp_range_strong(r, world, total) for r in range(world)]
            assert sum(c for _, c in got) == total
            pos = 0
            for b, c in got:
                assert b == pos
                pos += c
            assert max(c for _, c in got) - min(c for _, c in got) <= 1
    assert wpt_dist.spp_range_weak(3, 8, 64) == (192, 64)
    with pytest.raises(ValueError):
        wpt_dist.spp_range_strong(2, 2, 8)


def _worker(rank, world, port, q):
    here = os.path.dirname(os.path.abspath(__file__))
    root = os.path.dirname(here)
    for p in (os.path.join(root, "oracle"), os.path.join(root, "smallpt-enoki-optix_b200"), here):
        sys.path.insert(0, p)
    import oracle_py as O
    import scenes
    import util
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sd = scenes.cornell(8, colored=True)
    om, cam = util.oracle_mesh(sd), util.oracle_camera(sd)
    W = H = 24
    b, c = wpt_dist.spp_range_strong(rank, world, 9)
    film, seg = om.render("bvh", cam, W, H, b, c, 4, rng_mode=O.RNG_PHILOX, seed=11, threads=1)
    t = torch.from_numpy(film)
    n = wpt_dist.reduce_film(t, dst=0)
    segs = torch.tensor([seg], dtype=torch.int64)
    dist.reduce(segs, dst=0)
    if rank == 0:
        whole, wseg = om.render("bvh", cam, W, H, 0, 9, 4, rng_mode=O.RNG_PHILOX, seed=11, threads=1)
        q.put((n, bool(np.allclose(t.numpy(), whole, rtol=0, atol=1e-5)), int(segs[0]) == wseg))
    dist.destroy_process_group()


def test_two_rank_reduce_equals_single_rank():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == (2, True, True)

"""-m gpu, two GPUs (skipped on a one-GPU box; run under `gpurun --gpus 2`): the N>1 data plane of the product —
one process per GPU, samples sharded by the global index, the depth pre-pass sharded over the ranks with one
ncclAllGather inside wpt_render, ONE wpt_film_reduce (ncclReduce on libwpt's own communicator) — equals the one-GPU frame.
torch.distributed (gloo) only ships the 128-byte NCCL id."""
import os
import sys

import pytest

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, q):
    here = os.path.dirname(os.path.abspath(__file__))
    root = os.path.dirname(here)
    for p in (os.path.join(root, "oracle"), os.path.join(root, "smallpt-enoki-optix_b200"), here):
        sys.path.insert(0, p)
    import torch
    import torch.distributed as dist
    import scenes
    import wpt
    import wpt_dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    red = wpt_dist.FilmReducer(rank)
    sd = scenes.knob(120, 50)
    scene = wpt.scene_from(sd, device=rank)
    cam = wpt.camera_from(sd)
    W, H, depth = 160, 128, 6
    out = []
    # (sharded pre-pass?, samples): 15 samples split 8 / 7 — the ranks disagree on the ">= 8 samples" rule of the
    # unsharded pre-pass, the sharded one must still be entered by both (it is a collective)
    # with three or more ranks also a 3 x 2 image: ceil(6 / 4) = 2 pixels per rank leaves the last rank of four without
    # a single pixel of the pre-pass — it must launch nothing and still join the all-gather
    # (at least 8 samples per rank in the first two, or the unsharded pre-pass would not run and the retrace counts
    # could not be compared; 8 * world - 1 samples split 8, ..., 8, 7)
    cases = [(0, 12 * world, W, H), (1, 12 * world, W, H), (1, 8 * world - 1, W, H)] + ([(1, 16, 3, 2)] if world >= 3 else [])
    for shard, spp, W, H in cases:
        r = wpt.Renderer(scene, W, H, 8)
        if shard:
            r.set_depth_shard(red.comm, rank, world)
        b, c = wpt_dist.spp_range_strong(rank, world, spp)
        film = torch.zeros(3, W * H, device="cuda")
        r.render(cam, film, b, c, depth, seed=9)
        red.reduce(film, 0)
        st = r.finish()
        segs = torch.tensor([st["segments"], st["retraced"]], dtype=torch.int64)
        dist.reduce(segs, dst=0)
        out.append((film, segs, spp, W, H))
        r.close()
    # the exchange without a root: reduce-scatter, every rank scales its slice and copies it into ONE host film that all
    # ranks map (wpt_film_reduce_scatter / wpt_film_shard_finish / wpt_host_register); 13 x 7 pixels: 273 floats do not
    # divide by 2 or 4, so the padded tail is exercised
    import tempfile
    import numpy as np
    Ws, Hs, spp_s = 13, 7, 8 * world
    box = [tempfile.mkdtemp(prefix="wpt_test_", dir="/dev/shm") if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    shared = wpt_dist.SharedHostFilm(box[0], 3 * Ws * Hs, rank, dist.barrier)
    pad, lo, cnt = red.shard_layout(Ws * Hs)
    flat = torch.zeros(pad, device="cuda")
    rs = wpt.Renderer(scene, Ws, Hs, 8)
    b, c = wpt_dist.spp_range_strong(rank, world, spp_s)
    rs.render(cam, flat[:3 * Ws * Hs].view(3, Ws * Hs), b, c, depth, seed=9)
    red.reduce_scatter(flat, Ws * Hs)
    red.shard_finish(flat, Ws * Hs, 1.0 / spp_s, shared.array)
    rs.finish()
    torch.cuda.synchronize()
    dist.barrier()
    scatter_ok = None
    if rank == 0:
        whole = torch.zeros(3, Ws * Hs, device="cuda")
        rs.render(cam, whole, 0, spp_s, depth, seed=9)
        rs.finish()
        want = (whole / spp_s).cpu().numpy().reshape(-1)
        scatter_ok = bool(np.allclose(np.asarray(shared.array), want, rtol=0, atol=2e-6)) and float(want.sum()) > 0 \
            and pad % world == 0 and pad >= 3 * Ws * Hs
    dist.barrier()
    shared.close()
    rs.close()
    if rank == 0:
        import shutil
        shutil.rmtree(box[0], ignore_errors=True)
    if rank == 0:
        ok = []
        for f, s, spp, W, H in out:
            r = wpt.Renderer(scene, W, H, 8)
            whole = torch.zeros(3, W * H, device="cuda")
            r.render(cam, whole, 0, spp, depth, seed=9)
            st = r.finish()
            ok.append(bool(torch.allclose(f, whole, rtol=0, atol=2e-5)) and int(s[0]) == st["segments"])
        # sharded and unsharded pre-pass give the same seeds: the same rays are traced a second time
        q.put((ok, int(out[0][1][1]) == int(out[1][1][1]), float(whole.sum()) > 0 and scatter_ok))
    red.close()
    dist.destroy_process_group()


def _run(world):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs (run under gpurun --gpus %d)" % (world, world))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    return res


def test_two_rank_sharded_frame_equals_one_gpu():
    assert _run(2) == ([True, True, True], True, True)


def test_four_rank_sharded_frame_with_a_rank_that_owns_no_pixel():
    assert _run(4) == ([True, True, True, True], True, True)

#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 wavefront path tracer (BASELINE.json metric:
Mrays/s, all bounces, device-timed; workload C3 = configs[2]: procedural 1 007 232-triangle
terrain+spheres scene, 1920x1080, 64 spp, depth 8 — the configuration north_star's target is
quoted on; it fits one GPU).

One step = one whole 64-spp frame through the hot path (raygen -> depth x (trace, shade)), film
resident in HBM.  N>1 (torchrun, one rank per GPU): configs[2] literally — the SAME 64-spp frame split
over the ranks (strong scaling: rank g renders the global sample indices [g*64/N, (g+1)*64/N) of all
pixels) and the partial films summed onto rank 0 with one NCCL reduce per step (wpt_film_reduce of the C
ABI), inside the timed region.  --scaling weak gives every rank its own 64 samples instead.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C1..C5]

--impl reference times the reference's own CPU implementation of the path on the host cores:
the oracle's restated loop driving the real Embree 3.6.1 (oracle/_ref) where that was built,
else the oracle port; each step is a bounded sample of the same workload.
"""
import argparse
import ctypes as C
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "smallpt-enoki-optix_b200")
for p in (PKG, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

SEED = 0x5EED
# one string for both arms: the driver divides the two lines only if metric, unit and direction agree
METRIC = "Mrays/s (all bounces, device-timed)"


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


def workload(name, scale=1.0):
    import scenes
    gen, W, H, spp, depth = scenes.CONFIGS[name]
    return gen, W, H, spp, depth


def alg_bytes(n_tris):
    """Algorithmic bytes per ray segment (SURVEY §8d's formula with this layout):
    one root->leaf descent of D = ceil(log7(T/3)) 64-byte 7-wide nodes + one 3-triangle leaf of 64-byte
    blocks (the primitive id travels in the triangle block), plus the SoA record streams each kernel
    must move."""
    D = int(np.ceil(np.log(max(n_tris / 3.0, 7.0)) / np.log(7.0)))
    trav = D * 64 + 3 * 64
    trace = 32 + 16 + trav                # read (o,d) 2x16 B, write hit 16 B
    shade = 40 + 16 + 48 + 40             # read ray 40 + hit 16 + attributes 48, write next ray 40
    return dict(D=D, trav=trav, trace=trace, shade=shade, pipeline=trace + shade)


def traffic_per_launch(rays_per_launch):
    """dram bytes per launch of the dominant kernel, DERIVED from the newest committed ncu --set full capture
    (profiles/r*_trace_traffic.json: measured dram__bytes per ray x the rays one launch processes here).
    Returns (bytes, source file) — the line says where the number comes from."""
    import glob
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_trace_traffic.json")), reverse=True):
        try:
            t = json.load(open(f))
            return float(t["dram_bytes_per_ray_weighted"]) * rays_per_launch, os.path.relpath(f, ROOT)
        except Exception:
            continue
    return None, None


def ncu_counters():
    """the binding-limit counters of the dominant kernel from the newest committed ncu summary (a profile run is never
    a bench value; these are attached so that the line shows WHICH bound binds, SURVEY §8d's honesty note)"""
    import glob
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_trace_counters.json")), reverse=True):
        try:
            return json.load(open(f))
        except Exception:
            continue
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------------
def cpu_reference_run(sd, W, H, depth, spp, budget_s, threads=0):
    """The oracle's restated loop (reference RNG: per-pixel PCG32) on Embree 3.6.1 when oracle/_ref
    exists ("reference"), else on the oracle's own BVH ("port").  Returns (Mrays/s, segments, s, kind, cores)."""
    import oracle_py as O
    kind = "reference" if O.ref_available() else "port"
    backend = "embree" if kind == "reference" else "bvh"
    m = O.Mesh(sd.pos, sd.idx, sd.nrm, sd.nidx, sd.mat, sd.albedo)
    c = sd.camera
    cam = O.camera(c["look_from"], c["look_at"], c["up"], fov_y=c["fov_y"])
    m.backend(backend)  # build the CPU BVH outside the timed region (the reference does not time commit either)
    if threads <= 0:
        # all host cores this process may use; torchrun exports OMP_NUM_THREADS=1, which must not
        # shrink the CPU arm to one thread
        try:
            threads = len(os.sched_getaffinity(0))
        except AttributeError:
            threads = os.cpu_count() or 1
    cores = threads
    t0 = time.perf_counter()
    film, seg = m.render(backend, cam, W, H, 0, spp, depth, rng_mode=O.RNG_PCG32, threads=threads)
    dt = time.perf_counter() - t0
    return seg / dt / 1e6, seg, dt, kind, cores


def frame_config(args, sd, W, H, spp_total, depth):
    """what is rendered — the same dict, key for key and value for value, in both arms (the driver compares them); how an
    arm renders it is that arm's `engine` object"""
    return {"workload": f"{args.workload}:{sd.name}", "tris": sd.n_tris, "width": W, "height": H, "spp": spp_total,
            "depth": depth, "camera": "pinhole, the workload's default view", "albedo": "all materials white" if bool((sd.albedo == 1).all()) else "per-material greys",
            "l2": "GPU arm: no flush, the inputs of a step exceed the 126 MB L2 — ray/hit queues of a default pass %.0f MB "
                  "(96 B per path in flight) + the BVH; CPU arm: the same paths, host caches" % (
                      W * H * max(1, min(spp_total, 64, (1 << 28) // (W * H))) * 96 / 1e6)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import scenes
    gen, W, H, spp, depth = workload(args.workload)
    if args.spp:
        spp = args.spp
    sd = gen()
    # one step = the whole frame, all `spp` samples, like a step of the GPU arm (C3: 340 M segments, ~6 s on 16 cores,
    # ~3 s on 32: 25 steps stay within a few minutes).  --ref-spp bounds it for slower hosts.
    sample_spp = min(args.ref_spp, spp) if args.ref_spp > 0 else spp
    times, segs = [], 0
    kind, cores = "port", 1
    for i in range(args.warmup + args.steps):
        mr, seg, dt, kind, cores = cpu_reference_run(sd, W, H, depth, sample_spp, 0)
        if i >= args.warmup:
            times.append(dt); segs += seg
    tot = sum(times)
    val = segs / tot / 1e6
    sample = f"{args.workload} scene ({sd.n_tris} tris) {W}x{H}, depth {depth}, {sample_spp} spp per step (of {spp})"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "Mrays/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot / len(times),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": frame_config(args, sd, W, H, spp, depth),
            "engine": {"spp_per_step": sample_spp, "sample": sample, "threads": cores,
                       "timing": "host steady clock around the render (a CPU arm has no device clock); BVH build outside"},
            "ms_per_frame": 1e3 * tot / len(times) * spp / sample_spp, "segments_per_step": segs / len(times),
            "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample,
                             "mrays_per_core": val / cores,
                             "note": "the GPU/CPU ratio is a property of this box's host CPU (16 cores on one pool box, "
                                     "32 on another): compare Mrays/s per core across boxes"},
            "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# --------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    import scenes
    import wpt

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — libwpt has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    import wpt_dist as wdist
    reducer = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        # the data plane's own communicator inside libwpt.so; torch.distributed ships its 128-byte id and
        # does the barrier / max-over-ranks bookkeeping only
        reducer = wdist.FilmReducer(local)

    gen, W, H, spp, depth = workload(args.workload)
    if args.spp:
        spp = args.spp
    if args.scaling == "strong":
        # one frame of `spp` samples split over the ranks; spp below is this rank's share
        spp_total = spp
        spp_begin, spp = wdist.spp_range_strong(rank, world, spp_total)
        if spp == 0:
            raise SystemExit("bench.py: --scaling strong needs at least one sample per rank")
    else:
        spp_total = spp * world
        spp_begin, _ = wdist.spp_range_weak(rank, world, spp)   # disjoint, decorrelated (RNG keyed on global index)
    sd = gen()
    N = W * H
    t0 = time.perf_counter()
    # N > 1: ONE host SAH build for the box — rank 0 commits first and leaves the tree in /dev/shm, the others load it
    # (wpt_scene_commit_cached); untimed, like the reference's own build (src/main.cpp:366 precedes :373)
    cache = None
    if world > 1:
        # a directory of this run's own (0700, random name), announced by rank 0: nothing a crashed or foreign run left
        # in /dev/shm can be picked up, and the library checks what it loads anyway (csrc/api.cu tree_load)
        box = [tempfile.mkdtemp(prefix="wpt_tree_", dir="/dev/shm") if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        cache = os.path.join(box[0], "%s.bin" % args.workload)
        if rank != 0:
            dist.barrier()
    scene = wpt.scene_from(sd, device=local, cache_path=cache)
    if world > 1 and rank == 0:
        dist.barrier()
    commit_s = time.perf_counter() - t0
    bst = scene.stats()
    cam = wpt.camera_from(sd)
    if args.spp_per_pass <= 0:
        # paths in flight per pass: as many samples as fit 2^28 paths (96 B of queue each, 25 GB), at most
        # 64 — measured on C3 (profiles/): 4 -> 3349, 8 -> 3564, 16 -> 3679, 32 -> 3746, 64 -> 3803 Mrays/s;
        # late bounces hold few paths and a persistent-kernel launch has a fixed ramp-up/drain cost
        args.spp_per_pass = int(max(1, min(spp, 64, (1 << 28) // N)))
    r = wpt.Renderer(scene, W, H, args.spp_per_pass)
    # two lanes (profiles/r1_microbench.md): every trace launch at the kernel's own occupancy (8 CTAs/SM, 64 registers)
    r.set_tuning(args.variant, args.ctas_per_sm, args.overlap)
    r.set_depth_seeding(args.seeding)
    r.set_tail(args.tail)
    if reducer is not None and args.depth_shard:
        r.set_depth_shard(reducer.comm, rank, world)            # 1/N of the depth pre-pass per rank + one all-gather
    film = torch.zeros(3, N, device="cuda")

    def step():
        film.zero_()
        r.render(cam, film, spp_begin, spp, depth, seed=SEED)
        if reducer is not None:
            reducer.reduce(film, 0)                             # ONE ncclReduce of the partial films (wpt_film_reduce)
        wpt.film_scale(film, 1.0 / spp_total)
        return r.finish()

    if args.graph:
        # launch-bound small frames: the whole frame (and the reduce) recorded once into a CUDA graph, replayed per step
        step()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(graph, stream=side):
                film.zero_()
                r.render(cam, film, spp_begin, spp, depth, seed=SEED, stream=side)
                if reducer is not None:
                    reducer.reduce(film, 0, stream=side)
                wpt.film_scale(film, 1.0 / spp_total, stream=side)
        torch.cuda.current_stream().wait_stream(side)

        def step():
            graph.replay()
            return r.finish()

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    ev0.record()
    segs, launches = 0, 0
    for _ in range(args.steps):
        st = step()
        segs += st["segments"]; launches += st["kernel_launches"] + 1
    ev1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = ev0.elapsed_time(ev1)
    t = torch.tensor([ms, float(segs)], dtype=torch.float64, device="cuda")
    if world > 1:
        tm = t.clone(); dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ts = t.clone(); dist.all_reduce(ts, op=dist.ReduceOp.SUM)
        ms, segs_all = float(tm[0]), float(ts[1])
    else:
        segs_all = float(segs)
    value = segs_all / (ms * 1e-3) / 1e6

    if args.quick:
        if rank == 0:
            print(json.dumps({"quick": True, "value": value, "unit": "Mrays/s", "ms_per_step": ms / args.steps,
                              "n_gpus": world, "gpu_launches": launches, "clocks": clocks,
                              "variant": args.variant, "tail": args.tail, "overlap": args.overlap}), flush=True)
        if world > 1:
            reducer.close()
            dist.destroy_process_group()
        return 0

    # ---- e2e: the same frame through the host-buffer entry points, host clock around everything: film zero ->
    # render -> (N>1: reduce onto rank 0) -> 1/spp -> D2H of the planar film into PINNED host memory
    pinned = wpt.PinnedArray((3, N))
    host_film = pinned.array
    e_steps = max(1, args.steps)           # the same number of frames as the device-timed loop

    shared = None
    if world > 1 and args.e2e_scatter:
        # N > 1: no root.  ncclReduceScatter leaves 1/N of the summed film on every rank; each scales its slice and DMAs
        # it to its place in ONE host film that all ranks map (a /dev/shm file, page-locked in every process)
        pad, _, _ = reducer.shard_layout(N)
        film_rs = torch.zeros(pad, device="cuda")
        film_rs_planes = film_rs[:3 * N].view(3, N)
        shared = wdist.SharedHostFilm(os.path.dirname(cache), 3 * N, rank, dist.barrier)

    def e2e_step():
        if world == 1:
            _, st = r.render_host(cam, spp_begin, spp, depth, seed=SEED, out=host_film)     # wpt_render_host
            return st
        if shared is None:
            film.zero_()
            r.render(cam, film, spp_begin, spp, depth, seed=SEED)
            reducer.reduce(film, 0)
            if rank == 0:
                wpt.film_scale(film, 1.0 / spp_total)
                torch.from_numpy(host_film).copy_(film, non_blocking=True)      # pinned destination: plain DMA
            return r.finish()                                                 # synchronises the stream
        film_rs.zero_()
        r.render(cam, film_rs_planes, spp_begin, spp, depth, seed=SEED)
        reducer.reduce_scatter(film_rs, N)                                    # wpt_film_reduce_scatter
        reducer.shard_finish(film_rs, N, 1.0 / spp_total, shared.array)       # own slice: 1/spp + D2H into the shared film
        return r.finish()                                                     # synchronises the stream

    e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e_t0 = time.perf_counter()
    e_segs = 0
    for _ in range(e_steps):
        e_segs += e2e_step()["segments"]
    torch.cuda.synchronize()
    e_dt = time.perf_counter() - e_t0
    et = torch.tensor([e_dt, float(e_segs)], dtype=torch.float64, device="cuda")
    if world > 1:
        em = et.clone(); dist.all_reduce(em, op=dist.ReduceOp.MAX)
        es = et.clone(); dist.all_reduce(es, op=dist.ReduceOp.SUM)
        e_dt, e_segs_all = float(em[0]), float(es[1])
    else:
        e_segs_all = float(e_segs)
    e2e = {"value": e_segs_all / e_dt / 1e6, "unit": "Mrays/s",
           "h2d_bytes_per_step": C.sizeof(wpt.Camera) + C.sizeof(wpt.RenderConfig),
           "d2h_bytes_per_step": 3 * N * 4, "ms_per_frame": 1e3 * e_dt / e_steps, "steps": e_steps,
           "what": ("wpt_render_host: host camera/config in, film zero + full frame + 1/spp + D2H of the planar film "
                    "into pinned host memory" if world == 1 else
                    "per rank: film zero + this rank's samples; wpt_film_reduce onto rank 0; rank 0: 1/spp + D2H of the "
                    "reduced film into pinned host memory; host clock, max over ranks" if shared is None else
                    "per rank: film zero + this rank's samples; wpt_film_reduce_scatter (no root); every rank: 1/spp on its "
                    "1/N of the film + D2H of that slice into ONE host film shared by the ranks (/dev/shm mapping, page-locked "
                    "in every process); d2h bytes are the sum over the ranks; host clock, max over ranks")}

    # ---- sameness across N: the UN-normalised frame, reduced onto rank 0, summed per channel on the host in fp64.
    # Every contribution of an all-white scene is exactly 1.0 and a pixel holds at most spp of them, so the film is
    # integers and its sums are exact: they — and the segment count — must be the same numbers at N = 1, 2, 4, 8
    # (src/main.cpp:407,429 is the sum that is sharded).  One untimed frame.
    film.zero_()
    r.render(cam, film, spp_begin, spp, depth, seed=SEED)
    if reducer is not None:
        reducer.reduce(film, 0)
    cst = r.finish()
    cseg = torch.tensor([cst["segments"]], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(cseg, op=dist.ReduceOp.SUM)
    checksum = None
    if rank == 0:
        hf = film.cpu().numpy().astype(np.float64)
        sums = hf.sum(axis=1)
        checksum = {"film_checksum": [int(round(x)) for x in sums],
                    "exact": bool(np.all(hf == np.rint(hf))),
                    "escaped_paths": int(round(sums[0])) if np.all(hf[0] == np.rint(hf[0])) else None,
                    "segments": int(cseg.item()),
                    "what": "per-channel sums of the un-normalised reduced film (fp64 on the host) of one untimed frame; "
                            "`exact`: every pixel an integer (all-white scene), so the sums are too"}
        if shared is not None:
            # the host film the end-to-end leg assembled from the ranks' slices against the film reduced onto rank 0
            eh = shared.array.reshape(3, N).astype(np.float64) * spp_total
            checksum["e2e_film_checksum"] = [int(round(x)) for x in eh.sum(axis=1)]
            checksum["e2e_film_equals_reduced"] = bool(np.allclose(eh, hf, rtol=1e-6, atol=1e-6 * spp_total))

    # ---- roofline of the dominant kernel (trace), measured live with CUDA events per kernel
    roof = None
    if rank == 0:
        # per-kernel events need serial kernels: single lane, the kernel's own occupancy (8 CTAs/SM)
        r.set_tuning(args.variant, args.ctas_per_sm, 0)
        r.set_depth_shard(None, 0, 1)          # this frame runs on rank 0 alone: no collective inside wpt_render
        r.set_profile(True)
        film.zero_()
        r.render(cam, film, spp_begin, spp, depth, seed=SEED)
        pst = r.finish()
        r.set_profile(False)
        ab = alg_bytes(sd.n_tris)
        peak, which = measured_peak_gbs()
        wave = pst["tail_from"] if pst["tail_from"] >= 1 else depth
        n_trace = pst["passes"] * (wave + (1 if wave < depth else 0))
        if args.seeding and spp >= 8 and cam.c.lens_radius == 0.0:
            n_trace += 1 + pst["passes"]      # depth pre-pass + one second-trace launch per pass (wpt_render)
        trace_s = pst["ms_trace"] * 1e-3
        achieved = pst["segments"] * ab["trace"] / trace_s / 1e9
        traffic, traffic_src = traffic_per_launch(pst["segments"] / n_trace)
        shade_s = pst["ms_shade"] * 1e-3
        shaded = sum(pst["live"][:wave])           # segments that went through shade_kernel (the fused tail shades its own)
        # algorithmic bytes of those launches: every segment reads its ray record and hit (40 + 16 B); only the paths that go
        # on (the rays entering the next bounce) gather 48 B of attributes and write a 40 B record
        cont = sum(pst["live"][j + 1] for j in range(wave) if j + 1 < depth)
        shade_bytes = shaded * 56 + cont * 88
        roof = {"bound": "hbm", "kernel": "trace_wq", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "peak_source": which, "traffic": traffic, "traffic_source": traffic_src,
                "traffic_note": "derived: dram__bytes per ray of the committed ncu --set full capture x this run's rays per launch",
                "algorithmic_yardstick": True,
                "binding_limit": "instruction issue + L1 data-stage wavefronts (ncu, profiles/); the BVH is L2 resident, so "
                                 "`achieved` counts ALGORITHMIC bytes (SURVEY §8d), not DRAM traffic",
                "ncu_counters": ncu_counters(),
                "alg_bytes_per_ray": ab["trace"], "rays_per_launch": pst["segments"] / n_trace,
                "avg_launch_ms": pst["ms_trace"] / n_trace, "launches": n_trace,
                "trace_mrays_s": pst["segments"] / trace_s / 1e6,
                "trace_time_includes": "depth pre-pass, second traces and (fused tail) the in-kernel shading of bounces >= %d" % wave,
                "share_of_step": {"raygen": pst["ms_raygen"] / pst["ms_total"], "trace": pst["ms_trace"] / pst["ms_total"],
                                  "shade": pst["ms_shade"] / pst["ms_total"]},
                "shade": {"kernel": "shade_kernel", "bound": "hbm", "alg_bytes": "56 B per segment + 88 B per continuing path",
                          "achieved": shade_bytes / max(shade_s, 1e-9) / 1e9,
                          "frac": shade_bytes / max(shade_s, 1e-9) / 1e9 / peak, "segments": shaded, "continuing": cont,
                          "ms": pst["ms_shade"], "ncu": "profiles/r2_stream_kernels_ncu.txt"},
                "pipeline": {"alg_bytes_per_ray": ab["pipeline"],
                             "achieved": value / world * 1e6 * ab["pipeline"] / 1e9,
                             "frac": value / world * 1e6 * ab["pipeline"] / 1e9 / peak},
                "live_per_bounce": pst["live"][:depth]}

    # ---- CPU baseline beside it (rank 0, N=1 only; bounded sample)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        mr, seg, dt, kind, cores = cpu_reference_run(sd, W, H, depth, 1, 0)      # calibrate
        k = int(max(1, min(spp, round(12.0 / max(dt, 1e-3)))))                       # ~12 s of CPU work
        mr, seg, dt, kind, cores = cpu_reference_run(sd, W, H, depth, k, 0)
        cpu = {"value": mr, "unit": "Mrays/s", "cores": cores, "kind": kind, "mrays_per_core": mr / cores,
               "sample": f"same scene/camera, {W}x{H}, depth {depth}, {k} spp of {spp} ({seg} segments, {dt:.1f} s)"}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": frame_config(args, sd, W, H, spp_total, depth),
                "engine": {"spp_this_rank": spp, "spp_per_pass": args.spp_per_pass,
                           "parallelism": f"spp-shard x{world}", "reduce": "wpt_film_reduce (ncclReduce, own communicator)" if world > 1 else None,
                           "overlap_lanes": int(args.overlap), "cuda_graph": bool(args.graph), "depth_seeding": int(args.seeding),
                           "fused_tail_from": st["tail_from"], "trace_variant": args.variant,
                           "depth_prepass_sharded": bool(reducer is not None and args.depth_shard),
                           "l2": "inputs larger than L2: ray/hit queues %.0f MB per pass + %.0f MB BVH" % (
                               N * args.spp_per_pass * 96 / 1e6, (bst["node_bytes"] + bst["tri_bytes"]) / 1e6),
                           "bvh": {"nodes": bst["n_nodes"], "depth": bst["max_depth"], "sah": bst["sah_cost"],
                                   "build_ms": bst["build_ms"], "upload_ms": bst["upload_ms"]}},
                "film_checksum": checksum["film_checksum"], "checksum": checksum,
                "ms_per_frame": ms / args.steps, "segments_per_step": segs_all / args.steps,
                "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu,
                "commit_s": commit_s, "tree_shared_through_cache": bool(cache)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        if shared is not None:
            shared.close()
        if rank == 0 and cache:
            shutil.rmtree(os.path.dirname(cache), ignore_errors=True)
        reducer.close()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C3", choices=["C1", "C2", "C3", "C4", "C5"])
    ap.add_argument("--spp", type=int, default=0, help="override the workload's spp")
    ap.add_argument("--spp-per-pass", type=int, default=0, help="samples in flight per pass (0 = auto)")
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--ctas-per-sm", type=int, default=0)
    ap.add_argument("--overlap", type=int, default=1, help="1 = two lanes on two streams (shade of one overlaps trace of the other)")
    ap.add_argument("--scaling", choices=("weak", "strong"), default="strong",
                    help="strong (default): the frame's --spp samples split over the ranks; weak: every rank renders --spp samples")
    ap.add_argument("--tail", type=int, default=0, help="fused tail from this bounce on (0 = automatic, -1 = off)")
    ap.add_argument("--depth-shard", type=int, default=1, help="N>1: 1 = shard the depth pre-pass over the ranks (default)")
    ap.add_argument("--graph", action="store_true", help="replay the frame from a CUDA graph (single GPU)")
    ap.add_argument("--seeding", type=int, default=1, help="1 = depth seeding of the camera rays (default), 0 = off")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-scatter", type=int, default=0,
                    help="N>1 end-to-end leg: 1 = wpt_film_reduce_scatter + every rank copies its slice into one shared host film; "
                         "0 = wpt_film_reduce onto rank 0 + D2H there (default)")
    ap.add_argument("--ref-spp", type=int, default=0, help="--impl reference: samples per step (0 = the workload's, like the GPU arm)")
    ap.add_argument("--quick", action="store_true", help="profiling runs: timed loop only, short warm-up allowed")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours" and not args.quick:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())

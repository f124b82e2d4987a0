#!/usr/bin/env python
"""bench.py — point-frames/sec of the colorization hot path on N B200s (+ roofline, CPU baseline, end-to-end).

A "step" is one pass of the hot path over one batch: every keyframe of the workload projected, culled, sampled and
accumulated over the whole map cloud, the per-point accumulators exchanged across ranks (N>1) and finalized.
  value : whole-job point-frames/s with the cloud and all images already resident in HBM (CUDA events on the stream the
          kernels are launched on, max over ranks)
  e2e   : the same metric through the C ABI with HOST buffers: cloud upload, one image upload per frame and the
          read-back of the colours are inside the timed region
  roofline / cpu_baseline / parity : see DESIGN.md "Measurement"

Workloads (BASELINE.json configs):
  N = 1 : C2 (1e7 points x 200 frames @ 1920x1080), the configuration the metric is quoted on.
  N > 1 : C4 (1e8 points x 2000 frames @ 1920x1080) with the FRAMES SHARDED over the ranks ("scaling": "strong"), as
          BASELINE.json config 4 names it; `--scaling weak` keeps round 1's mode (C2 per rank).  In strong mode rank 0 also
          times the WHOLE frame set on its own GPU (`single_gpu_same_workload`) so that the line carries its own N=1 point.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference] [--workload C2] [--sample bilinear]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "point_frames_per_sec_colorized"
UNIT = "point-frames/s"
FRAME_BLOCK = 25     # strong scaling: frames are dealt to the ranks in blocks of 25 consecutive keyframes (pose locality + balance)


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def _host_threads() -> int:
    """All the host threads this process may use.  torch.distributed.run exports OMP_NUM_THREADS=1 to its workers; the CPU
    arm ignores it (it is the only thing running on the host when it runs) and passes this count to the oracle explicitly."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def _plan(args, world):
    """Workload of this run: (name, points, frames of the whole job, W, H, scaling, image period)."""
    from pointcloudprocessor_b200 import synth
    scaling = args.scaling or ("strong" if world > 1 else "single")
    name = args.workload or ("C4" if scaling == "strong" else "C2")
    n, f, w, h = synth.CONFIGS[name]
    if args.points:
        n = args.points
    if args.frames:
        f = args.frames
    if scaling == "weak":
        total = f * world
        period = f
    else:
        total = f
        period = args.image_period or (40 if f > 400 else f)
        period = min(period, f)
    return dict(name=name, n=n, frames_total=total, frames_per_rank_weak=f, W=w, H=h, scaling=scaling, period=period, block=args.frame_block or FRAME_BLOCK)


def _frames_of_rank(plan, rank, world):
    """Global keyframe numbers of one rank."""
    F = plan["frames_total"]
    if plan["scaling"] == "weak":
        f = plan["frames_per_rank_weak"]
        return list(range(rank * f, (rank + 1) * f))
    from pointcloudprocessor_b200.shard import frames_of_rank
    return frames_of_rank(F, rank, world, "dealt", plan.get("block", FRAME_BLOCK))


def _config(plan, world, sample):
    """The `config` object: identical in the native and the reference arm of the same launch."""
    n, F, W, H = plan["n"], plan["frames_total"], plan["W"], plan["H"]
    mode = {"single": "one GPU", "strong": f"frames sharded over {world} GPUs in blocks of {plan.get('block', FRAME_BLOCK)} (fixed total work)",
            "weak": f"{plan['frames_per_rank_weak']} frames per GPU x {world} GPUs (work grows with N)"}[plan["scaling"]]
    return {
        "workload": f"{plan['name']}: {n} points x {F} frames @ {W}x{H}, colorize + z-buffer culling, SUM accumulators, {sample} sampling",
        "points": n, "frames": F, "image": [W, H], "n_gpus": world,
        "parallelism": f"{mode}; cloud replicated; accumulators exchanged once per step (NCCL over NVLink)" if world > 1 else mode,
        "images": f"{plan['period']} distinct synthetic images, frame g reads image g mod {plan['period']}" if plan["period"] < F else "one synthetic image per frame",
        "l2": "inputs larger than L2 (cloud 16 B x N + one image per frame); no explicit flush",
    }


class ClockSampler:
    """nvidia-smi clocks during the timed region (B200_PROFILING.md "clocks DURING the timed region")."""
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.tmp = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "50",
                                          "-i", str(index)], stdout=self.tmp, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.tmp.flush()
        rows = [r.strip().split(",") for r in open(self.tmp.name).read().strip().splitlines() if r.strip()]
        os.unlink(self.tmp.name)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except Exception:
                continue
            for k, nm in enumerate(names):
                if len(r) > 2 + k and r[2 + k].strip().lower().startswith("active"):
                    reasons.add(nm)
        if sm:
            out["sm_mhz"] = statistics.median(sm)
            out["sm_max_mhz"] = max(mx)
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


# ---------------------------------------------------------------------------------------------------------------------
# CPU legs: the oracle (the reference itself cannot be compiled on this box) on all host cores.
# These two functions and the parity checks below are the only places where bench.py touches oracle/.
# ---------------------------------------------------------------------------------------------------------------------
def _oracle_accumulate(cloud, poses, images_of, frames, W, H, intr, dist, threads, bilinear):
    """Runs the oracle over the given global frames; returns (seconds, u32 accumulators)."""
    from oracle import pyoracle as O
    oc = O.Colorizer(O.camera(intr, dist), W, H, cloud, bilinear=bilinear)
    t0 = time.perf_counter()
    for g in frames:
        oc.frame(poses[g], images_of(g), threads=threads)
    dt = time.perf_counter() - t0
    acc = oc.accumulators()[0]
    oc.close()
    return dt, acc


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return 0  # the CPU arm is one process on the host's cores; the other ranks of a torchrun launch have nothing to do
    from oracle import pyoracle as O
    from pointcloudprocessor_b200 import synth
    O.build()
    plan = _plan(args, max(world, args.gpus))
    n, F, W, H = plan["n"], plan["frames_total"], plan["W"], plan["H"]
    intr, dist = synth.make_camera(W, H)
    cloud = synth.make_cloud(n)
    poses, _ = synth.make_poses(F)
    threads = _host_threads()
    bilinear = args.sample == "bilinear"
    cam = O.camera(intr, dist)
    oc = O.Colorizer(cam, W, H, cloud, bilinear=bilinear)
    img0 = synth.make_image(0, W, H)
    t0 = time.perf_counter()
    oc.frame(poses[0], img0, threads=threads)
    t1 = time.perf_counter() - t0
    total_steps = args.steps + args.warmup
    k = int(max(1, min(F, plan["period"], (150.0 / total_steps) / max(t1, 1e-6))))  # whole run <= ~2.5 min
    images = [synth.make_image(i, W, H) for i in range(k)]
    times = []
    for s in range(total_steps):
        oc.close()
        oc = O.Colorizer(cam, W, H, cloud, bilinear=bilinear)
        t0 = time.perf_counter()
        for i in range(k):
            oc.frame(poses[i], images[i], threads=threads)
        oc.finalize_sum()
        dt = time.perf_counter() - t0
        if s >= args.warmup:
            times.append(dt)
    oc.close()
    total = sum(times)
    value = n * k * len(times) / total
    sample = f"first {k} of {F} frames over the full {n}-point cloud per step, OpenMP over points on {threads} threads"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak" if plan["scaling"] == "weak" else "strong" if world > 1 else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": _config(plan, max(world, args.gpus), args.sample),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "sample_frames_per_step": k,
    }
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------------------------------
# native arm
# ---------------------------------------------------------------------------------------------------------------------
def run_native(args):
    import torch
    import torch.distributed as dist

    from pointcloudprocessor_b200 import api, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the CUDA library is the only implementation (no CPU fallback)")
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py: --gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
    torch.cuda.set_device(local)
    numa = _bind_to_gpu_numa_node(local) if world > 1 else None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    plan = _plan(args, world)
    n, F, W, H, P = plan["n"], plan["frames_total"], plan["W"], plan["H"], plan["period"]
    strong = plan["scaling"] == "strong"
    intr, dcoef = synth.make_camera(W, H)
    cam = api.create_camera("pinhole", intr, dcoef)
    sample_mode = api.SAMPLE_BILINEAR if args.sample == "bilinear" else api.SAMPLE_NEAREST
    prm = api.default_params(image_size=(W, H), sample_mode=sample_mode)
    poses_all, _ = synth.make_poses(F)
    mine = _frames_of_rank(plan, rank, world)
    f = len(mine)
    cloud_np = synth.make_cloud(n)

    # images: frame g reads image g mod P.  Rank 0 of a strong run also times the whole frame set alone, so it holds all P.
    need = sorted({g % P for g in (range(F) if strong else mine)})   # strong mode: the shares are re-dealt after the warm-up, keep every image
    slot_of = {im: k for k, im in enumerate(need)}
    img_bytes = H * W * 3
    h_cloud = torch.empty((n, 4), dtype=torch.float32).pin_memory()
    h_cloud.numpy()[...] = cloud_np
    # the e2e leg hands the map over the way the reference holds it: pcl::PointXYZI records of 32 bytes {x, y, z, 1, intensity, pad x 3}
    pcl_layout = args.cloud_layout == "pcl"
    if pcl_layout:
        h_cloud32 = torch.zeros((n, 8), dtype=torch.float32).pin_memory()
        h32 = h_cloud32.numpy()
        h32[:, :3] = cloud_np[:, :3]
        h32[:, 3] = 1.0
        h32[:, 4] = cloud_np[:, 3]
    h_imgs = torch.empty((len(need), H, W, 3), dtype=torch.uint8).pin_memory()
    for k, im in enumerate(need):
        synth.make_image(im, W, H, out=h_imgs.numpy()[k])
    h_rgb = torch.empty((n, 3), dtype=torch.uint8).pin_memory()
    h_cnt = torch.empty((n,), dtype=torch.int32).pin_memory()
    d_cloud = h_cloud.cuda()
    d_imgs = h_imgs.cuda()
    d_rgb = torch.empty((n, 3), dtype=torch.uint8, device="cuda")
    d_cnt = torch.empty((n,), dtype=torch.int32, device="cuda")

    def dimg(g):
        return d_imgs.data_ptr() + slot_of[g % P] * img_bytes

    def himg(g):
        return h_imgs.data_ptr() + slot_of[g % P] * img_bytes

    def himg_np(g):
        return h_imgs.numpy()[slot_of[g % P]]

    stream = torch.cuda.Stream()
    gc = api.Colorizer(cam, prm, device=local)
    gc.set_stream(stream.cuda_stream)
    gc.set_cloud_device(d_cloud.data_ptr(), n)
    if world > 1:
        ids = [gc.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        gc.nccl_init(ids[0], rank, world)

    ex0 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ex1 = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]

    mine_poses = np.ascontiguousarray(poses_all[mine])
    mine_ptrs = np.array([dimg(g) for g in mine], np.uint64)

    def step_device(timed_step=None):
        gc.reset_accumulators()
        gc.submit_frames_device(mine_poses, mine_ptrs, W * 3)   # the keyframe loop in one call (same work as per-frame calls)
        if world > 1:
            if timed_step is not None:
                gc.join()                      # the exchange interval starts when the last frame's kernels are done
                ex0[timed_step].record(stream)
            gc.exchange_finalize_device(d_rgb.data_ptr(), d_cnt.data_ptr())
            if timed_step is not None:
                ex1[timed_step].record(stream)
        else:
            gc.finalize_device(d_rgb.data_ptr(), d_cnt.data_ptr())

    # N > 1: the map crosses the host link once in total — every rank uploads its slice, the slices are all-gathered over NVLink
    # (pcp_set_cloud_sharded); --cloud-upload replicated: every rank uploads the whole map (round 2's first form, host-link bound)
    sharded_upload = world > 1 and args.cloud_upload == "sharded"
    slice_first, _slice_count = api.cloud_slice(n, rank, world)

    def step_e2e():
        stride, ioff = (32, 16) if pcl_layout else (16, 12)   # pcl::PointXYZI records, or float4
        base = (h_cloud32 if pcl_layout else h_cloud).data_ptr()
        if sharded_upload:
            gc.set_cloud_sharded_ptr(base + slice_first * stride, n, stride=stride, intensity_offset=ioff)
        else:
            gc.set_cloud_ptr(base, n, stride=stride, intensity_offset=ioff)   # H2D cloud (resets the accumulators)
        for g in mine:
            gc.submit_frame_ptr(poses_all[g], himg(g), W * 3)  # H2D image per frame, overlapped on the copy stream
        if world > 1:
            gc.exchange_finalize_device(d_rgb.data_ptr(), d_cnt.data_ptr())
            if rank == 0:                                 # the job's output is read back once, by rank 0 (it writes the PCD)
                with torch.cuda.stream(stream):           # the stream the exchange was enqueued on
                    h_rgb.copy_(d_rgb, non_blocking=True)
                    h_cnt.copy_(d_cnt, non_blocking=True)
            gc.synchronize()
        else:
            gc.finalize_ptr(h_rgb.data_ptr(), h_cnt.data_ptr())  # D2H colours + observation counts (synchronises)

    # ---- statistics pass (untimed): visible fraction for the algorithmic-byte count --------------------------------
    gc.reset_accumulators()
    nvis = 0
    stat_frames = mine[:16]
    stat_sum = {"chunks_fetched": 0, "active_chunks": 0, "records": 0}
    for g in stat_frames:
        r = gc.submit_frame_device(poses_all[g], dimg(g), W * 3, want_counts=True)
        nvis += r["n_visible"]
        c = gc.debug_counters()          # the device counters of that frame (untimed pass)
        for k in stat_sum:
            stat_sum[k] += c[k]
    f_vis = nvis / (max(len(stat_frames), 1) * n)

    # ---- value: HBM-resident ------------------------------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step_device()
    gc.synchronize()
    torch.cuda.synchronize()
    barrier()

    # ---- strong scaling: cost-aware dealing of the frame blocks (untimed) -----------------------------------------------
    # The blocks of a path do not cost the same (what the camera sees changes along it) and eight GPUs under load do not run
    # at the same speed: with equal block counts the ranks' frame times differ by > 10 % and the fast ranks wait in the
    # exchange.  So, after the warm-up: (1) every rank times each of ITS blocks once (events around the block, untimed step);
    # (2) the ranks all-gather the per-block costs and deal the blocks longest-first to the rank that would finish earliest
    # (shard.deal_by_cost); (3) one more untimed step measures each rank's time on its new share, which gives the ranks' relative
    # speeds, and the deal is repeated with them.  The partition is a deterministic function of the gathered numbers, identical
    # on every rank; any partition gives bit-identical sums (integer accumulators), which the parity check below verifies.
    rebalance = None
    if strong and world > 1 and not args.no_rebalance:
        from pointcloudprocessor_b200.shard import deal_by_cost
        B_ = plan["block"]
        n_blocks = (F + B_ - 1) // B_
        frames_of_block = lambda bb: list(range(bb * B_, min((bb + 1) * B_, F)))
        my_blocks = sorted({g // B_ for g in mine})
        evs = []
        gc.reset_accumulators()
        for bb in my_blocks:
            fr = frames_of_block(bb)
            ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            gc.join()
            ea.record(stream)
            gc.submit_frames_device(poses_all[fr], np.array([dimg(g) for g in fr], np.uint64), W * 3)
            gc.join()
            eb.record(stream)
            evs.append((bb, ea, eb))
        gc.exchange_finalize_device(d_rgb.data_ptr(), d_cnt.data_ptr())
        gc.synchronize()
        torch.cuda.synchronize()
        t = torch.zeros(n_blocks, device="cuda", dtype=torch.float64)
        for bb, ea, eb in evs:
            t[bb] = ea.elapsed_time(eb)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        costs = [float(x) for x in t.cpu()]
        speed = [1.0] * world
        hist = []
        for it in range(2):
            owner = deal_by_cost(costs, world, speed)
            mine = [g for bb in range(n_blocks) if owner[bb] == rank for g in frames_of_block(bb)]
            f = len(mine)
            mine_poses = np.ascontiguousarray(poses_all[mine])
            mine_ptrs = np.array([dimg(g) for g in mine], np.uint64)
            predicted = [sum(c for bb, c in enumerate(costs) if owner[bb] == r) for r in range(world)]
            ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            gc.reset_accumulators()
            ea.record(stream)
            gc.submit_frames_device(mine_poses, mine_ptrs, W * 3)
            gc.join()
            eb.record(stream)
            gc.exchange_finalize_device(d_rgb.data_ptr(), d_cnt.data_ptr())
            gc.synchronize()
            torch.cuda.synchronize()
            m = torch.zeros(world, device="cuda", dtype=torch.float64)
            m[rank] = ea.elapsed_time(eb)
            dist.all_reduce(m, op=dist.ReduceOp.SUM)
            measured = [float(x) for x in m.cpu()]
            hist.append({"blocks_by_rank": [owner.count(r) for r in range(world)], "frames_ms_by_rank": [round(x, 3) for x in measured],
                         "speed_used": [round(x, 4) for x in speed]})
            rel = [predicted[r] / max(measured[r], 1e-9) for r in range(world)]   # work per millisecond, relative
            mean_rel = sum(rel) / world
            speed = [speed[r] * 0 + rel[r] / mean_rel for r in range(world)]
        barrier()
        rebalance = {"block_cost_ms_min_mean_max": [round(min(costs), 3), round(sum(costs) / n_blocks, 3), round(max(costs), 3)], "iterations": hist,
                     "what": "blocks of %d keyframes timed once each, dealt longest-first to the earliest-finishing rank (shard.deal_by_cost), repeated once "
                             "with the ranks' measured relative speeds; untimed steps after the warm-up; results do not depend on the partition" % B_}

    sampler = ClockSampler(local) if rank == 0 else None
    # per-kernel events inside the timed region (informative: "kernel_ms_per_frame_in_timed_region") only for large clouds: they are four
    # event records per frame on the submitting thread and keep the library from putting several small frames into one launch, and for
    # clouds of <= 4e6 points the submitting thread is what bounds the frame rate
    timed_profile = n > 4_000_000
    gc.profile_enable(timed_profile)
    l0 = gc.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    ev0.record(stream)
    for s in range(args.steps):
        step_device(timed_step=s)
    ev1.record(stream)
    gc.synchronize()
    torch.cuda.synchronize()
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = gc.launch_count() - l0
    prof = gc.profile_read()
    xphases = gc.exchange_times() if world > 1 else None   # phases of this rank's last exchange of the timed region
    gc.profile_enable(False)
    clocks = sampler.stop() if sampler else None
    ms_rank = ms
    ms = max_over_ranks(ms)
    value = n * F * args.steps / (ms * 1e-3)
    exchange_ms = exchange_min = exchange_mean = None
    if world > 1:
        mine_ex = sum(a.elapsed_time(b) for a, b in zip(ex0, ex1)) / args.steps
        exchange_ms = max_over_ranks(mine_ex)
        # the rank that arrives LAST waits for nobody: the minimum over ranks is the cost of the exchange itself, the rest of the
        # maximum is time spent waiting for slower ranks (frame-cost imbalance between the shards)
        t = torch.tensor([mine_ex, -mine_ex, mine_ex], device="cuda", dtype=torch.float64)
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        exchange_min, exchange_mean = -float(tmax[1].item()), float(tsum[2].item()) / world
        frames_mine = ms_rank / args.steps - mine_ex      # this rank's frames alone (its step minus its exchange interval)
        t2 = torch.tensor([frames_mine, -frames_mine], device="cuda", dtype=torch.float64)
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        frames_max, frames_min = float(t2[0].item()), -float(t2[1].item())

    # ---- e2e: host buffers through the C ABI --------------------------------------------------------------------------
    e2e_steps = max(1, min(args.steps, 10 if n <= 20_000_000 else 3))
    step_e2e()
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    gc.synchronize()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e_value = n * F * e2e_steps / e2e_s
    h2d = (1 if sharded_upload else world) * n * (32 if pcl_layout else 16) + F * img_bytes   # the cloud once (sharded) or per rank; every frame's image once
    d2h = n * 3 + n * 4                        # colours + counts, read back once (rank 0 when N > 1)
    e2e_matches_device = None
    if rank == 0:
        rgb_e2e = h_rgb.numpy().copy()
        e2e_matches_device = bool(np.array_equal(rgb_e2e, d_rgb.cpu().numpy()))

    # ---- parity against the oracle (every N) --------------------------------------------------------------------------
    # N = 1: the frames the cpu_baseline sample timed.  N > 1: the first --parity-frames frames of EVERY rank go through the
    # sharded path and the exchange; rank 0 compares the exchanged accumulators with the oracle run over the same global frames.
    cpu = None
    parity = None
    bilinear = args.sample == "bilinear"
    threads = _host_threads()
    if not args.no_cpu:
        if world == 1:
            from oracle import pyoracle as O
            img0 = himg_np(mine[0])
            oc = O.Colorizer(O.camera(intr, dcoef), W, H, cloud_np, bilinear=bilinear)
            t0 = time.perf_counter()
            oc.frame(poses_all[mine[0]], img0, threads=threads)
            t1 = time.perf_counter() - t0
            oc.close()
            k = int(max(1, min(f, args.cpu_budget / max(t1, 1e-6))))
            frames_chk = mine[:k]
            dt, acc_o = _oracle_accumulate(cloud_np, poses_all, himg_np, frames_chk, W, H, intr, dcoef, threads, bilinear)
            cpu = {"value": n * k / dt, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": f"first {k} of {F} frames over the full {n}-point cloud, OpenMP over points, {dt:.1f} s"}
            gc.reset_accumulators()
            for g in frames_chk:
                gc.submit_frame_device(poses_all[g], dimg(g), W * 3)
            acc_g, _, _ = gc.accumulators()
            parity = {"frames": k, "accumulators_bit_exact_vs_oracle": bool(np.array_equal(acc_g, acc_o))}
        else:
            kp = max(1, args.parity_frames)
            gc.reset_accumulators()
            for g in mine[:kp]:
                gc.submit_frame_device(poses_all[g], dimg(g), W * 3)
            gc.allreduce()
            gc.synchronize()
            parts = [None] * world
            dist.all_gather_object(parts, [int(g) for g in mine[:kp]])   # the frames every rank actually put through
            if rank == 0:
                acc_g, _, _ = gc.accumulators()
                frames_chk = sorted(g for part in parts for g in part)
                imgs_chk = {}

                def img_of(g):
                    im = g % P
                    if im not in imgs_chk:
                        imgs_chk[im] = himg_np(g) if im in slot_of else synth.make_image(im, W, H)
                    return imgs_chk[im]

                dt, acc_o = _oracle_accumulate(cloud_np, poses_all, img_of, frames_chk, W, H, intr, dcoef, threads, bilinear)
                parity = {"frames": len(frames_chk), "frames_per_rank": kp, "ranks": world, "oracle_threads": threads, "oracle_seconds": round(dt, 1),
                          "what": "exchanged accumulators of the first frames of every rank vs the oracle over the same global frames",
                          "accumulators_bit_exact_vs_oracle": bool(np.array_equal(acc_g, acc_o))}
            barrier()

    # ---- strong scaling: the same frame set on ONE GPU (rank 0; the other ranks wait) ---------------------------------
    single = None
    if strong and world > 1:
        if rank == 0:
            g1 = api.Colorizer(cam, prm, device=local)
            g1.set_stream(stream.cuda_stream)
            g1.set_cloud_device(d_cloud.data_ptr(), n)
            allf = list(range(F))

            all_ptrs = np.array([dimg(g) for g in allf], np.uint64)

            def step_single():
                g1.reset_accumulators()
                g1.submit_frames_device(poses_all[allf], all_ptrs, W * 3)
                g1.finalize_device(d_rgb.data_ptr(), d_cnt.data_ptr())

            step_single()
            g1.synchronize()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ss = 2
            s0.record(stream)
            for _ in range(ss):
                step_single()
            s1.record(stream)
            g1.synchronize()
            torch.cuda.synchronize()
            sms = s0.elapsed_time(s1) / ss
            single = {"n_gpus": 1, "value": n * F / (sms * 1e-3), "ms_per_step": sms, "steps": ss,
                      "note": "the whole frame set of this run on rank 0's GPU alone (no exchange), timed after the N-GPU region"}
            g1.close()
        barrier()

    # ---- roofline of the dominant kernel ------------------------------------------------------------------------------
    # The timed region above runs frames_in_flight frames concurrently (kernels of neighbouring frames overlap), so the
    # per-kernel durations recorded there include time shared with other kernels.  The roofline numbers come from a second
    # timed pass over the same frames with frames_in_flight = 1 (one kernel at a time), CUDA events on the launching stream.
    roofline = None
    if rank == 0:
        peak, peak_src = _peaks()
        frames_prof = max(prof["frames"], 1)
        kern_ms_pipelined = {"scan": prof["scan_ms"] / frames_prof, "resolve": prof["resolve_ms"] / frames_prof} if timed_profile else None
        prm1 = api.default_params(image_size=(W, H), frames_in_flight=1, sample_mode=sample_mode)
        gc1 = api.Colorizer(cam, prm1, device=local)
        gc1.set_stream(stream.cuda_stream)
        gc1.set_cloud_device(d_cloud.data_ptr(), n)
        serial_steps = max(1, min(args.steps, 5 if n <= 20_000_000 else 1))

        def step_serial():
            gc1.reset_accumulators()
            gc1.submit_frames_device(mine_poses, mine_ptrs, W * 3)
            gc1.finalize_device(d_rgb.data_ptr(), d_cnt.data_ptr())

        for _ in range(2 if n <= 20_000_000 else 1):
            step_serial()
        gc1.synchronize()
        gc1.profile_enable(True)
        evs0, evs1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        evs0.record(stream)
        for _ in range(serial_steps):
            step_serial()
        evs1.record(stream)
        gc1.synchronize()
        torch.cuda.synchronize()
        serial_ms_per_step = evs0.elapsed_time(evs1) / serial_steps
        prof1 = gc1.profile_read()
        gc1.close()
        frames_prof1 = max(prof1["frames"], 1)
        kern_ms = {"scan": prof1["scan_ms"] / frames_prof1, "resolve": prof1["resolve_ms"] / frames_prof1}
        dominant = max(kern_ms, key=kern_ms.get)
        Wc, Hc = W // prm.downsample, H // prm.downsample
        scan_bytes = 16.0 * n                                    # pass 1 reads every point once as float4
        pipe_bytes = n * (32.0 + 32.0 * f_vis) + 3.0 * W * H + 16.0 * Wc * Hc   # whole per-frame pipeline, §8d formula
        scan_gbs = scan_bytes / (kern_ms["scan"] * 1e-3) / 1e9 if kern_ms["scan"] > 0 else 0.0
        frame_ms = sum(kern_ms.values())
        pipe_gbs = pipe_bytes / (frame_ms * 1e-3) / 1e9 if frame_ms > 0 else 0.0
        ms_step_rank = ms / args.steps
        roofline = {
            "bound": "hbm", "kernel": "classify_kernel + scan_ring_kernel (pass 1)" if dominant == "scan" else dominant, "achieved": scan_gbs, "peak": peak,
            "unit": "GB/s", "frac": scan_gbs / peak, "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": scan_bytes,
            "kernel_ms_per_frame": kern_ms, "kernel_share_of_frame": {k: v / frame_ms for k, v in kern_ms.items()} if frame_ms > 0 else None,
            "pipeline": {"algorithmic_bytes_per_frame": pipe_bytes, "f_vis": f_vis, "achieved": pipe_gbs, "frac": pipe_gbs / peak},
            "note": "achieved = ALGORITHMIC bytes (every point of the cloud, SURVEY.md 8d) / time; the pre-binned scan never reads chunks that are "
                    "certainly culled, so DRAM traffic (`traffic`, ncu) is below the algorithmic bytes and frac is not capped at 1",
            "measured_on": f"second timed pass on rank 0, {serial_steps} steps with frames_in_flight=1 (kernels not overlapped): {serial_ms_per_step:.3f} ms/step",
            "kernel_ms_per_frame_in_timed_region": kern_ms_pipelined,
            "pipeline_in_timed_region": {"achieved": pipe_bytes * f / (ms_step_rank * 1e-3) / 1e9, "frac": pipe_bytes * f / (ms_step_rank * 1e-3) / 1e9 / peak,
                                         "note": "algorithmic bytes of this rank's frames / ms_per_step of the headline (overlapped) run"},
        }
        # DRAM bytes of pass 1 from THIS run's device counters: every fetched chunk is one 1 KB bulk copy; classify_kernel reads the chunk
        # boxes and ids (36 B per chunk) and writes the 16-byte list entries the scouts read back; records are written once (32 B)
        ns = max(len(stat_frames), 1)
        n_chunks_total = (n + 63) // 64
        counted = {k: v / ns for k, v in stat_sum.items()}
        in_run = counted["chunks_fetched"] * 1024.0 + n_chunks_total * 36.0 + counted["active_chunks"] * 32.0 + counted["records"] * 32.0
        roofline["traffic_in_run"] = {"bytes_per_launch": in_run, "frac_of_peak": in_run / (kern_ms["scan"] * 1e-3) / 1e9 / peak if kern_ms["scan"] > 0 else None,
                                      "per_frame_counters": counted,
                                      "how": "chunks fetched x 1 KB + chunks x 36 B (boxes, ids) + active entries x 32 B (written, read) + records x 32 B, "
                                             "from the frame counters of this run's statistics pass"}
        tr = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tr):
            try:
                tj = json.load(open(tr))
                roofline["traffic"] = tj.get("scan_kernel_dram_bytes_per_launch")
                roofline["traffic_source"] = tj.get("source_note", "profiles/traffic.json: ncu --set full capture of this command at C2 (dram__bytes_read.sum + dram__bytes_write.sum per launch)")
                if roofline["traffic"] and kern_ms["scan"] > 0 and plan["name"] == "C2":
                    roofline["dram_frac"] = roofline["traffic"] / (kern_ms["scan"] * 1e-3) / 1e9 / peak
            except Exception:
                pass

    gc.close()
    if rank == 0:
        cfg = _config(plan, world, args.sample)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong" if strong and world > 1 else "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": cfg,
            "host": {"affinity": (f"rank bound to the {numa} CPUs NVML lists for its GPU" if numa else "unbound"), "threads": threads},
            "frames_in_flight": (int(prm.frames_in_flight) or "auto (12 slots and 4 frames per launch for clouds <= 4e6 points, else 3 slots)"), "frames_per_rank": f,
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": e2e_steps,
                    "colours_equal_device_path": e2e_matches_device,
                    "cloud_layout": "pcl::PointXYZI, 32-byte records (the reference's)" if pcl_layout else "float4 {x,y,z,intensity}",
                    "cloud_upload": ("each rank uploads 1/N of the map, ncclAllGather over NVLink (pcp_set_cloud_sharded)" if sharded_upload
                                     else ("every rank uploads the whole map" if world > 1 else "whole map"))},
            "roofline": roofline,
        }
        if exchange_ms is not None:
            line["exchange"] = {"ms_per_step": exchange_ms, "what": "reduce-scatter of the u32 sums (16 B/point), finalize of the rank's slice, all-gather of colours + counts (8 B/point); "
                                "CUDA events around it on the launching stream, max over ranks", "share_of_step": exchange_ms / (ms / args.steps),
                                "ms_min_over_ranks": exchange_min, "ms_mean_over_ranks": exchange_mean,
                                "frames_ms_per_step_min_max_over_ranks": [frames_min, frames_max],
                                "phases_ms_rank0_last_step": xphases,
                                "note": "min over ranks = the exchange itself (the last rank to arrive waits for nobody); max - min = waiting for slower ranks"}
        if single is not None:
            line["single_gpu_same_workload"] = single
        if rebalance is not None:
            line["rebalance"] = rebalance
        if cpu is not None:
            line["cpu_baseline"] = cpu
        if parity is not None:
            line["parity"] = parity
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def _bind_to_gpu_numa_node(local):
    """Multi-rank runs: pin this process to the CPUs NVML names as closest to its GPU BEFORE the pinned staging buffers are
    allocated, so that first touch puts them on the GPU's NUMA node (the end-to-end path is bound by the host link).  Best effort."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (m >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="", help="C1..C5 (default: C2 on one GPU, C4 sharded on several)")
    ap.add_argument("--scaling", default="", choices=["", "strong", "weak"], help="N>1: strong = C4 frames sharded (default), weak = C2 per rank")
    ap.add_argument("--points", type=int, default=0)
    ap.add_argument("--frames", type=int, default=0)
    ap.add_argument("--image-period", type=int, default=0)
    ap.add_argument("--no-rebalance", action="store_true", help="strong scaling: keep the equal-count dealing of the frame blocks")
    ap.add_argument("--frame-block", type=int, default=0, help="strong scaling: consecutive keyframes per dealt block (default 25)")
    ap.add_argument("--sample", default="nearest", choices=["nearest", "bilinear"])
    ap.add_argument("--cloud-upload", default="sharded", choices=["sharded", "replicated"],
                    help="e2e leg at N > 1: each rank uploads its slice of the map and the slices are all-gathered, or every rank uploads all of it")
    ap.add_argument("--cloud-layout", default="pcl", choices=["pcl", "float4"], help="record layout of the cloud handed over in the e2e leg")
    ap.add_argument("--parity-frames", type=int, default=2, help="N>1: frames per rank of the oracle check of the exchanged accumulators")
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for the cpu_baseline sample")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_native(args)


if __name__ == "__main__":
    sys.exit(main())

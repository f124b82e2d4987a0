#!/usr/bin/env python
"""bench.py — point-frames/sec of the colorization hot path on N B200s (+ roofline, CPU baseline, end-to-end).

A "step" is one pass of the hot path over one batch: every keyframe of the workload projected, culled, sampled and
accumulated over the whole map cloud, the per-point accumulators all-reduced across ranks (N>1) and finalized.
  value : whole-job point-frames/s with the cloud and all images already resident in HBM (CUDA events on the stream the
          kernels are launched on, max over ranks)
  e2e   : the same metric through the C ABI with HOST buffers: cloud upload, one image upload per frame and the
          read-back of the colours are inside the timed region
  roofline / cpu_baseline : see DESIGN.md "Measurement"

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference] [--workload C2]
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


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def _workload(args):
    from pointcloudprocessor_b200 import synth
    n, f, w, h = synth.CONFIGS[args.workload]
    if args.points:
        n = args.points
    if args.frames:
        f = args.frames
    return n, f, w, h


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
# reference arm: the CPU oracle (the reference itself cannot be compiled on this box) on all host cores
# ---------------------------------------------------------------------------------------------------------------------
def _cpu_sample(cloud, poses, images, W, H, intr, dist, threads, budget_s, max_frames):
    """Times the OpenMP oracle on the first k frames of the workload (k sized from a one-frame calibration)."""
    from oracle import pyoracle as O
    cam = O.camera(intr, dist)
    oc = O.Colorizer(cam, W, H, cloud)
    t0 = time.perf_counter()
    oc.frame(poses[0], images[0], threads=threads)
    t1 = time.perf_counter() - t0
    k = int(max(1, min(max_frames, budget_s / max(t1, 1e-6))))
    oc.close()
    oc = O.Colorizer(cam, W, H, cloud)
    t0 = time.perf_counter()
    for i in range(k):
        oc.frame(poses[i % len(poses)], images[i % len(images)], threads=threads)
    oc.finalize_sum()
    dt = time.perf_counter() - t0
    acc = oc.accumulators()[0]
    oc.close()
    return k, dt, acc


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import pyoracle as O
    from pointcloudprocessor_b200 import synth
    O.build()
    n, f, W, H = _workload(args)
    intr, dist = synth.make_camera(W, H)
    cloud = synth.make_cloud(n)
    poses, _ = synth.make_poses(f)
    threads = O.max_threads()
    cam = O.camera(intr, dist)
    oc = O.Colorizer(cam, W, H, cloud)
    img0 = synth.make_image(0, W, H)
    t0 = time.perf_counter()
    oc.frame(poses[0], img0, threads=threads)
    t1 = time.perf_counter() - t0
    total_steps = args.steps + args.warmup
    k = int(max(1, min(f, (150.0 / total_steps) / max(t1, 1e-6))))  # whole run <= ~2.5 min
    images = [synth.make_image(i, W, H) for i in range(k)]
    times = []
    for s in range(total_steps):
        oc.close()
        oc = O.Colorizer(cam, W, H, cloud)
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
    sample = f"first {k} of {f} frames over the full {n}-point cloud per step, OpenMP over points"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": f"{args.workload}: {n} points x {f} frames @ {W}x{H}, colorize + z-buffer culling", "sample_frames_per_step": k},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
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

    n, f, W, H = _workload(args)
    intr, dcoef = synth.make_camera(W, H)
    cam = api.create_camera("pinhole", intr, dcoef)
    prm = api.default_params(image_size=(W, H))
    # weak scaling: every rank colours its own f frames (a different stretch of the path) of the same replicated map
    poses_all, _ = synth.make_poses(f * world)
    poses = poses_all[rank * f:(rank + 1) * f]
    cloud_np = synth.make_cloud(n)

    # pinned host staging (e2e) and device-resident copies (value)
    h_cloud = torch.empty((n, 4), dtype=torch.float32).pin_memory()
    h_cloud.numpy()[...] = cloud_np
    h_imgs = torch.empty((f, H, W, 3), dtype=torch.uint8).pin_memory()
    for i in range(f):
        synth.make_image(rank * f + i, W, H, out=h_imgs.numpy()[i])
    h_rgb = torch.empty((n, 3), dtype=torch.uint8).pin_memory()
    h_cnt = torch.empty((n,), dtype=torch.int32).pin_memory()
    d_cloud = h_cloud.cuda()
    d_imgs = h_imgs.cuda()
    d_rgb = torch.empty((n, 3), dtype=torch.uint8, device="cuda")
    d_cnt = torch.empty((n,), dtype=torch.int32, device="cuda")

    stream = torch.cuda.Stream()
    gc = api.Colorizer(cam, prm, device=local)
    gc.set_stream(stream.cuda_stream)
    gc.set_cloud_device(d_cloud.data_ptr(), n)
    if world > 1:
        ids = [gc.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        gc.nccl_init(ids[0], rank, world)
    img_bytes = H * W * 3
    img_ptrs = [d_imgs.data_ptr() + i * img_bytes for i in range(f)]
    himg_ptrs = [h_imgs.data_ptr() + i * img_bytes for i in range(f)]

    def step_device():
        gc.reset_accumulators()
        for i in range(f):
            gc.submit_frame_device(poses[i], img_ptrs[i], W * 3)
        if world > 1:
            gc.allreduce()
        gc.finalize_device(d_rgb.data_ptr(), d_cnt.data_ptr())

    def step_e2e():
        gc.set_cloud_ptr(h_cloud.data_ptr(), n)          # H2D cloud (resets the accumulators)
        for i in range(f):
            gc.submit_frame_ptr(poses[i], himg_ptrs[i], W * 3)  # H2D image per frame, overlapped on the copy stream
        if world > 1:
            gc.allreduce()
        gc.finalize_ptr(h_rgb.data_ptr(), h_cnt.data_ptr())  # D2H colours + observation counts (synchronises)

    # ---- statistics pass (untimed): visible fraction for the algorithmic-byte count --------------------------------
    gc.reset_accumulators()
    nvis = 0
    stat_frames = min(f, 16)
    for i in range(stat_frames):
        r = gc.submit_frame_device(poses[i], img_ptrs[i], W * 3, want_counts=True)
        nvis += r["n_visible"]
    f_vis = nvis / (stat_frames * n)

    # ---- value: HBM-resident ------------------------------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step_device()
    gc.synchronize()
    torch.cuda.synchronize()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    gc.profile_enable(True)
    l0 = gc.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    ev0.record(stream)
    for _ in range(args.steps):
        step_device()
    ev1.record(stream)
    gc.synchronize()
    torch.cuda.synchronize()
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = gc.launch_count() - l0
    prof = gc.profile_read()
    gc.profile_enable(False)
    clocks = sampler.stop() if sampler else None
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = n * f * world * args.steps / (ms * 1e-3)

    # ---- e2e: host buffers through the C ABI --------------------------------------------------------------------------
    e2e_steps = max(1, min(args.steps, 10))
    step_e2e()
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    gc.synchronize()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    barrier()
    e2e_value = n * f * world * e2e_steps / e2e_s
    h2d = n * 16 + f * img_bytes
    d2h = n * 3 + n * 4
    rgb_e2e = h_rgb.numpy().copy()
    rgb_dev = d_rgb.cpu().numpy()
    e2e_matches_device = bool(np.array_equal(rgb_e2e, rgb_dev))

    # ---- roofline of the dominant kernel ------------------------------------------------------------------------------
    # The timed region above runs frames_in_flight frames concurrently (kernels of neighbouring frames overlap), so the
    # per-kernel durations recorded there include time shared with other kernels.  The roofline numbers come from a second
    # timed pass over the same workload with frames_in_flight = 1 (one kernel at a time), CUDA events on the launching stream.
    peak, peak_src = _peaks()
    frames_prof = max(prof["frames"], 1)
    kern_ms_pipelined = {"scan": prof["scan_ms"] / frames_prof, "resolve": prof["resolve_ms"] / frames_prof}
    prm1 = api.default_params(image_size=(W, H), frames_in_flight=1)
    gc1 = api.Colorizer(cam, prm1, device=local)
    gc1.set_stream(stream.cuda_stream)
    gc1.set_cloud_device(d_cloud.data_ptr(), n)
    serial_steps = max(1, min(args.steps, 5))

    def step_serial():
        gc1.reset_accumulators()
        for i in range(f):
            gc1.submit_frame_device(poses[i], img_ptrs[i], W * 3)
        gc1.finalize_device(d_rgb.data_ptr(), d_cnt.data_ptr())

    for _ in range(2):
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
    roofline = {
        "bound": "hbm", "kernel": "classify_kernel + scan_binned_kernel (pass 1)" if dominant == "scan" else dominant, "achieved": scan_gbs, "peak": peak, "unit": "GB/s",
        "frac": scan_gbs / peak, "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": scan_bytes,
        "kernel_ms_per_frame": kern_ms, "kernel_share_of_frame": {k: v / frame_ms for k, v in kern_ms.items()} if frame_ms > 0 else None,
        "pipeline": {"algorithmic_bytes_per_frame": pipe_bytes, "f_vis": f_vis, "achieved": pipe_gbs, "frac": pipe_gbs / peak},
        "note": "achieved = ALGORITHMIC bytes (every point of the cloud, SURVEY.md 8d) / time; the pre-binned scan never reads chunks that are "
                "certainly culled, so DRAM traffic (`traffic`, ncu) is below the algorithmic bytes and frac is not capped at 1",
        "measured_on": f"second timed pass, {serial_steps} steps with frames_in_flight=1 (kernels not overlapped): {serial_ms_per_step:.3f} ms/step",
        "kernel_ms_per_frame_in_timed_region": kern_ms_pipelined,
        "pipeline_in_timed_region": {"achieved": pipe_bytes * f / (ms / args.steps * 1e-3) / 1e9, "frac": pipe_bytes * f / (ms / args.steps * 1e-3) / 1e9 / peak,
                                     "note": "algorithmic bytes of all frames / ms_per_step of the headline (overlapped) run"},
    }
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):
        try:
            roofline["traffic"] = json.load(open(tr)).get("scan_kernel_dram_bytes_per_launch")
        except Exception:
            pass

    # ---- CPU baseline (rank 0, N=1 only) + parity of the sample -------------------------------------------------------
    cpu = None
    parity = None
    if world == 1 and not args.no_cpu:
        from oracle import pyoracle as O
        threads = O.max_threads()
        images_np = h_imgs.numpy()
        k, dt, acc_o = _cpu_sample(cloud_np, poses, images_np, W, H, intr, dcoef, threads, args.cpu_budget, f)
        cpu = {"value": n * k / dt, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"first {k} of {f} frames over the full {n}-point cloud, OpenMP over points, {dt:.1f} s"}
        gc.reset_accumulators()
        for i in range(k):
            gc.submit_frame_device(poses[i], img_ptrs[i], W * 3)
        acc_g, _, _ = gc.accumulators()
        parity = {"frames": k, "accumulators_bit_exact_vs_oracle": bool(np.array_equal(acc_g, acc_o))}

    gc.close()
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"{args.workload}: {n} points x {f} frames @ {W}x{H} per GPU, colorize + z-buffer culling, SUM accumulators",
                       "points": n, "frames_per_gpu": f, "image": [W, H], "parallelism": f"frames sharded x{world}, cloud replicated, NCCL allreduce u32",
                       "frames_in_flight": int(prm.frames_in_flight),
                       "l2": "inputs larger than L2 (cloud 16 B x N + one image per frame); no explicit flush",
                       "host_affinity": (f"rank bound to the {numa} CPUs NVML lists for its GPU" if numa else "unbound")},
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": e2e_steps,
                    "colours_equal_device_path": e2e_matches_device},
            "roofline": roofline,
        }
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
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--points", type=int, default=0)
    ap.add_argument("--frames", type=int, default=0)
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for the cpu_baseline sample")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_native(args)


if __name__ == "__main__":
    sys.exit(main())

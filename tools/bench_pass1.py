#!/usr/bin/env python
"""Kernel-variant measurement of the per-frame kernels (pass 1 = classify + scan, resolve) on one GPU.

For the library named by PCP_LIB_PATH (default: the in-tree libpcp_b200.so) it prints ONE JSON line with
  serial   : per-frame kernel times with frames_in_flight = 1 (CUDA events between the kernels, nothing overlapped)
  pipelined: ms per frame with frames_in_flight = 3 (the headline mode)
  sha      : SHA-1 of the u32 accumulators after all frames (equal across variants <=> same colours; the default library is
             checked against the oracle by tests/ and bench.py)
   PCP_LIB_PATH=pointcloudprocessor_b200/var_x.so python tools/bench_pass1.py [--workload C2] [--frames 60] [--reps 5]
"""
import argparse
import hashlib
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch

from pointcloudprocessor_b200 import _capi, api, synth


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--points", type=int, default=0)
    ap.add_argument("--frames", type=int, default=60)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--sample", default="nearest", choices=["nearest", "bilinear"])
    ap.add_argument("--tag", default="")
    ap.add_argument("--flight", type=int, default=3, help="an extra frames_in_flight setting to time (1 and 3 always run)")
    ap.add_argument("--group", type=int, default=0, help="frames_per_launch of the extra setting (0 = the library's choice)")
    ap.add_argument("--per-frame-calls", action="store_true", help="one pcp_submit_frame_device call per frame instead of the loop call")
    a = ap.parse_args()
    n, f, W, H = synth.CONFIGS[a.workload]
    n = a.points or n
    f = min(a.frames, f)
    intr, dc = synth.make_camera(W, H)
    cam = api.create_camera("pinhole", intr, dc)
    poses, _ = synth.make_poses(f)
    cloud = torch.from_numpy(synth.make_cloud(n)).cuda()
    imgs = torch.empty((f, H, W, 3), dtype=torch.uint8)
    for i in range(f):
        synth.make_image(i, W, H, out=imgs.numpy()[i])
    d_imgs = imgs.cuda()
    ib = H * W * 3
    out = {"lib": os.path.basename(_capi.LIB_PATH), "tag": a.tag, "workload": f"{a.workload}: {n} x {f} @ {W}x{H}", "sample": a.sample}
    stream = torch.cuda.Stream()
    sm = api.SAMPLE_BILINEAR if a.sample == "bilinear" else api.SAMPLE_NEAREST
    for flight in (1, 3) + ((a.flight,) if a.flight not in (1, 3) else ()):  # --flight 0: slots and frames per launch chosen by the library
        prm = api.default_params(image_size=(W, H), frames_in_flight=flight, sample_mode=sm,
                                 frames_per_launch=(a.group if flight == a.flight else 1))
        gc = api.Colorizer(cam, prm, device=0)
        gc.set_stream(stream.cuda_stream)
        gc.set_cloud_device(cloud.data_ptr(), n)

        ptrs = np.array([d_imgs.data_ptr() + i * ib for i in range(f)], np.uint64)

        def step():
            gc.reset_accumulators()
            if a.per_frame_calls:
                for i in range(f):
                    gc.submit_frame_device(poses[i], d_imgs.data_ptr() + i * ib, W * 3)
            else:
                gc.submit_frames_device(poses[:f], ptrs, W * 3)

        for _ in range(2):
            step()
        gc.synchronize()
        gc.profile_enable(flight == 1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        import time
        e0.record(stream)
        t0 = time.perf_counter()
        for _ in range(a.reps):
            step()
        host_us = 1e6 * (time.perf_counter() - t0) / (a.reps * f)  # time the submitting thread spends per frame (launches, events)
        e1.record(stream)
        gc.synchronize()
        torch.cuda.synchronize()
        per_frame_us = 1e3 * e0.elapsed_time(e1) / (a.reps * f)
        if flight == 1:
            p = gc.profile_read()
            fr = max(p["frames"], 1)
            out["serial"] = {"frame_us": round(per_frame_us, 2), "pass1_us": round(1e3 * p["scan_ms"] / fr, 2), "resolve_us": round(1e3 * p["resolve_ms"] / fr, 2),
                             "pass1_frac_of_hbm": round(16.0 * n / (p["scan_ms"] / fr * 1e-3) / 1e9 / 6538.0, 4)}
            acc, _, _ = gc.accumulators()
            out["sha"] = hashlib.sha1(acc.tobytes()).hexdigest()[:16]
            out["counters_last_frame"] = gc.debug_counters()
        else:
            out["pipelined" if flight == 3 else f"pipelined_flight{flight}_group{a.group}"] = {"frame_us": round(per_frame_us, 2), "host_submit_us": round(host_us, 2), "pf_per_s": float(f"{n / (per_frame_us * 1e-6):.4g}")}
            acc, _, _ = gc.accumulators()
            out["sha3"] = hashlib.sha1(acc.tobytes()).hexdigest()[:16]
        gc.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()

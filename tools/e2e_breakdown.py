#!/usr/bin/env python
"""Where the end-to-end (host buffers through the C ABI) time of one C2 step goes: cloud upload + pre-binning, the frame loop with
per-frame image uploads, finalize + download; plus raw pinned H2D/D2H bandwidth and the scan statistics of a PCP_STATS build.
   python tools/e2e_breakdown.py [--points N] [--frames F]"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch

from pointcloudprocessor_b200 import api, synth


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=10_000_000)
    ap.add_argument("--frames", type=int, default=200)
    a = ap.parse_args()
    n, f, W, H = a.points, a.frames, 1920, 1080
    intr, dc = synth.make_camera(W, H)
    cam = api.create_camera("pinhole", intr, dc)
    prm = api.default_params(image_size=(W, H))
    poses, _ = synth.make_poses(f)
    h_cloud = torch.empty((n, 4), dtype=torch.float32).pin_memory()
    h_cloud.numpy()[...] = synth.make_cloud(n)
    h_imgs = torch.empty((f, H, W, 3), dtype=torch.uint8).pin_memory()
    for i in range(f):
        synth.make_image(i, W, H, out=h_imgs.numpy()[i])
    h_rgb = torch.empty((n, 3), dtype=torch.uint8).pin_memory()
    h_cnt = torch.empty((n,), dtype=torch.int32).pin_memory()
    # raw copies
    d = torch.empty_like(h_imgs, device="cuda")
    for _ in range(2):
        torch.cuda.synchronize(); t = time.perf_counter(); d.copy_(h_imgs, non_blocking=True); torch.cuda.synchronize()
        dt = time.perf_counter() - t
    print(f"raw H2D pinned {h_imgs.numel() / 1e9:.2f} GB: {dt * 1e3:.1f} ms = {h_imgs.numel() / dt / 1e9:.1f} GB/s")
    t = time.perf_counter(); h_imgs.copy_(d, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t
    print(f"raw D2H pinned: {dt * 1e3:.1f} ms = {h_imgs.numel() / dt / 1e9:.1f} GB/s")
    del d
    gc = api.Colorizer(cam, prm, device=0)
    ib = H * W * 3
    for rep in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        gc.set_cloud_ptr(h_cloud.data_ptr(), n); gc.synchronize(); t1 = time.perf_counter()
        for i in range(f):
            gc.submit_frame_ptr(poses[i], h_imgs.data_ptr() + i * ib, W * 3)
        t2 = time.perf_counter(); gc.synchronize(); t3 = time.perf_counter()
        gc.finalize_ptr(h_rgb.data_ptr(), h_cnt.data_ptr()); t4 = time.perf_counter()
        print(f"rep {rep}: set_cloud {1e3 * (t1 - t0):.1f} ms | submit loop (host) {1e3 * (t2 - t1):.1f} ms, drained after {1e3 * (t3 - t1):.1f} ms | "
              f"finalize+D2H {1e3 * (t4 - t3):.1f} ms | total {1e3 * (t4 - t0):.1f} ms")
    # statistics (meaningful with a -DPCP_STATS build)
    tot = {}
    for i in range(0, f, max(1, f // 20)):
        gc.submit_frame_ptr(poses[i], h_imgs.data_ptr() + i * ib, W * 3)
        c = gc.debug_counters()
        for k, v in c.items():
            tot.setdefault(k, []).append(v)
    print("mean counters per frame:", {k: float(np.mean(v)) for k, v in tot.items()}, "chunks", (n + 127) // 128)
    gc.close()


if __name__ == "__main__":
    main()

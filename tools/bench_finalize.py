#!/usr/bin/env python
"""pcp_finalize_device on one GPU: the packed two-kernel form against the one-kernel gather (PCP_DEBUG_FINALIZE_ONEPASS), CUDA events.

Prints one JSON line per cloud size: us per call of both forms (colours + counts, colours only), the algorithmic bytes (16 B sums +
4 B inv + 3 B colour [+ 4 B count] per point) as a fraction of the HBM peak, and whether both forms produced the same bytes.
   python tools/bench_finalize.py [--points 10000000 100000000] [--frames 20] [--reps 20]
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
    ap.add_argument("--points", type=int, nargs="+", default=[10_000_000])
    ap.add_argument("--frames", type=int, default=20)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--order", default="random", choices=["random", "voxel"],
                    help="caller's point order: synth's (random) or the order of a voxel-grid filter's output (cells in raster order)")
    ap.add_argument("--peak", type=float, default=6452.0, help="HBM copy peak, GB/s (MEASURED_PEAKS.json on this pool)")
    a = ap.parse_args()
    W, H = 1920, 1080
    intr, dc = synth.make_camera(W, H)
    cam = api.create_camera("pinhole", intr, dc)
    poses, _ = synth.make_poses(a.frames)
    imgs = torch.empty((a.frames, H, W, 3), dtype=torch.uint8)
    for i in range(a.frames):
        synth.make_image(i, W, H, out=imgs.numpy()[i])
    d_imgs = imgs.cuda()
    ptrs = np.array([d_imgs.data_ptr() + i * H * W * 3 for i in range(a.frames)], np.uint64)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.Stream()  # the context's stream: events on any other stream would not bracket its kernels
    for n in a.points:
        host = synth.make_cloud(n)
        if a.order == "voxel":  # pcl::VoxelGrid writes its output cell by cell: neighbours in the file are neighbours in space
            cell = np.floor(host[:, :3] / 0.5).astype(np.int64)
            cell -= cell.min(axis=0)
            ext = cell.max(axis=0) + 1
            host = host[np.argsort((cell[:, 2] * ext[1] + cell[:, 1]) * ext[0] + cell[:, 0], kind="stable")]
        cloud = torch.from_numpy(np.ascontiguousarray(host)).cuda()
        rgb = torch.empty(n * 3 + 16, dtype=torch.uint8, device="cuda")
        cnt = torch.empty(n, dtype=torch.int32, device="cuda")
        out = {"points": n, "frames": a.frames, "order": a.order}
        shas = {}
        for name, flags in (("two_pass", _capi.DEBUG_FINALIZE_TWOPASS), ("one_pass", _capi.DEBUG_FINALIZE_ONEPASS)):
            prm = api.default_params(image_size=(W, H), debug_flags=flags)
            gc = api.Colorizer(cam, prm, device=0)
            gc.set_stream(stream.cuda_stream)
            gc.set_cloud_device(cloud.data_ptr(), n)
            gc.submit_frames_device(poses, ptrs, W * 3)
            gc.synchronize()
            cases = [("rgb_count", rgb.data_ptr(), cnt.data_ptr()), ("rgb", rgb.data_ptr(), 0)]
            if name == "one_pass":
                cases.append(("bytes_rgb_count", rgb.data_ptr() + 1, cnt.data_ptr()))  # unaligned: the one-point-per-thread kernel of round 1
            for what, rp, cp in cases:
                for _ in range(3):
                    gc.finalize_device(rp, cp)
                gc.synchronize()
                tot = 0.0
                for _ in range(a.reps):
                    flush.zero_()  # the accumulators come from HBM, as after a long run of frames
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(stream)
                    gc.finalize_device(rp, cp)
                    e1.record(stream)
                    torch.cuda.synchronize()
                    tot += e0.elapsed_time(e1)
                us = 1e3 * tot / a.reps
                bytes_pt = 23 + (4 if cp else 0)
                out[f"{name}_{what}_us"] = round(us, 1)
                out[f"{name}_{what}_frac"] = round(bytes_pt * n / (us * 1e-6) / 1e9 / a.peak, 3)
            gc.finalize_device(rgb.data_ptr(), cnt.data_ptr())
            gc.synchronize()
            torch.cuda.synchronize()
            shas[name] = hashlib.sha1(rgb[: n * 3].cpu().numpy().tobytes() + cnt.cpu().numpy().tobytes()).hexdigest()[:16]
            out[f"{name}_coloured"] = int((cnt > 0).sum().item())
            gc.close()
        out["same_bytes"] = shas["two_pass"] == shas["one_pass"]
        out["sha"] = shas["two_pass"]
        print(json.dumps(out), flush=True)
        del cloud, rgb, cnt


if __name__ == "__main__":
    main()

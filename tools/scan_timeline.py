#!/usr/bin/env python
"""Where the time of ONE scan_ring_kernel launch goes, CTA by CTA (needs the -DPCP_TIMELINE build of the library: PCP_LIB_PATH).

Every CTA stamps %globaltimer at: start, after the programmatic-launch wait, scout done with its static share, scout done with the
pool, worker warp 0's first chunk arrived, worker warp 0 done.  Frames run one at a time (frames_in_flight = 1); the stamps of the
last frame are read back.  Prints percentiles over the CTAs, in us from the first CTA's start.
   python - <<< 'from pointcloudprocessor_b200 import build; build.build_library(force=True, defines=("PCP_TIMELINE",), out="pointcloudprocessor_b200/var_timeline.so")'
   PCP_LIB_PATH=$PWD/pointcloudprocessor_b200/var_timeline.so python tools/scan_timeline.py [--workload C2] [--frames 12]
"""
import argparse
import ctypes as C
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
    ap.add_argument("--frames", type=int, default=12)
    a = ap.parse_args()
    n, f, W, H = synth.CONFIGS[a.workload]
    f = min(a.frames, f)
    intr, dc = synth.make_camera(W, H)
    cam = api.create_camera("pinhole", intr, dc)
    poses, _ = synth.make_poses(f)
    cloud = torch.from_numpy(synth.make_cloud(n)).cuda()
    imgs = torch.empty((f, H, W, 3), dtype=torch.uint8)
    for i in range(f):
        synth.make_image(i, W, H, out=imgs.numpy()[i])
    d_imgs = imgs.cuda()
    L = _capi.load()
    L.pcp_debug_timeline.argtypes = [C.c_void_p, C.c_size_t]
    prm = api.default_params(image_size=(W, H), frames_in_flight=1)
    out = []
    with api.Colorizer(cam, prm, device=0) as gc:
        gc.set_cloud_device(cloud.data_ptr(), n)
        for rep in range(3):
            gc.reset_accumulators()
            for i in range(f):
                gc.submit_frame_device(poses[i], d_imgs.data_ptr() + i * H * W * 3, W * 3)
                gc.synchronize()
            grid = 148 * 4
            tl = np.zeros((grid, 8), np.uint64)
            assert L.pcp_debug_timeline(tl.ctypes.data, grid) == 0
            t = tl[:, :6].astype(np.int64)
            t0 = t[:, 0].min()
            us = (t - t0) / 1e3
            names = ["cta start", "after launch wait", "scout: static share done", "scout: pool done", "worker 0: first chunk here", "worker 0: done"]
            row = {"rep": rep, "counters": gc.debug_counters()}
            for k, nm in enumerate(names):
                v = us[:, k]
                row[nm] = {"min": round(float(v.min()), 2), "p50": round(float(np.median(v)), 2), "p90": round(float(np.percentile(v, 90)), 2), "max": round(float(v.max()), 2)}
            row["chunks published per CTA (static / total)"] = {"static_mean": float(tl[:, 6].mean()), "total_mean": float(tl[:, 7].mean()),
                                                               "total_min": int(tl[:, 7].min()), "total_max": int(tl[:, 7].max())}
            sm_end = np.maximum(us[:, 3], us[:, 5])
            row["kernel span us (first start to last milestone)"] = round(float(sm_end.max()), 2)
            out.append(row)
    print(json.dumps(out[-1], indent=1))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Scan statistics of a PCP_STATS build (PCP_LIB_PATH=pointcloudprocessor_b200/var_stats.so): per frame, the active chunks, the
chunks skipped whole, the points kept by stages A0 / A1, records and visible points, averaged over the C2 path.
   PCP_LIB_PATH=... python tools/scan_stats.py [--points N] [--frames F] [--flight K]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np

from pointcloudprocessor_b200 import api, synth


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=10_000_000)
    ap.add_argument("--frames", type=int, default=40)
    ap.add_argument("--flight", type=int, default=3)
    a = ap.parse_args()
    W, H = 1920, 1080
    intr, dc = synth.make_camera(W, H)
    cam = api.create_camera("pinhole", intr, dc)
    prm = api.default_params(image_size=(W, H))
    prm.frames_in_flight = a.flight
    poses, _ = synth.make_poses(a.frames)
    cloud = synth.make_cloud(a.points)
    img = synth.make_image(0, W, H)
    tot = {}
    with api.Colorizer(cam, prm, device=0) as gc:
        gc.set_cloud(cloud)
        for i in range(a.frames):
            gc.submit_frame(poses[i], img)
            c = gc.debug_counters()
            if i >= a.flight:  # the first frame of every slot has no seed points
                for k, v in c.items():
                    tot[k] = tot.get(k, 0) + v
    m = a.frames - a.flight
    print({k: round(v / m) for k, v in tot.items()}, "seed" if os.environ.get("PCP_NO_SEED") != "1" else "no seed")


if __name__ == "__main__":
    main()

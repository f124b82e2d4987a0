#!/usr/bin/env python
"""How far the reference's kd-tree re-indexing (src/PointCloudProcessor.cpp:557-593, oracle reindex="radius") is from carrying
the point index (oracle reindex="carry" = the CUDA path, decision D11), on small clouds (the radius mode is brute force).

The reference sends every coloured camera-frame point back to the world with a FLOAT c2w and gives the observation to every
map point within 1e-5 m of the result.  Near the origin the float round trip lands within 1e-5 m and both modes colour the same
points; at map-scale coordinates one float ulp is already > 1e-5 m / 3 and most observations find no map point at all, so the
reference leaves those points uncoloured and removePointsWithNoColor (hpp:238-252) deletes them.

    python tools/reindex_divergence.py [--points 20000] [--frames 6]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np

from oracle import pyoracle as O
from pointcloudprocessor_b200 import synth


def run(cloud, poses, W, H, intr, dist, mode):
    oc = O.Colorizer(O.camera(intr, dist), W, H, cloud, keep_obs=True, reindex=mode)
    for i, p in enumerate(poses):
        oc.frame(p, synth.make_image(i, W, H))
    rgb = oc.finalize_top5()
    cnt = oc.obs_counts()
    oc.close()
    return rgb, cnt


def compare(cloud, poses, W, H, intr, dist):
    rc, cc = run(cloud, poses, W, H, intr, dist, "carry")
    rr, cr = run(cloud, poses, W, H, intr, dist, "radius")
    seen = cc > 0
    both = seen & (cr > 0)
    d = np.abs(rc[both].astype(int) - rr[both].astype(int)).max(axis=1) if both.any() else np.zeros(0, int)
    return {"points_observed_carry": int(seen.sum()), "points_observed_radius": int((cr > 0).sum()),
            "observed_by_carry_lost_by_radius": int((seen & (cr == 0)).sum()), "observed_only_by_radius": int((~seen & (cr > 0)).sum()),
            "observations_carry": int(cc.sum()), "observations_radius": int(cr.sum()),
            "colour_differs_among_common": int((d > 0).sum()), "max_level_difference": int(d.max()) if d.size else 0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=20_000)
    ap.add_argument("--frames", type=int, default=6)
    a = ap.parse_args()
    W, H = 320, 240
    intr, dist = synth.make_camera(W, H)
    cloud = synth.make_cloud(a.points)
    poses, _ = synth.make_poses(a.frames)
    out = {"scene": f"{a.points} points x {a.frames} frames @ {W}x{H}, TOP5"}
    out["corridor (|xyz| <= 20 m)"] = compare(cloud, poses, W, H, intr, dist)
    for shift in (100.0, 1000.0):  # the same scene moved away from the origin: map-scale coordinates
        c2 = cloud.copy(); c2[:, 0] += np.float32(shift); c2[:, 1] += np.float32(shift)
        p2 = poses.copy(); p2[:, 0] += shift; p2[:, 1] += shift
        out[f"shifted by {shift:g} m in x and y"] = compare(c2, p2, W, H, intr, dist)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()

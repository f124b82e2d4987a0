#!/usr/bin/env python
"""Information only (SURVEY.md §8 row a-3'): how far the z-buffer visibility this build implements (view_culling.cpp:52-174, the
oracle) is from the Katz hidden-point removal the reference has switched on (view_culling.cpp:266-334): spherical flip with
R = 9e4, convex hull of the flipped points and the origin (qhull through scipy), visible = hull vertices.  CPU only.

  python tools/hpr_agreement.py [points] [frames]        default: the C1 shape (1e6 points, 1280x720), 3 frames
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def katz_hpr(pts_cam, cam, W, H, oracle, R=90000.0):
    """-> boolean mask over pts_cam, as hidden_points_removal returns indices."""
    from scipy.spatial import ConvexHull
    p = pts_cam.astype(np.float64)
    front = p[:, 2] / np.maximum(np.linalg.norm(p, axis=1), 1e-300) >= 0.0          # :277
    uv = oracle.project(cam, p)                                                       # :283
    iu, iv = np.trunc(uv[:, 0]), np.trunc(uv[:, 1])
    with np.errstate(invalid="ignore"):
        inside = front & (iu >= 0) & (iv >= 0) & (iu < W) & (iv < H)                  # :284-288
    idx = np.nonzero(inside)[0]
    q = p[idx]
    nrm = np.linalg.norm(q, axis=1, keepdims=True)
    flipped = q + 2.0 * (R - nrm) * q / nrm                                           # :292
    hull = ConvexHull(np.vstack([flipped, np.zeros((1, 3))]))                         # :295-306, "qhull " = default options
    v = hull.vertices[hull.vertices < len(idx)]
    out = np.zeros(len(p), bool)
    out[idx[v]] = True
    return out, inside


def main():
    from oracle import pyoracle as oracle
    from pointcloudprocessor_b200 import synth
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000
    f = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    W, H = 1280, 720
    intr, dist = synth.make_camera(W, H)
    cam = oracle.camera(intr, dist)
    cloud = synth.make_cloud(n, 42)
    poses, _ = synth.make_poses(20)
    print(f"{n} points, {W}x{H}; per frame: in image | z-buffer visible | HPR visible | both | Jaccard | z-buffer points HPR also keeps")
    for k in np.linspace(0, 19, f).astype(int):
        w2c, _ = oracle.pose_matrices(poses[k])
        pc = oracle.transform(cloud[:, :3].copy(), w2c)
        z = np.zeros(n, bool)
        z[oracle.view_culling(cam, W, H, pc)] = True
        h, inside = katz_hpr(pc, cam, W, H, oracle)
        both = int((z & h).sum())
        print(f"frame {k:2d}: {int(inside.sum()):8d} | {int(z.sum()):8d} | {int(h.sum()):8d} | {both:8d} | {both / max(1, int((z | h).sum())):.3f} | {both / max(1, int(z.sum())):.3f}")


if __name__ == "__main__":
    main()

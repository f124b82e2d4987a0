"""Runs pieces of the REFERENCE's own Python source, read from /root/reference at generation time (never copied into the repo).

generate_distance_masks (scripts/genNormAndDistanceMask.py:233-265) is the one function of the reference on or next to the hot
path that can execute in this container: its body needs only numpy and cv2, but the module imports open3d / plantcv at the top,
so a plain `import` fails.  The method is cut out of the module's syntax tree (`ast`), compiled on its own and called on a stub
`self` that carries exactly the attributes the body reads.  What comes back was computed by the reference's statements, not by a
transcription of them.

Only tests/golden/make_golden.py (fixture generation, in the container that has /root/reference) and the CPU test that re-checks the
fixture when the reference tree is present call this.  Nothing here runs on the GPU box.
"""
from __future__ import annotations

import ast
import copy
import logging
import os

import numpy as np

REFERENCE_SCRIPT = "/root/reference/PointCloudProcessor/scripts/genNormAndDistanceMask.py"


def available() -> bool:
    return os.path.exists(REFERENCE_SCRIPT)


def _extract_method(path: str, name: str):
    """(function object compiled from the reference's source of method `name`, first line, last line)."""
    src = open(path).read()
    tree = ast.parse(src)
    for node in ast.walk(tree):
        if isinstance(node, ast.ClassDef):
            for item in node.body:
                if isinstance(item, ast.FunctionDef) and item.name == name:
                    mod = ast.Module(body=[item], type_ignores=[])
                    ast.fix_missing_locations(mod)
                    import cv2  # the body's only non-numpy dependency
                    ns = {"np": np, "cv2": cv2, "logger": logging.getLogger("reference"), "copy": copy}
                    exec(compile(mod, path, "exec"), ns)  # noqa: S102 - the reference's own statements are what must run
                    return ns[name], item.lineno, item.end_lineno
    raise LookupError(f"{name} not found in {path}")


class _Pcd:
    """Stand-in for the open3d cloud: the body only evaluates np.asarray(pcd.points)."""

    def __init__(self, points):
        self.points = np.asarray(points, np.float64)  # open3d stores doubles


class _Frame:
    def __init__(self, points, H, W, ts):
        self.pcd = _Pcd(points)
        self.crack_mask = np.zeros((H, W), np.uint8)
        self.frame_timestamp = ts
        self.dist_mask = None
        self.pts_3d_mask = None

    def add_dist_mask(self, m):
        self.dist_mask = m

    def add_pts_3d_mask(self, m):
        self.pts_3d_mask = m


class _Self:
    def __init__(self, frames, K, D):
        self.frames = frames
        self.intrinsic_matrix = K
        self.distortion_coefficients = D


def reference_distance_masks(points_cam, W, H, intr, dist):
    """The reference's generate_distance_masks on one frame's camera-frame cloud -> (distance_mask HxW f32, points_3d_mask HxWx3 f32,
    (first, last) source lines that were executed)."""
    fn, l0, l1 = _extract_method(REFERENCE_SCRIPT, "generate_distance_masks")
    K = np.array([[intr[0], 0.0, intr[2]], [0.0, intr[1], intr[3]], [0.0, 0.0, 1.0]], np.float64)
    D = np.asarray(list(dist) + [0.0] * (5 - len(dist)), np.float64)
    fr = _Frame(points_cam, H, W, 0.0)
    fn(_Self([fr], K, D))
    return fr.dist_mask, fr.pts_3d_mask, (l0, l1)

"""ctypes binding of the CPU oracle (oracle/pcp_oracle.cpp).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module; nothing under pointcloudprocessor_b200/ does.  Parity is UNPINNED by the reference (it has no tests
or golden vectors and cannot be built here); see the header of pcp_oracle.cpp for the decision log.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libpcp_oracle.so")


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "pcp_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"], env={**os.environ, "CXX": "g++"})
    return _LIB_PATH


class OrcCamera(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("fx", "fy", "cx", "cy", "k1", "k2", "p1", "p2", "k3")]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_max_threads.restype = C.c_int
        L.orc_view_culling.restype = C.c_int
        L.orc_colorizer_create.restype = C.c_void_p
        L.orc_nid.restype = C.c_int
        L.orc_nid_gate.restype = C.c_int
        _lib = L
    return _lib


def _p(a, t=C.c_void_p):
    return None if a is None else a.ctypes.data_as(t)


def camera(intr, dist) -> OrcCamera:
    d = list(dist) + [0.0] * (5 - len(dist))
    return OrcCamera(float(intr[0]), float(intr[1]), float(intr[2]), float(intr[3]), *[float(x) for x in d[:5]])


def max_threads() -> int:
    return int(lib().orc_max_threads())


def pose_matrices(pose7, T_opt=None):
    """-> (w2c, c2w) as 4x4 float32 numpy (row/col indexed normally)."""
    pose7 = np.ascontiguousarray(pose7, np.float64)
    w2c = np.zeros(16, np.float32)
    c2w = np.zeros(16, np.float32)
    T = None if T_opt is None else np.ascontiguousarray(np.asarray(T_opt, np.float64).T.reshape(-1))  # col-major
    lib().orc_pose_matrices(_p(pose7), _p(T), _p(w2c), _p(c2w))
    return w2c.reshape(4, 4).T.copy(), c2w.reshape(4, 4).T.copy()


def transform(xyz, M):
    xyz = np.ascontiguousarray(xyz, np.float32)
    n, s = xyz.shape
    Mc = np.ascontiguousarray(np.asarray(M, np.float32).T.reshape(-1))
    out = np.empty((n, 3), np.float32)
    lib().orc_transform(_p(xyz), C.c_size_t(n), C.c_size_t(s), _p(Mc), _p(out))
    return out


def project(cam: OrcCamera, xyz):
    xyz = np.ascontiguousarray(xyz, np.float64)
    uv = np.empty((xyz.shape[0], 2), np.float64)
    lib().orc_project(C.byref(cam), _p(xyz), C.c_size_t(xyz.shape[0]), _p(uv))
    return uv


def view_culling(cam, W, H, pts_cam, T=None, ds=14, tol=0.05, depth_culling=True, want_maps=False):
    pts_cam = np.ascontiguousarray(pts_cam, np.float32)
    n = pts_cam.shape[0]
    out = np.empty(max(n, 1), np.int32)
    Wc, Hc = W // ds, H // ds
    dm = np.empty((Hc, Wc), np.float32)
    im = np.empty((Hc, Wc), np.int32)
    ncand = C.c_int(0)
    Tc = None if T is None else np.ascontiguousarray(np.asarray(T, np.float64).T.reshape(-1))
    m = lib().orc_view_culling(C.byref(cam), W, H, ds, C.c_double(tol), int(depth_culling), _p(pts_cam), C.c_size_t(n), _p(Tc),
                               _p(out), _p(dm), _p(im), C.byref(ncand))
    if want_maps:
        return out[:m].copy(), dm, im, ncand.value
    return out[:m].copy()


def color_pixels(cam, W, H, pts_cam):
    pts_cam = np.ascontiguousarray(pts_cam, np.float32)
    pix = np.empty(pts_cam.shape[0], np.int32)
    lib().orc_color_pixels(C.byref(cam), W, H, _p(pts_cam), C.c_size_t(pts_cam.shape[0]), _p(pix))
    return pix


def distance_masks(cam, W, H, pts_cam):
    """generate_distance_masks (scripts/genNormAndDistanceMask.py:233-265) -> (distance HxW f32, xyz HxWx3 f32, index HxW i32)."""
    pts_cam = np.ascontiguousarray(np.asarray(pts_cam, np.float32)[:, :3])
    dist = np.empty((H, W), np.float32)
    xyz = np.empty((H, W, 3), np.float32)
    idx = np.empty((H, W), np.int32)
    lib().orc_distance_masks(C.byref(cam), W, H, _p(pts_cam), C.c_size_t(pts_cam.shape[0]), _p(dist), _p(xyz), _p(idx))
    return dist, xyz, idx


def scores(pts_cam, pose7):
    pts_cam = np.ascontiguousarray(pts_cam, np.float32)
    pose7 = np.ascontiguousarray(pose7, np.float64)
    out = np.empty((pts_cam.shape[0], 3), np.float32)
    lib().orc_scores(_p(pts_cam), C.c_size_t(pts_cam.shape[0]), _p(pose7), _p(out))
    return out


class Colorizer:
    """The loop body of pcdColorizationAndSmooth (src/PointCloudProcessor.cpp:488-595) on the CPU."""

    def __init__(self, cam, W, H, xyzi, cull_size=None, ds=14, tol=0.05, keep_obs=False, bilinear=False, reindex="carry"):
        self.cam, self.W, self.H = cam, W, H
        self.xyzi = np.ascontiguousarray(xyzi, np.float32)
        assert self.xyzi.ndim == 2 and self.xyzi.shape[1] == 4
        self.n = self.xyzi.shape[0]
        cw, ch = cull_size if cull_size is not None else (W, H)
        self._h = C.c_void_p(lib().orc_colorizer_create(C.byref(cam), W, H, cw, ch, ds, C.c_double(tol), _p(self.xyzi),
                                                        C.c_size_t(self.n), int(keep_obs)))
        if bilinear:
            lib().orc_colorizer_set_sample_mode(self._h, 1)
        if reindex not in ("carry", "radius"):
            raise ValueError("reindex must be 'carry' or 'radius'")
        if reindex == "radius":  # src/PointCloudProcessor.cpp:557-593 literally (brute force: small clouds only); decision D11
            assert keep_obs, "reindex='radius' re-indexes the observation lists: keep_obs=True"
            lib().orc_colorizer_set_reindex(self._h, 1)

    def close(self):
        if self._h:
            lib().orc_colorizer_destroy(self._h)
            self._h = None

    __del__ = close

    def frame(self, pose7, bgr, mask=None, T_opt=None, want_vis=False, want_pix=False, threads=1):
        pose7 = np.ascontiguousarray(pose7, np.float64)
        assert bgr.dtype == np.uint8 and bgr.shape[0] == self.H and bgr.shape[1] == self.W and bgr.shape[2] == 3
        assert bgr.strides[2] == 1 and bgr.strides[1] == 3
        T = None if T_opt is None else np.ascontiguousarray(np.asarray(T_opt, np.float64).T.reshape(-1))
        vis = np.empty(self.n, np.uint8) if want_vis else None
        pix = np.empty(self.n, np.int32) if want_pix else None
        counts = np.zeros(3, np.int32)
        mp = 0
        if mask is not None:
            assert mask.dtype == np.uint8 and mask.shape == (self.H, self.W) and mask.strides[1] == 1
            mp = mask.strides[0]
        lib().orc_colorizer_frame(self._h, _p(pose7), _p(T), _p(bgr), C.c_size_t(bgr.strides[0]), _p(mask), C.c_size_t(mp), _p(vis),
                                  _p(pix), int(threads), _p(counts))
        return {"vis": vis, "pix": pix, "n_candidates": int(counts[0]), "n_visible": int(counts[1]), "n_colored": int(counts[2])}

    def accumulators(self):
        acc = np.empty((self.n, 4), np.uint32)
        mh = np.empty(self.n, np.uint32)
        mm = np.empty(self.n, np.uint32)
        lib().orc_colorizer_get_acc(self._h, _p(acc), _p(mh), _p(mm))
        return acc, mh, mm

    def obs_counts(self):
        """Observations per point (RGBCloud::getPointData(i).rgbScores.size()); needs keep_obs."""
        c = np.empty(self.n, np.uint32)
        lib().orc_colorizer_obs_counts(self._h, _p(c))
        return c

    def finalize_sum(self):
        rgb = np.empty((self.n, 3), np.uint8)
        lib().orc_colorizer_finalize_sum(self._h, _p(rgb))
        return rgb

    def finalize_top5(self, top_m=5):
        rgb = np.empty((self.n, 3), np.uint8)
        lib().orc_colorizer_finalize_top5(self._h, _p(rgb), int(top_m))
        return rgb


def hsv_roundtrip(bgr, sat_scale=1.0, val_scale=1.0, block=32):
    """generateColorMap's 8-bit BGR->HSV->BGR image preparation (src/PointCloudProcessor.cpp:722-741)."""
    bgr = np.ascontiguousarray(bgr, np.uint8)
    H, W, _ = bgr.shape
    out = np.empty_like(bgr)
    lib().orc_hsv_roundtrip(_p(bgr), C.c_size_t(bgr.strides[0]), _p(out), C.c_size_t(out.strides[0]), W, H, C.c_float(sat_scale),
                            C.c_float(val_scale), int(block))
    return out


def nid(cam, image_f64, xyzi, se3, bins=16, want_hist=False):
    image_f64 = np.ascontiguousarray(image_f64, np.float64)
    H, W = image_f64.shape
    xyzi = np.ascontiguousarray(xyzi, np.float32)
    se3 = np.ascontiguousarray(se3, np.float64)
    out = C.c_double(0.0)
    hist = np.zeros((bins, bins), np.float64)
    counts = np.zeros(bins, np.float64)
    ok = lib().orc_nid(C.byref(cam), _p(image_f64), W, H, _p(xyzi), C.c_size_t(xyzi.shape[0]), bins, _p(se3), C.byref(out), _p(hist),
                       _p(counts))
    if want_hist:
        return out.value, bool(ok), hist, counts
    return out.value, bool(ok)


def nid_grad(cam, image_f64, xyzi, se3, bins=16):
    """NIDCost::operator()<ceres::Jet<double,7>>: -> (ok, nid, grad[7]) over the ambient Sophus parameters."""
    img = np.ascontiguousarray(image_f64, np.float64)
    H, W = img.shape
    xyzi = np.ascontiguousarray(xyzi, np.float32)
    se3 = np.ascontiguousarray(se3, np.float64)
    out = np.zeros(1, np.float64)
    g = np.zeros(7, np.float64)
    L = lib()
    L.orc_nid_grad.restype = C.c_int
    ok = L.orc_nid_grad(C.byref(cam), _p(img), W, H, _p(xyzi), C.c_size_t(xyzi.shape[0]), int(bins), _p(se3), _p(out), _p(g))
    return bool(ok), float(out[0]), g


def nid_gate(init_se3, se3) -> bool:
    a = np.ascontiguousarray(init_se3, np.float64)
    b = np.ascontiguousarray(se3, np.float64)
    return bool(lib().orc_nid_gate(_p(a), _p(b)))

"""Python harness over the C ABI (numpy in, numpy out).  Names follow the reference's C++ API:
PointCloudProcessor (include/PointCloudProcessor.hpp), vlcal::ViewCulling (include/vlcal/calib/view_culling.hpp),
vlcal::NIDCost (include/vlcal/costs/nid_cost.hpp), camera::create_camera (include/camera/create_camera.hpp).

Every call goes through libpcp_b200.so; nothing here computes on the CPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi
from ._capi import ACCUM_SUM, ACCUM_TOP5, SAMPLE_BILINEAR, SAMPLE_NEAREST, Camera, FrameOut, Params


class PcpError(RuntimeError):
    pass


def create_camera(camera_model: str, intrinsics, distortion_coeffs) -> Camera:
    """camera::create_camera (src/camera/create_camera.cpp:33-49): only "pinhole" exists; missing distortion
    coefficients are zero-filled (create_camera.cpp:23-26)."""
    if camera_model != "pinhole":
        raise PcpError(f"error: unknown camera model {camera_model}")
    if len(intrinsics) != 4:
        raise PcpError("error: num of intrinsic parameters mismatch!!")
    d = [float(x) for x in list(distortion_coeffs)[:5]]
    d += [0.0] * (5 - len(d))
    return Camera(float(intrinsics[0]), float(intrinsics[1]), float(intrinsics[2]), float(intrinsics[3]), (C.c_double * 5)(*d))


def default_params(image_size=None, cull_size=None, **kw) -> Params:
    p = Params()
    _capi.load().pcp_default_params(C.byref(p))
    if image_size is not None:
        p.image_width, p.image_height = int(image_size[0]), int(image_size[1])
        p.cull_width, p.cull_height = p.image_width, p.image_height
    if cull_size is not None:
        p.cull_width, p.cull_height = int(cull_size[0]), int(cull_size[1])
    for k, v in kw.items():
        if not hasattr(p, k):
            raise PcpError(f"unknown param {k}")
        setattr(p, k, v)
    return p


def _ptr(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def cloud_slice(n_total: int, rank: int, world: int):
    """(first, count) of rank's slice of an n_total-point map (pcp_cloud_slice)."""
    f, c = C.c_size_t(0), C.c_size_t(0)
    if _capi.load().pcp_cloud_slice(n_total, rank, world, C.byref(f), C.byref(c)) != 0:
        raise PcpError("pcp_cloud_slice: bad arguments")
    return int(f.value), int(c.value)


class Colorizer:
    """Device-resident state of pcdColorizationAndSmooth (src/PointCloudProcessor.cpp:474-602) on one GPU."""

    def __init__(self, camera: Camera, params: Params, device: int = 0):
        self._L = _capi.load()
        self._h = C.c_void_p()
        rc = self._L.pcp_create(C.byref(self._h), int(device), C.byref(camera), C.byref(params))
        if rc != _capi.PCP_OK:
            msg = self._L.pcp_last_error(None)
            self._h = None
            raise PcpError(f"pcp_create failed ({rc}): {msg.decode() if msg else ''}")
        self.params = params
        self.camera = camera
        self.n = 0
        self._keep = []

    # -- lifecycle -------------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._L.pcp_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc, what):
        if rc != _capi.PCP_OK:
            msg = self._L.pcp_last_error(self._h)
            raise PcpError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    @property
    def handle(self):
        return self._h

    def set_stream(self, cuda_stream_ptr):
        self._check(self._L.pcp_set_stream(self._h, C.c_void_p(cuda_stream_ptr or 0)), "pcp_set_stream")

    def synchronize(self):
        self._check(self._L.pcp_synchronize(self._h), "pcp_synchronize")

    def launch_count(self) -> int:
        return int(self._L.pcp_launch_count(self._h))

    def hsv_roundtrip(self, bgr: np.ndarray, saturation_scale: float = 1.0, brightness_scale: float = 1.0) -> np.ndarray:
        """generateColorMap's 8-bit BGR->HSV->BGR image preparation (src/PointCloudProcessor.cpp:722-741) on the device."""
        if bgr.dtype != np.uint8 or bgr.ndim != 3 or bgr.shape[2] != 3 or bgr.strides[2] != 1 or bgr.strides[1] != 3:
            raise PcpError("hsv_roundtrip: bgr must be HxWx3 uint8 with packed pixels")
        out = np.empty((bgr.shape[0], bgr.shape[1], 3), np.uint8)
        self._check(self._L.pcp_hsv_roundtrip(self._h, _ptr(bgr), bgr.strides[0], _ptr(out), out.strides[0], bgr.shape[1], bgr.shape[0],
                                              C.c_float(saturation_scale), C.c_float(brightness_scale)), "pcp_hsv_roundtrip")
        return out

    def debug_counters(self) -> dict:
        """Device counters of the last submitted frame (see pcp_debug_counters)."""
        out = np.zeros(16, np.uint32)
        self._check(self._L.pcp_debug_counters(self._h, _ptr(out)), "pcp_debug_counters")
        keys = ("records", "candidates", "visible", "colored", "active_chunks", "chunks_skipped", "chunks_fetched", "kept_a1")
        d = {k: int(out[i]) for i, k in enumerate(keys)}
        d["order_locality"] = float(out[14]) / float(out[15]) if out[15] else None
        return d

    # -- cloud -----------------------------------------------------------------------------------------
    def set_cloud(self, points: np.ndarray, intensity_offset: int | None = None):
        """points: (N,4) float32 {x,y,z,intensity} or any (N,k) float32 record array (pcl::PointXYZI: k=8,
        intensity at float index 4)."""
        pts = np.ascontiguousarray(points, np.float32)
        if pts.ndim != 2 or pts.shape[1] < 3:
            raise PcpError("set_cloud: expected (N,k>=3) float32")
        stride = pts.shape[1] * 4
        if intensity_offset is None:
            intensity_offset = 12 if pts.shape[1] == 4 else (16 if pts.shape[1] >= 5 else 8)
        self._check(self._L.pcp_set_cloud(self._h, _ptr(pts), pts.shape[0], stride, intensity_offset), "pcp_set_cloud")
        self.n = pts.shape[0]

    def set_cloud_device(self, dptr: int, n: int):
        self._check(self._L.pcp_set_cloud_device(self._h, C.c_void_p(dptr), n), "pcp_set_cloud_device")
        self.n = n

    # -- frames ----------------------------------------------------------------------------------------
    @staticmethod
    def _mat16(T):
        return None if T is None else np.ascontiguousarray(np.asarray(T, np.float64).T.reshape(-1))

    def submit_frame(self, pose7, bgr: np.ndarray, mask: np.ndarray | None = None, T_opt=None, want_vis=False, want_pix=False,
                     want_counts=False, count_candidates=True, keep_pix_on_device=False):
        pose = np.ascontiguousarray(pose7, np.float64)
        T = self._mat16(T_opt)
        if bgr.dtype != np.uint8 or bgr.ndim != 3 or bgr.shape[2] != 3 or bgr.strides[2] != 1 or bgr.strides[1] != 3:
            raise PcpError("submit_frame: bgr must be HxWx3 uint8 with packed pixels")
        if bgr.shape[0] != self.params.image_height or bgr.shape[1] != self.params.image_width:
            raise PcpError("submit_frame: image size differs from params")
        mp = 0
        if mask is not None:
            if mask.dtype != np.uint8 or mask.shape != bgr.shape[:2] or mask.strides[1] != 1:
                raise PcpError("submit_frame: mask must be HxW uint8 of the image size")
            mp = mask.strides[0]
        out = None
        vis = pix = None
        if want_vis or want_pix or want_counts or keep_pix_on_device:
            out = FrameOut()
            out.flags = (0 if count_candidates else _capi.OUT_SKIP_CANDIDATES) | (_capi.OUT_PIXELS_ON_DEVICE if keep_pix_on_device else 0)
            if want_vis:
                vis = np.zeros((self.n + 31) // 32, np.uint32)
                out.visible_bits = vis.ctypes.data
            if want_pix:
                pix = np.empty(self.n, np.int32)
                out.pixel_index = pix.ctypes.data
        rc = self._L.pcp_submit_frame(self._h, _ptr(pose), _ptr(T), _ptr(bgr), bgr.strides[0], _ptr(mask), mp,
                                      C.byref(out) if out is not None else None)
        self._check(rc, "pcp_submit_frame")
        if out is None:
            self._keep.append((bgr, mask))  # the copy is asynchronous: keep the host buffers alive until synchronize()
            if len(self._keep) > 8:
                self._keep.pop(0)
            return None
        res = {"n_candidates": int(out.n_candidates), "n_visible": int(out.n_visible), "n_colored": int(out.n_colored)}
        if vis is not None:
            res["vis_bits"] = vis
            res["vis"] = np.unpackbits(vis.view(np.uint8), bitorder="little")[: self.n]
        if pix is not None:
            res["pix"] = pix
        return res

    MASK_RECORD = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("b", "u1"), ("g", "u1"), ("r", "u1"), ("a", "u1"), ("mask", "<u2"), ("pad", "<u2")])

    def frame_mask_cloud(self) -> np.ndarray:
        """generateSegmentMap (cpp:769-816) for the frame just submitted (alone, with a mask, with want_pix or keep_pix_on_device):
        the points that received a pixel as PointXYZRGBMask records (camera-frame xyz, bgr, mask; mask 255 -> red), map order."""
        m = C.c_size_t(0)
        self._check(self._L.pcp_frame_mask_cloud(self._h, None, 0, C.byref(m)), "pcp_frame_mask_cloud")
        out = np.empty(m.value, self.MASK_RECORD)
        if m.value:
            self._check(self._L.pcp_frame_mask_cloud(self._h, out.ctypes.data, m.value, C.byref(m)), "pcp_frame_mask_cloud")
        return out

    def submit_frame_device(self, pose7, d_bgr: int, pitch: int, d_mask: int = 0, mask_pitch: int = 0, T_opt=None, want_counts=False):
        pose = np.ascontiguousarray(pose7, np.float64)
        T = self._mat16(T_opt)
        out = FrameOut() if want_counts else None
        rc = self._L.pcp_submit_frame_device(self._h, _ptr(pose), _ptr(T), C.c_void_p(d_bgr), pitch, C.c_void_p(d_mask or 0), mask_pitch,
                                             C.byref(out) if out is not None else None)
        self._check(rc, "pcp_submit_frame_device")
        if out is not None:
            return {"n_candidates": int(out.n_candidates), "n_visible": int(out.n_visible), "n_colored": int(out.n_colored)}
        return None

    def submit_frames_device(self, poses7, d_bgr_ptrs, pitch: int, T_opt=None, d_mask_ptrs=None, mask_pitch: int = 0):
        """The keyframe loop over device-resident images in one call (frame i: poses7[i], image pointer d_bgr_ptrs[i])."""
        poses = np.ascontiguousarray(poses7, np.float64).reshape(-1, 7)
        ptrs = np.ascontiguousarray(d_bgr_ptrs, np.uint64)
        assert ptrs.shape[0] == poses.shape[0]
        mptrs = None if d_mask_ptrs is None else np.ascontiguousarray(d_mask_ptrs, np.uint64)
        T = self._mat16(T_opt)
        rc = self._L.pcp_submit_frames_device(self._h, poses.shape[0], _ptr(poses), _ptr(T), _ptr(ptrs), pitch, _ptr(mptrs), mask_pitch)
        self._check(rc, "pcp_submit_frames_device")

    def set_cloud_ptr(self, host_ptr: int, n: int, stride: int = 16, intensity_offset: int = 12):
        """set_cloud from a raw (e.g. pinned) host pointer."""
        self._check(self._L.pcp_set_cloud(self._h, C.c_void_p(host_ptr), n, stride, intensity_offset), "pcp_set_cloud")
        self.n = n

    def set_cloud_sharded_ptr(self, slice_host_ptr: int, n_total: int, stride: int = 16, intensity_offset: int = 12):
        """Multi-GPU hand-over (after nccl_init): this rank uploads only its slice of the map (`cloud_slice`), the slices are
        all-gathered over NVLink.  Collective: every rank calls it."""
        self._check(self._L.pcp_set_cloud_sharded(self._h, C.c_void_p(slice_host_ptr), n_total, stride, intensity_offset), "pcp_set_cloud_sharded")
        self.n = n_total

    def set_cloud_sharded(self, points: np.ndarray, rank: int, world: int, intensity_offset: int | None = None):
        """`points`: the WHOLE (N,k) float32 array; only this rank's slice of it is uploaded."""
        pts = np.ascontiguousarray(points, np.float32)
        stride = pts.shape[1] * 4
        if intensity_offset is None:
            intensity_offset = 12 if pts.shape[1] == 4 else (16 if pts.shape[1] >= 5 else 8)
        first, count = cloud_slice(pts.shape[0], rank, world)
        self._cloud_keepalive = pts
        self.set_cloud_sharded_ptr(pts.ctypes.data + first * stride, pts.shape[0], stride, intensity_offset)

    def submit_frame_ptr(self, pose7, bgr_ptr: int, pitch: int, mask_ptr: int = 0, mask_pitch: int = 0, T_opt=None):
        """submit_frame from raw (e.g. pinned) host pointers; asynchronous."""
        pose = np.ascontiguousarray(pose7, np.float64)
        T = self._mat16(T_opt)
        rc = self._L.pcp_submit_frame(self._h, _ptr(pose), _ptr(T), C.c_void_p(bgr_ptr), pitch, C.c_void_p(mask_ptr or 0), mask_pitch, None)
        self._check(rc, "pcp_submit_frame")

    def finalize_ptr(self, rgb_ptr: int, count_ptr: int = 0):
        self._check(self._L.pcp_finalize(self._h, C.c_void_p(rgb_ptr or 0), C.c_void_p(count_ptr or 0)), "pcp_finalize")

    # -- ViewCulling -----------------------------------------------------------------------------------
    def cull(self, w2c=None, T=None) -> np.ndarray:
        idx = np.empty(max(self.n, 1), np.int32)
        m = C.c_size_t(0)
        w = None if w2c is None else np.ascontiguousarray(np.asarray(w2c, np.float32).T.reshape(-1))
        rc = self._L.pcp_cull(self._h, _ptr(w), _ptr(self._mat16(T)), _ptr(idx), C.byref(m))
        self._check(rc, "pcp_cull")
        return idx[: m.value].copy()

    def cull_pose(self, pose7, T_opt=None) -> np.ndarray:
        pose = np.ascontiguousarray(pose7, np.float64)
        idx = np.empty(max(self.n, 1), np.int32)
        m = C.c_size_t(0)
        rc = self._L.pcp_cull_pose(self._h, _ptr(pose), _ptr(self._mat16(T_opt)), _ptr(idx), C.byref(m))
        self._check(rc, "pcp_cull_pose")
        return idx[: m.value].copy()

    def splat_masks_pose(self, pose7, T_opt=None, want_xyz=True, want_index=True):
        """generate_distance_masks (scripts/genNormAndDistanceMask.py:233-265) of the frame's culled cloud:
        -> (distance HxW f32, xyz HxWx3 f32 | None, index HxW i32 | None, pixels filled)."""
        pose = np.ascontiguousarray(pose7, np.float64)
        W, H = int(self.params.image_width), int(self.params.image_height)
        dist = np.empty((H, W), np.float32)
        xyz = np.empty((H, W, 3), np.float32) if want_xyz else None
        idx = np.empty((H, W), np.int32) if want_index else None
        m = C.c_size_t(0)
        rc = self._L.pcp_splat_masks_pose(self._h, _ptr(pose), _ptr(self._mat16(T_opt)), _ptr(dist), _ptr(xyz), _ptr(idx), C.byref(m))
        self._check(rc, "pcp_splat_masks_pose")
        return dist, xyz, idx, int(m.value)

    # -- accumulators / finalize -----------------------------------------------------------------------
    def accumulators(self):
        acc = np.empty((self.n, 4), np.uint32)
        mh = np.empty(self.n, np.uint32)
        mm = np.empty(self.n, np.uint32)
        self._check(self._L.pcp_get_accumulators(self._h, _ptr(acc), _ptr(mh), _ptr(mm)), "pcp_get_accumulators")
        return acc, mh, mm

    def mask_accumulators(self):
        """(hits of mask == 255, max mask value) per point; also available after exchange_finalize."""
        mh = np.empty(self.n, np.uint32)
        mm = np.empty(self.n, np.uint32)
        self._check(self._L.pcp_get_accumulators(self._h, None, _ptr(mh), _ptr(mm)), "pcp_get_accumulators")
        return mh, mm

    def accumulators_device(self) -> int:
        return int(self._L.pcp_accumulators_device(self._h) or 0)

    def reset_accumulators(self):
        self._check(self._L.pcp_reset_accumulators(self._h), "pcp_reset_accumulators")

    def finalize(self, want_count=True):
        rgb = np.empty((self.n, 3), np.uint8)
        cnt = np.empty(self.n, np.uint32) if want_count else None
        self._check(self._L.pcp_finalize(self._h, _ptr(rgb), _ptr(cnt)), "pcp_finalize")
        return (rgb, cnt) if want_count else rgb

    def finalize_compact(self) -> np.ndarray:
        """smoothColors + removePointsWithNoColor with the compaction on the device: structured array of the coloured points
        (x, y, z float32, b, g, r, a uint8 = pcl::PointXYZRGB's 16 bytes) in the caller's point order."""
        dt = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("b", "u1"), ("g", "u1"), ("r", "u1"), ("a", "u1")])
        m = C.c_size_t(0)
        self._check(self._L.pcp_finalize_compact(self._h, None, 0, C.byref(m)), "pcp_finalize_compact")
        out = np.empty(m.value, dt)
        if m.value:
            self._check(self._L.pcp_finalize_compact(self._h, _ptr(out), m.value, C.byref(m)), "pcp_finalize_compact")
        return out

    def finalize_device(self, d_rgb: int, d_count: int = 0):
        self._check(self._L.pcp_finalize_device(self._h, C.c_void_p(d_rgb or 0), C.c_void_p(d_count or 0)), "pcp_finalize_device")

    def profile_enable(self, on: bool = True):
        self._check(self._L.pcp_profile_enable(self._h, int(on)), "pcp_profile_enable")

    def profile_read(self):
        ms = (C.c_double * 3)()
        fr = C.c_uint64(0)
        self._check(self._L.pcp_profile_read(self._h, ms, C.byref(fr)), "pcp_profile_read")
        return {"init_ms": ms[0], "scan_ms": ms[1], "resolve_ms": ms[2], "frames": int(fr.value)}

    # -- multi-GPU -------------------------------------------------------------------------------------
    def nccl_unique_id(self) -> bytes:
        buf = (C.c_uint8 * 128)()
        self._check(self._L.pcp_nccl_unique_id(buf), "pcp_nccl_unique_id")
        return bytes(buf)

    def nccl_init(self, id128: bytes, rank: int, nranks: int):
        buf = (C.c_uint8 * 128).from_buffer_copy(id128)
        self._check(self._L.pcp_nccl_init(self._h, buf, rank, nranks), "pcp_nccl_init")

    def set_frame_numbering(self, next_index: int, stride: int = 1):
        """Global keyframe numbers of the frames submitted from now on (rank r of G, interleaved sharding: (r, G))."""
        self._check(self._L.pcp_set_frame_numbering(self._h, int(next_index), int(stride)), "pcp_set_frame_numbering")

    def allreduce(self):
        self._check(self._L.pcp_allreduce(self._h), "pcp_allreduce")

    def exchange_finalize(self, want_count=True):
        """Reduce-scatter of the sums + finalize of the rank's slice + all-gather of the colours: (rgb, count) of the whole job."""
        rgb = np.empty((self.n, 3), np.uint8)
        cnt = np.empty(self.n, np.uint32) if want_count else None
        self._check(self._L.pcp_exchange_finalize(self._h, _ptr(rgb), _ptr(cnt)), "pcp_exchange_finalize")
        return (rgb, cnt) if want_count else rgb

    def exchange_finalize_device(self, d_rgb: int, d_count: int = 0):
        self._check(self._L.pcp_exchange_finalize_device(self._h, C.c_void_p(d_rgb or 0), C.c_void_p(d_count or 0)), "pcp_exchange_finalize_device")

    def exchange_times(self):
        """ms of {reduce-scatter, slice finalize, all-gather, unpack} of the last exchange_finalize made with profiling enabled."""
        ms = (C.c_double * 4)()
        self._check(self._L.pcp_debug_exchange_times(self._h, ms), "pcp_debug_exchange_times")
        return {"reduce_scatter_ms": ms[0], "finalize_slice_ms": ms[1], "all_gather_ms": ms[2], "unpack_ms": ms[3]}

    def join(self):
        """The context's stream waits (on the device) for every frame submitted so far."""
        self._check(self._L.pcp_join(self._h), "pcp_join")


def load_pcd_xyzi(path) -> np.ndarray:
    """pcl::io::loadPCDFile<pcl::PointXYZI> (src/PointCloudProcessor.cpp:112): (N,4) float32 {x,y,z,intensity}."""
    L = _capi.load()
    ptr = C.POINTER(C.c_float)()
    n = C.c_size_t(0)
    rc = L.pcp_pcd_read_xyzi(str(path).encode(), C.byref(ptr), C.byref(n))
    if rc != 0:
        raise PcpError("Couldn't read point cloud file.", rc)
    try:
        return np.ctypeslib.as_array(ptr, shape=(n.value, 4)).copy() if n.value else np.zeros((0, 4), np.float32)
    finally:
        L.pcp_io_free(ptr)


def save_pcd_xyzi_ascii(path, xyzi: np.ndarray):
    """pcl::io::savePCDFileASCII (src/PointCloudProcessor.cpp:135)."""
    a = np.ascontiguousarray(xyzi, np.float32)
    rc = _capi.load().pcp_pcd_write_xyzi(str(path).encode(), _ptr(a), a.shape[0])
    if rc != 0:
        raise PcpError(f"cannot write {path}", rc)


def imread(path, grayscale=False) -> np.ndarray:
    """cv::imread(path[, IMREAD_GRAYSCALE]): HxWx3 B,G,R or HxW uint8 (JPEG through nvJPEG, PNG / PGM / PPM on the host)."""
    L = _capi.load()
    ptr = C.POINTER(C.c_uint8)()
    w, h, c = C.c_int(0), C.c_int(0), C.c_int(0)
    rc = L.pcp_image_read(str(path).encode(), int(grayscale), C.byref(ptr), C.byref(w), C.byref(h), C.byref(c))
    if rc != 0:
        raise PcpError(f"Failed to read image from: {path}", rc)
    try:
        shape = (h.value, w.value) if c.value == 1 else (h.value, w.value, c.value)
        return np.ctypeslib.as_array(ptr, shape=shape).copy()
    finally:
        L.pcp_io_free(ptr)


class PointCloudProcessor:
    """PointCloudProcessor(pointCloudPath, odometryPath, imagesFolder, maskImageFolder, outputPath, enableMLS, enableNIDOptimize,
    enableInitialGuessManual).process()   (include/PointCloudProcessor.hpp:17-27, src/PointCloudProcessor.cpp:1007-1032)."""

    def __init__(self, pointCloudPath, odometryPath, imagesFolder, maskImageFolder="", outputPath=".", enableMLS=False, enableNIDOptimize=False,
                 enableInitialGuessManual=False, camera: Camera | None = None, cull_size=None, T_camera_lidar_optimized=None, device=0,
                 skip_outputs=False, sample_mode=SAMPLE_NEAREST, binary_output=False, nid_grey_plane=False):
        self._keep = [str(pointCloudPath).encode(), str(odometryPath).encode(), str(imagesFolder).encode(), str(maskImageFolder or "").encode(),
                      str(outputPath).encode()]
        a = _capi.ProcessArgs()
        a.point_cloud_path, a.odometry_path, a.images_folder, a.mask_image_folder, a.output_path = self._keep
        a.enable_mls, a.enable_nid_optimize, a.enable_initial_guess_manual = int(enableMLS), int(enableNIDOptimize), int(enableInitialGuessManual)
        a.device = device
        self._cam = camera
        if camera is not None:
            a.camera = C.pointer(camera)
        if cull_size is not None:
            a.cull_width, a.cull_height = int(cull_size[0]), int(cull_size[1])
        self._T = None
        if T_camera_lidar_optimized is not None:
            self._T = np.ascontiguousarray(np.asarray(T_camera_lidar_optimized, np.float64).T.reshape(-1))
            a.T_camera_lidar_optimized = self._T.ctypes.data_as(C.POINTER(C.c_double))
        a.skip_outputs = int(skip_outputs)
        a.sample_mode = int(sample_mode)
        a.nid_grey_plane = int(nid_grey_plane)
        a.binary_output = int(binary_output)  # cloudInWorldWithRGB.pcd as DATA binary instead of the reference's ASCII
        self._args = a
        self.stats = None

    def process(self) -> dict:
        st = _capi.ProcessStats()
        err = C.create_string_buffer(1024)
        rc = _capi.load().pcp_process(C.byref(self._args), C.byref(st), err, len(err))
        self.stats = {k: getattr(st, k) for k, _ in st._fields_}
        if rc != 0:
            raise PcpError(err.value.decode(errors="replace") or f"pcp_process failed ({rc})", rc)
        return self.stats


def pose_matrices(pose7, T_opt=None):
    """(w2c, c2w) float32 4x4 as derived at src/PointCloudProcessor.cpp:495-519 (host helper of the library)."""
    pose = np.ascontiguousarray(pose7, np.float64)
    T = Colorizer._mat16(T_opt)
    w = np.zeros(16, np.float32)
    c = np.zeros(16, np.float32)
    rc = _capi.load().pcp_pose_matrices(_ptr(pose), _ptr(T), _ptr(w), _ptr(c))
    if rc != _capi.PCP_OK:
        raise PcpError("pcp_pose_matrices failed")
    return w.reshape(4, 4).T.copy(), c.reshape(4, 4).T.copy()


class ViewCulling:
    """vlcal::ViewCulling(proj, image_size, params).cull(points, T_camera_lidar) — z-buffer variant
    (src/vlcal/calib/view_culling.cpp:52-174).  Returns the culled cloud like the reference (view_culling.cpp:176-185)."""

    def __init__(self, proj: Camera, image_size, enable_depth_buffer_culling: bool = True, device: int = 0):
        p = default_params(image_size=image_size)
        p.enable_depth_buffer_culling = int(enable_depth_buffer_culling)
        self._c = Colorizer(proj, p, device)

    def cull(self, points: np.ndarray, T_camera_lidar=None, return_indices=False):
        self._c.set_cloud(points)
        idx = self._c.cull(None, T_camera_lidar)
        return idx if return_indices else np.ascontiguousarray(points)[idx]

    def close(self):
        self._c.close()


class NIDCost:
    """vlcal::NIDCost(proj, normalized_image, points, bins) (include/vlcal/costs/nid_cost.hpp:26-40)."""

    def __init__(self, ctx: Colorizer, normalized_image: np.ndarray, points: np.ndarray, bins: int = 16):
        self._ctx = ctx
        self._L = ctx._L
        pts = np.ascontiguousarray(points, np.float32)
        ioff = 12 if pts.shape[1] == 4 else 16
        self._h = C.c_void_p()
        if normalized_image.dtype == np.uint8:
            # the 8-bit image itself (pcp_nid_create_u8): HxW grey plane, or HxWx3 B,G,R read the way NIDCost reads it (the first W
            # interleaved values of every row, nid_cost.hpp:61, :88); converted to [0,1] and binned on the device
            img = np.ascontiguousarray(normalized_image)
            if img.ndim not in (2, 3):
                raise PcpError("NIDCost: uint8 image must be HxW or HxWx3")
            rc = self._L.pcp_nid_create_u8(ctx.handle, _ptr(pts), pts.shape[0], pts.shape[1] * 4, ioff, _ptr(img), img.strides[0], img.shape[1],
                                           img.shape[0], int(bins), C.byref(self._h))
            ctx._check(rc, "pcp_nid_create_u8")
            return
        img = np.ascontiguousarray(normalized_image, np.float64)
        if img.ndim != 2:
            raise PcpError("NIDCost: normalized_image must be a single-channel HxW float64 plane")
        rc = self._L.pcp_nid_create(ctx.handle, _ptr(pts), pts.shape[0], pts.shape[1] * 4, ioff, _ptr(img), img.shape[1], img.shape[0],
                                    int(bins), C.byref(self._h))
        ctx._check(rc, "pcp_nid_create")

    def __call__(self, se3_params):
        """NIDCost::operator()<double>: se3 (7,) or (B,7) in Sophus order -> (nid, ok)."""
        se3 = np.ascontiguousarray(np.atleast_2d(se3_params), np.float64)
        B = se3.shape[0]
        out = np.empty(B, np.float64)
        ok = np.empty(B, np.uint8)
        self._ctx._check(self._L.pcp_nid_eval(self._h, _ptr(se3), B, _ptr(out), _ptr(ok)), "pcp_nid_eval")
        return out, ok.astype(bool)

    def evaluate_with_gradient(self, se3_params):
        """NIDCost::operator()<ceres::Jet<double,7>>: -> (nid (B,), grad (B,7) over {qx,qy,qz,qw,tx,ty,tz}, ok (B,))."""
        se3 = np.ascontiguousarray(np.atleast_2d(se3_params), np.float64)
        B = se3.shape[0]
        out = np.empty(B, np.float64)
        grad = np.empty((B, 7), np.float64)
        ok = np.empty(B, np.uint8)
        self._ctx._check(self._L.pcp_nid_eval_grad(self._h, _ptr(se3), B, _ptr(out), _ptr(grad), _ptr(ok)), "pcp_nid_eval_grad")
        return out, grad, ok.astype(bool)

    def close(self):
        if self._h:
            self._L.pcp_nid_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def multi_nid_with_gradient(costs, init_se3, se3_params):
    """MultiNIDCost::operator()<ceres::Jet<double,7>> (visual_camera_calibration.cpp:93-124, :204-233): -> (sum, grad, ok)."""
    L = _capi.load()
    se3 = np.ascontiguousarray(np.atleast_2d(se3_params), np.float64)
    init = np.ascontiguousarray(init_se3, np.float64)
    B = se3.shape[0]
    arr = (C.c_void_p * len(costs))(*[c._h for c in costs])
    out = np.empty(B, np.float64)
    grad = np.empty((B, 7), np.float64)
    ok = np.empty(B, np.uint8)
    rc = L.pcp_nid_eval_multi_grad(arr, len(costs), _ptr(init), _ptr(se3), B, _ptr(out), _ptr(grad), _ptr(ok))
    costs[0]._ctx._check(rc, "pcp_nid_eval_multi_grad")
    return out, grad, ok.astype(bool)


def bfgs_options(**kw) -> _capi.BfgsOptions:
    """ceres::GradientProblemSolver::Options at their defaults with line_search_direction_type = BFGS."""
    o = _capi.BfgsOptions()
    _capi.load().pcp_bfgs_default_options(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise AttributeError(k)
        setattr(o, k, v)
    return o


def _summary_dict(st):
    d = {k: getattr(st, k) for k, _ in st._fields_}
    if "message" in d:
        d["message"] = d["message"].decode(errors="replace")
    return d


def se3_bfgs(fn, init_se3, options: _capi.BfgsOptions | None = None):
    """ceres::Solve over Sophus::Manifold<SE3> for a host first-order function: fn(se3[7]) -> (value, grad7, ok).
    Returns (se3, summary)."""
    def cb(_user, x, value, grad):
        v, g, ok = fn(np.array([x[i] for i in range(7)], np.float64))
        value[0] = float(v)
        for i in range(7):
            grad[i] = float(g[i])
        return 1 if ok else 0
    cfn = _capi.SE3_COST_FN(cb)
    init = np.ascontiguousarray(init_se3, np.float64)
    out = np.empty(7, np.float64)
    st = _capi.BfgsSummary()
    rc = _capi.load().pcp_se3_bfgs(cfn, None, _ptr(init), C.byref(options) if options is not None else None, _ptr(out), C.byref(st))
    if rc != _capi.PCP_OK:
        raise PcpError(f"pcp_se3_bfgs failed ({rc})", rc)
    return out, _summary_dict(st)


def calibrate(costs, init_T_camera_lidar=None, max_outer_iterations=None, delta_trans_thresh=None, delta_rot_thresh=None, **inner):
    """vlcal::VisualCameraCalibration(proj, keyframes, params).calibrate(init_T_camera_lidar)
    (src/vlcal/calib/visual_camera_calibration.cpp:49-84, :142-252) over NIDCost objects.  Returns (T 4x4, summary)."""
    L = _capi.load()
    p = _capi.CalibParams()
    L.pcp_calib_default_params(C.byref(p))
    if max_outer_iterations is not None:
        p.max_outer_iterations = int(max_outer_iterations)
    if delta_trans_thresh is not None:
        p.delta_trans_thresh = float(delta_trans_thresh)
    if delta_rot_thresh is not None:
        p.delta_rot_thresh = float(delta_rot_thresh)
    for k, v in inner.items():
        if not hasattr(p.inner, k):
            raise AttributeError(k)
        setattr(p.inner, k, v)
    T0 = np.eye(4) if init_T_camera_lidar is None else np.asarray(init_T_camera_lidar, np.float64)
    t_in = np.ascontiguousarray(T0.T.reshape(-1))
    t_out = np.empty(16, np.float64)
    st = _capi.CalibSummary()
    arr = (C.c_void_p * len(costs))(*[c._h for c in costs])
    rc = L.pcp_calibrate(arr, len(costs), _ptr(t_in), C.byref(p), _ptr(t_out), C.byref(st))
    costs[0]._ctx._check(rc, "pcp_calibrate")
    return t_out.reshape(4, 4).T.copy(), _summary_dict(st)


def multi_nid(costs, init_se3, se3_params):
    """MultiNIDCost::operator()<double> (src/vlcal/calib/visual_camera_calibration.cpp:93-124)."""
    L = _capi.load()
    se3 = np.ascontiguousarray(np.atleast_2d(se3_params), np.float64)
    init = np.ascontiguousarray(init_se3, np.float64)
    B = se3.shape[0]
    arr = (C.c_void_p * len(costs))(*[c._h for c in costs])
    out = np.empty(B, np.float64)
    ok = np.empty(B, np.uint8)
    rc = L.pcp_nid_eval_multi(arr, len(costs), _ptr(init), _ptr(se3), B, _ptr(out), _ptr(ok))
    costs[0]._ctx._check(rc, "pcp_nid_eval_multi")
    return out, ok.astype(bool)


def host_alloc(nbytes: int) -> int:
    """Pinned host memory (pcp_host_alloc); returns the address."""
    p = C.c_void_p()
    if _capi.load().pcp_host_alloc(C.byref(p), nbytes) != _capi.PCP_OK:
        raise PcpError("pcp_host_alloc failed")
    return int(p.value)


def host_free(ptr: int):
    _capi.load().pcp_host_free(C.c_void_p(ptr))


def host_array(ptr: int, shape, dtype):
    """numpy view of a raw host allocation."""
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    buf = (C.c_uint8 * n).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).reshape(shape)

"""ctypes binding of libpcp_b200.so (include/pcp/pcp.h).  No torch types cross this boundary.

The library is never rebuilt at import time and there is no CPU fallback: a missing library or a missing
CUDA device raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PCP_LIB_PATH") or os.path.join(_HERE, "libpcp_b200.so")  # PCP_LIB_PATH: kernel-variant experiments

PCP_OK, PCP_EINVAL, PCP_EIO, PCP_ECUDA, PCP_ENCCL, PCP_ENOMEM = range(6)
ACCUM_SUM, ACCUM_TOP5 = 0, 1
SAMPLE_NEAREST, SAMPLE_BILINEAR = 0, 1


class Camera(C.Structure):
    _fields_ = [("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double), ("dist", C.c_double * 5)]


class Params(C.Structure):
    _fields_ = [
        ("image_width", C.c_int), ("image_height", C.c_int), ("cull_width", C.c_int), ("cull_height", C.c_int),
        ("downsample", C.c_int), ("depth_tolerance", C.c_double), ("enable_depth_buffer_culling", C.c_int),
        ("accum_mode", C.c_int), ("sample_mode", C.c_int), ("top_m", C.c_int), ("prebin", C.c_int), ("frames_in_flight", C.c_int), ("hsv_adjust", C.c_int), ("saturation_scale", C.c_float),
        ("brightness_scale", C.c_float), ("hsv_vector_block", C.c_int), ("debug_flags", C.c_int), ("frames_per_launch", C.c_int),
    ]


class FrameOut(C.Structure):
    _fields_ = [("visible_bits", C.c_void_p), ("pixel_index", C.c_void_p), ("n_candidates", C.c_int64), ("n_visible", C.c_int64),
                ("n_colored", C.c_int64), ("flags", C.c_uint32), ("reserved", C.c_uint32)]


OUT_SKIP_CANDIDATES = 1
OUT_SKIP_CANDIDATES, OUT_PIXELS_ON_DEVICE = 1, 2
DEBUG_NO_PREFILTER, DEBUG_NO_SEED, DEBUG_FINALIZE_ONEPASS, DEBUG_FINALIZE_TWOPASS = 1, 2, 4, 8


# every symbol include/pcp/pcp.h declares
class ProcessArgs(C.Structure):
    _fields_ = [
        ("point_cloud_path", C.c_char_p), ("odometry_path", C.c_char_p), ("images_folder", C.c_char_p), ("mask_image_folder", C.c_char_p),
        ("output_path", C.c_char_p), ("enable_mls", C.c_int), ("enable_nid_optimize", C.c_int), ("enable_initial_guess_manual", C.c_int),
        ("device", C.c_int), ("camera", C.POINTER(Camera)), ("cull_width", C.c_int), ("cull_height", C.c_int),
        ("T_camera_lidar_optimized", C.POINTER(C.c_double)), ("skip_outputs", C.c_int), ("sample_mode", C.c_int), ("nid_grey_plane", C.c_int), ("binary_output", C.c_int), ("reserved", C.c_int * 7),
    ]


class ProcessStats(C.Structure):
    _fields_ = [
        ("n_points", C.c_int64), ("n_cropped", C.c_int64), ("n_frames", C.c_int64), ("n_keyframes", C.c_int64), ("n_colored_points", C.c_int64),
        ("n_mask_points", C.c_int64), ("n_visible_total", C.c_int64), ("seconds_load", C.c_double), ("seconds_cull_pass", C.c_double),
        ("seconds_color_pass", C.c_double), ("seconds_write", C.c_double), ("seconds_total", C.c_double), ("seconds_nid", C.c_double),
        ("nid_outer_iterations", C.c_int64), ("nid_inner_iterations", C.c_int64), ("nid_cost_evaluations", C.c_int64), ("nid_converged", C.c_int64),
        ("nid_initial_cost", C.c_double), ("nid_final_cost", C.c_double), ("T_camera_lidar", C.c_double * 16),
    ]


SYMBOLS = [
    "pcp_default_params", "pcp_abi_version", "pcp_create", "pcp_destroy", "pcp_last_error", "pcp_set_stream", "pcp_get_stream",
    "pcp_synchronize", "pcp_host_alloc", "pcp_host_free", "pcp_frame_mask_cloud", "pcp_set_cloud", "pcp_set_cloud_sharded", "pcp_cloud_slice", "pcp_set_cloud_device", "pcp_num_points",
    "pcp_submit_frame", "pcp_submit_frame_device", "pcp_cull", "pcp_cull_pose", "pcp_pose_matrices", "pcp_get_accumulators",
    "pcp_accumulators_device", "pcp_reset_accumulators", "pcp_nccl_unique_id", "pcp_nccl_init", "pcp_allreduce", "pcp_finalize", "pcp_finalize_device", "pcp_profile_enable", "pcp_profile_read",
    "pcp_nid_create", "pcp_nid_destroy", "pcp_nid_eval", "pcp_nid_eval_multi", "pcp_nid_eval_grad", "pcp_nid_eval_multi_grad", "pcp_launch_count", "pcp_debug_counters", "pcp_hsv_roundtrip", "pcp_process", "pcp_pcd_read_xyzi", "pcp_pcd_write_xyzi", "pcp_image_read", "pcp_io_free",
    "pcp_bfgs_default_options", "pcp_se3_bfgs", "pcp_calib_default_params", "pcp_calibrate", "pcp_set_frame_numbering", "pcp_splat_masks_pose",
    "pcp_join", "pcp_exchange_finalize", "pcp_exchange_finalize_device", "pcp_finalize_compact", "pcp_pcd_write_xyzrgb", "pcp_submit_frames_device", "pcp_compact_bits_noalloc", "pcp_nid_create_u8", "pcp_debug_exchange_times",
]
class BfgsOptions(C.Structure):
    _fields_ = [
        ("max_num_iterations", C.c_int), ("function_tolerance", C.c_double), ("gradient_tolerance", C.c_double), ("parameter_tolerance", C.c_double),
        ("sufficient_function_decrease", C.c_double), ("sufficient_curvature_decrease", C.c_double), ("max_line_search_step_expansion", C.c_double),
        ("min_line_search_step_contraction", C.c_double), ("max_line_search_step_contraction", C.c_double), ("min_line_search_step_size", C.c_double),
        ("max_num_line_search_step_size_iterations", C.c_int), ("max_num_line_search_direction_restarts", C.c_int), ("verbose", C.c_int),
        ("reserved", C.c_int * 5),
    ]


class BfgsSummary(C.Structure):
    _fields_ = [("iterations", C.c_int), ("termination", C.c_int), ("num_cost_evaluations", C.c_int), ("num_line_search_steps", C.c_int),
                ("initial_cost", C.c_double), ("final_cost", C.c_double), ("message", C.c_char * 96)]


class CalibParams(C.Structure):
    _fields_ = [("max_outer_iterations", C.c_int), ("delta_trans_thresh", C.c_double), ("delta_rot_thresh", C.c_double), ("inner", BfgsOptions)]


class CalibSummary(C.Structure):
    _fields_ = [("outer_iterations", C.c_int), ("converged", C.c_int), ("inner_iterations", C.c_int), ("cost_evaluations", C.c_int),
                ("last_inner_termination", C.c_int), ("initial_cost", C.c_double), ("final_cost", C.c_double), ("delta_t", C.c_double),
                ("delta_r", C.c_double)]


SE3_COST_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double))

_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(the CUDA library is the only implementation; there is no CPU fallback)")
    L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    vp, sz, i32, dbl = C.c_void_p, C.c_size_t, C.c_int, C.c_double
    L.pcp_default_params.argtypes = [C.POINTER(Params)]
    L.pcp_default_params.restype = None
    L.pcp_abi_version.restype = i32
    L.pcp_create.argtypes = [C.POINTER(vp), i32, C.POINTER(Camera), C.POINTER(Params)]
    L.pcp_destroy.argtypes = [vp]
    L.pcp_destroy.restype = None
    L.pcp_last_error.argtypes = [vp]
    L.pcp_last_error.restype = C.c_char_p
    L.pcp_set_stream.argtypes = [vp, vp]
    L.pcp_get_stream.argtypes = [vp]
    L.pcp_get_stream.restype = vp
    L.pcp_synchronize.argtypes = [vp]
    L.pcp_join.argtypes = [vp]
    L.pcp_compact_bits_noalloc.argtypes = [vp, vp, sz, vp, vp]
    L.pcp_submit_frames_device.argtypes = [vp, sz, vp, vp, vp, sz, vp, sz]
    L.pcp_finalize_compact.argtypes = [vp, vp, sz, C.POINTER(sz)]
    L.pcp_pcd_write_xyzrgb.argtypes = [C.c_char_p, vp, sz, i32]
    L.pcp_exchange_finalize.argtypes = [vp, vp, vp]
    L.pcp_exchange_finalize_device.argtypes = [vp, vp, vp]
    L.pcp_host_alloc.argtypes = [C.POINTER(vp), sz]
    L.pcp_host_free.argtypes = [vp]
    L.pcp_set_cloud.argtypes = [vp, vp, sz, sz, sz]
    L.pcp_set_cloud_device.argtypes = [vp, vp, sz]
    L.pcp_frame_mask_cloud.argtypes = [vp, vp, sz, C.POINTER(sz)]
    L.pcp_set_cloud_sharded.argtypes = [vp, vp, sz, sz, sz]
    L.pcp_cloud_slice.argtypes = [sz, C.c_int, C.c_int, C.POINTER(sz), C.POINTER(sz)]
    L.pcp_num_points.argtypes = [vp]
    L.pcp_num_points.restype = sz
    L.pcp_submit_frame.argtypes = [vp, vp, vp, vp, sz, vp, sz, C.POINTER(FrameOut)]
    L.pcp_submit_frame_device.argtypes = [vp, vp, vp, vp, sz, vp, sz, C.POINTER(FrameOut)]
    L.pcp_cull.argtypes = [vp, vp, vp, vp, C.POINTER(sz)]
    L.pcp_cull_pose.argtypes = [vp, vp, vp, vp, C.POINTER(sz)]
    L.pcp_pose_matrices.argtypes = [vp, vp, vp, vp]
    L.pcp_splat_masks_pose.argtypes = [vp, vp, vp, vp, vp, vp, C.POINTER(sz)]
    L.pcp_get_accumulators.argtypes = [vp, vp, vp, vp]
    L.pcp_accumulators_device.argtypes = [vp]
    L.pcp_accumulators_device.restype = vp
    L.pcp_reset_accumulators.argtypes = [vp]
    L.pcp_nccl_unique_id.argtypes = [vp]
    L.pcp_nccl_init.argtypes = [vp, vp, i32, i32]
    L.pcp_allreduce.argtypes = [vp]
    L.pcp_set_frame_numbering.argtypes = [vp, i32, i32]
    L.pcp_finalize.argtypes = [vp, vp, vp]
    L.pcp_finalize_device.argtypes = [vp, vp, vp]
    L.pcp_profile_enable.argtypes = [vp, i32]
    L.pcp_profile_read.argtypes = [vp, vp, vp]
    L.pcp_nid_create.argtypes = [vp, vp, sz, sz, sz, vp, i32, i32, i32, C.POINTER(vp)]
    L.pcp_nid_create_u8.argtypes = [vp, vp, sz, sz, sz, vp, sz, i32, i32, i32, C.POINTER(vp)]
    L.pcp_nid_destroy.argtypes = [vp]
    L.pcp_nid_destroy.restype = None
    L.pcp_nid_eval.argtypes = [vp, vp, i32, vp, vp]
    L.pcp_nid_eval_multi.argtypes = [C.POINTER(vp), i32, vp, vp, i32, vp, vp]
    L.pcp_debug_counters.argtypes = [vp, vp]
    L.pcp_pcd_read_xyzi.argtypes = [C.c_char_p, C.POINTER(C.POINTER(C.c_float)), C.POINTER(sz)]
    L.pcp_pcd_write_xyzi.argtypes = [C.c_char_p, vp, sz]
    L.pcp_image_read.argtypes = [C.c_char_p, i32, C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]
    L.pcp_io_free.argtypes = [vp]
    L.pcp_io_free.restype = None
    L.pcp_process.argtypes = [C.POINTER(ProcessArgs), C.POINTER(ProcessStats), C.c_char_p, sz]
    L.pcp_hsv_roundtrip.argtypes = [vp, vp, sz, vp, sz, i32, i32, C.c_float, C.c_float]
    L.pcp_nid_eval_grad.argtypes = [vp, vp, i32, vp, vp, vp]
    L.pcp_nid_eval_multi_grad.argtypes = [C.POINTER(vp), i32, vp, vp, i32, vp, vp, vp]
    L.pcp_bfgs_default_options.argtypes = [C.POINTER(BfgsOptions)]
    L.pcp_bfgs_default_options.restype = None
    L.pcp_se3_bfgs.argtypes = [SE3_COST_FN, vp, vp, C.POINTER(BfgsOptions), vp, C.POINTER(BfgsSummary)]
    L.pcp_calib_default_params.argtypes = [C.POINTER(CalibParams)]
    L.pcp_calib_default_params.restype = None
    L.pcp_calibrate.argtypes = [C.POINTER(vp), i32, vp, C.POINTER(CalibParams), vp, C.POINTER(CalibSummary)]
    L.pcp_launch_count.argtypes = [vp]
    L.pcp_launch_count.restype = C.c_uint64
    _lib = L
    return L

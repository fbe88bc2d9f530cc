"""ctypes binding of the C ABI (include/acmmp_b200.h).  No compute happens here."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ACMMP_B200_LIB") or os.path.join(_HERE, "libacmmp_b200.so")   # override: kernel-variant experiments

MAX_SRC_VIEWS = 32
COMPAT_REF_SM100 = 0
COMPAT_UPSTREAM = 1
TEX_R32F = 0   # acmmp_tex_format
TEX_AUTO = 1


class Camera(C.Structure):
    """= reference `struct Camera` (acmmp_definitions.h:47-55), 100 bytes."""
    _fields_ = [("K", C.c_float * 9), ("R", C.c_float * 9), ("t", C.c_float * 3),
                ("height", C.c_int32), ("width", C.c_int32),
                ("depth_min", C.c_float), ("depth_max", C.c_float)]

    @staticmethod
    def make(K, R, t, width, height, depth_min, depth_max):
        c = Camera()
        c.K[:] = [float(x) for x in np.asarray(K, np.float32).reshape(9)]
        c.R[:] = [float(x) for x in np.asarray(R, np.float32).reshape(9)]
        c.t[:] = [float(x) for x in np.asarray(t, np.float32).reshape(3)]
        c.width, c.height = int(width), int(height)
        c.depth_min, c.depth_max = float(depth_min), float(depth_max)
        return c

    def copy(self):
        c = Camera()
        C.memmove(C.byref(c), C.byref(self), C.sizeof(Camera))
        return c


class Params(C.Structure):
    """= reference `struct PatchMatchParams` (ACMMP.h:32-56), flags as int32."""
    _fields_ = [("max_iterations", C.c_int32), ("patch_size", C.c_int32), ("num_images", C.c_int32),
                ("max_image_size", C.c_int32), ("radius_increment", C.c_int32),
                ("sigma_spatial", C.c_float), ("sigma_color", C.c_float), ("top_k", C.c_int32),
                ("baseline", C.c_float), ("depth_min", C.c_float), ("depth_max", C.c_float),
                ("disparity_min", C.c_float), ("disparity_max", C.c_float),
                ("scaled_cols", C.c_float), ("scaled_rows", C.c_float),
                ("geom_consistency", C.c_int32), ("planar_prior", C.c_int32), ("multi_geometry", C.c_int32),
                ("hierarchy", C.c_int32), ("upsample", C.c_int32), ("seeded", C.c_int32)]


assert C.sizeof(Camera) == 100

EXPORTS = [
    "acmmp_create", "acmmp_destroy", "acmmp_last_error", "acmmp_version", "acmmp_default_params",
    "acmmp_set_geom_consistency", "acmmp_set_hierarchy", "acmmp_set_planar_prior_params",
    "acmmp_set_compat_mode", "acmmp_set_wave_emulation", "acmmp_set_texture_format", "acmmp_texture_is_u8", "acmmp_get_params", "acmmp_set_problem", "acmmp_set_depth_map",
    "acmmp_set_warm_start", "acmmp_set_hierarchy_inputs", "acmmp_set_seeded_planes",
    "acmmp_set_planar_prior", "acmmp_build_planar_prior", "acmmp_run", "acmmp_run_init",
    "acmmp_run_sweep", "acmmp_run_sweep_snapshot", "acmmp_run_tail", "acmmp_get_result", "acmmp_get_maps",
    "acmmp_get_depth_device", "acmmp_get_state", "acmmp_set_state", "acmmp_width", "acmmp_height",
    "acmmp_sync", "acmmp_launch_count", "acmmp_last_timing", "acmmp_jbu_depth", "acmmp_resize_linear",
    "acmmp_rng_init_states", "acmmp_measure_fp32_peak", "acmmp_measure_tex_peak", "acmmp_measure_tex_peak_u8", "acmmp_tex_sample", "acmmp_tex_sample_u8", "acmmp_get_planar_prior", "acmmp_delaunay", "acmmp_set_delaunay_threads", "acmmp_prior_raster_noalloc", "acmmp_timing_collect",
    "acmmp_count_ncc", "acmmp_ncc_count", "acmmp_scan_count_ncc", "acmmp_scan_ncc_count",
    "acmmp_fusion_create", "acmmp_fusion_destroy", "acmmp_fusion_last_error", "acmmp_fusion_set_view", "acmmp_fusion_run",
    "acmmp_fusion_get_points", "acmmp_fusion_get_mask", "acmmp_fusion_stats",
    "acmmp_fusion_set_view_second", "acmmp_fusion_run_prior_aware",
    "acmmp_jpeg_info", "acmmp_decode_jpeg", "acmmp_jpeg_last_error",
    "acmmp_pair_scores", "acmmp_pairs_last_error",
    "acmmp_scan_set_prior_maps", "acmmp_prior_planes_from_maps", "acmmp_seeded_last_error",
    "acmmp_scan_create", "acmmp_scan_destroy", "acmmp_scan_last_error", "acmmp_scan_set_image", "acmmp_scan_set_image_u8",
    "acmmp_scan_set_shard", "acmmp_scan_set_texture_format", "acmmp_scan_level_is_u8", "acmmp_scan_set_sequential_geom", "acmmp_scan_set_schedule", "acmmp_nccl_version", "acmmp_nccl_get_unique_id", "acmmp_scan_comm_init",
    "acmmp_scan_comm_init_all", "acmmp_scan_comm_destroy", "acmmp_scan_all_gather", "acmmp_scan_collect_view", "acmmp_scan_run_dataflow", "acmmp_scan_dataflow_stats", "acmmp_scan_dataflow_trace", "acmmp_plan_dataflow", "acmmp_scan_keep_photometric", "acmmp_set_box_staging", "acmmp_scan_set_box_staging", "acmmp_scan_level_stats", "acmmp_scan_get_photometric_depth", "acmmp_scan_num_levels", "acmmp_scan_begin_level",
    "acmmp_scan_run_stage", "acmmp_scan_exchange", "acmmp_scan_view_size", "acmmp_scan_get_view", "acmmp_scan_view_device",
    "acmmp_scan_timing", "acmmp_scan_sync", "acmmp_scan_reset", "acmmp_scan_host_stats", "acmmp_scan_mark", "acmmp_scan_elapsed",
]

_lib = None


class AcmmpError(RuntimeError):
    pass


def load():
    """dlopen the CUDA library; raises if it was not built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise AcmmpError(f"{LIB_PATH} not built: run `make` or __graft_entry__.build() (no CPU fallback exists)")
    lib = C.CDLL(LIB_PATH)
    vp, fp, up = C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_uint32)
    lib.acmmp_create.argtypes = [C.POINTER(vp), C.c_int]
    lib.acmmp_destroy.argtypes = [vp]
    lib.acmmp_destroy.restype = None
    lib.acmmp_last_error.argtypes = [vp]
    lib.acmmp_last_error.restype = C.c_char_p
    lib.acmmp_version.restype = C.c_char_p
    lib.acmmp_default_params.argtypes = [C.POINTER(Params)]
    lib.acmmp_default_params.restype = None
    lib.acmmp_set_geom_consistency.argtypes = [vp, C.c_int]
    lib.acmmp_set_hierarchy.argtypes = [vp]
    lib.acmmp_set_planar_prior_params.argtypes = [vp]
    lib.acmmp_set_compat_mode.argtypes = [vp, C.c_int]
    lib.acmmp_set_wave_emulation.argtypes = [vp, C.c_int]
    lib.acmmp_set_texture_format.argtypes = [vp, C.c_int]
    lib.acmmp_texture_is_u8.argtypes = [vp]
    lib.acmmp_get_params.argtypes = [vp, C.POINTER(Params)]
    lib.acmmp_set_problem.argtypes = [vp, C.c_int, C.POINTER(Camera), C.POINTER(vp), C.c_int]
    lib.acmmp_set_depth_map.argtypes = [vp, C.c_int, vp, C.c_int, C.c_int, C.c_int]
    lib.acmmp_set_warm_start.argtypes = [vp, fp, fp, fp]
    lib.acmmp_set_hierarchy_inputs.argtypes = [vp, fp, fp, C.c_int, C.c_int, fp]
    lib.acmmp_set_seeded_planes.argtypes = [vp, fp]
    lib.acmmp_set_planar_prior.argtypes = [vp, fp, up]
    lib.acmmp_build_planar_prior.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.acmmp_get_planar_prior.argtypes = [vp, fp, up]
    lib.acmmp_delaunay.argtypes = [C.POINTER(C.c_int32), C.c_int, C.POINTER(C.c_int32), C.c_int]
    lib.acmmp_set_delaunay_threads.argtypes = [C.c_int]
    lib.acmmp_prior_raster_noalloc.argtypes = [C.c_int, C.POINTER(Camera), C.c_float, C.c_float, vp, vp, C.c_int, vp, vp, vp, C.c_char_p, C.c_size_t]
    lib.acmmp_run.argtypes = [vp, C.c_uint64]
    lib.acmmp_run_init.argtypes = [vp, C.c_uint64]
    lib.acmmp_run_sweep.argtypes = [vp, C.c_int, C.c_int]
    lib.acmmp_run_sweep_snapshot.argtypes = [vp, C.c_int, C.c_int, fp, fp]
    lib.acmmp_run_tail.argtypes = [vp]
    lib.acmmp_get_result.argtypes = [vp, fp, fp, up]
    lib.acmmp_get_maps.argtypes = [vp, fp, fp, fp]
    lib.acmmp_get_depth_device.argtypes = [vp, C.POINTER(vp)]
    lib.acmmp_get_state.argtypes = [vp, fp, fp, up, up]
    lib.acmmp_set_state.argtypes = [vp, fp, fp, up, up]
    lib.acmmp_width.argtypes = [vp]
    lib.acmmp_height.argtypes = [vp]
    lib.acmmp_sync.argtypes = [vp]
    lib.acmmp_launch_count.argtypes = [vp]
    lib.acmmp_launch_count.restype = C.c_uint64
    lib.acmmp_last_timing.argtypes = [vp, C.POINTER(C.c_float * 4)]
    lib.acmmp_timing_collect.argtypes = [vp, C.POINTER(C.c_double * 4), C.POINTER(C.c_int)]
    lib.acmmp_jbu_depth.argtypes = [C.c_int, fp, C.c_int, C.c_int, fp, C.c_int, C.c_int, fp]
    lib.acmmp_resize_linear.argtypes = [C.c_int, fp, C.c_int, C.c_int, fp, C.c_int, C.c_int]
    lib.acmmp_rng_init_states.argtypes = [C.c_int, C.c_uint64, C.c_int, C.c_int, up]
    lib.acmmp_measure_fp32_peak.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.acmmp_measure_tex_peak.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.acmmp_measure_tex_peak_u8.argtypes = [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.acmmp_tex_sample.argtypes = [C.c_int, fp, C.c_int, C.c_int, fp, fp, C.c_int, fp]
    lib.acmmp_tex_sample_u8.argtypes = [C.c_int, fp, C.c_int, C.c_int, fp, fp, C.c_int, fp]
    i32p = C.POINTER(C.c_int32)
    lib.acmmp_scan_create.argtypes = [C.POINTER(vp), C.c_int, C.c_int, C.POINTER(Camera), i32p, i32p]
    lib.acmmp_scan_destroy.argtypes = [vp]
    lib.acmmp_scan_destroy.restype = None
    lib.acmmp_scan_last_error.argtypes = [vp]
    lib.acmmp_scan_last_error.restype = C.c_char_p
    lib.acmmp_scan_set_image.argtypes = [vp, C.c_int, vp, C.c_int]
    lib.acmmp_scan_set_image_u8.argtypes = [vp, C.c_int, vp, C.c_int]
    lib.acmmp_scan_set_shard.argtypes = [vp, C.c_int, C.c_int]
    lib.acmmp_scan_set_sequential_geom.argtypes = [vp, C.c_int]
    lib.acmmp_scan_set_schedule.argtypes = [vp, C.c_int, C.c_int]
    lib.acmmp_nccl_get_unique_id.argtypes = [C.POINTER(C.c_uint8)]
    lib.acmmp_scan_comm_init.argtypes = [vp, C.POINTER(C.c_uint8)]
    lib.acmmp_scan_comm_init_all.argtypes = [C.POINTER(vp), C.c_int]
    lib.acmmp_scan_comm_destroy.argtypes = [vp]
    lib.acmmp_scan_all_gather.argtypes = [vp]
    lib.acmmp_scan_collect_view.argtypes = [vp, C.c_int, C.c_int, vp, vp]
    lib.acmmp_scan_run_dataflow.argtypes = [vp, C.c_uint64]
    lib.acmmp_scan_level_stats.argtypes = [vp, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.POINTER(C.c_int)]
    lib.acmmp_scan_dataflow_trace.argtypes = [vp, C.POINTER(C.c_float), C.c_int]
    lib.acmmp_set_box_staging.argtypes = [vp, C.c_int]
    lib.acmmp_scan_set_box_staging.argtypes = [vp, C.c_int]
    lib.acmmp_scan_keep_photometric.argtypes = [vp, C.c_int]
    lib.acmmp_scan_get_photometric_depth.argtypes = [vp, C.c_int, vp]
    lib.acmmp_scan_dataflow_stats.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.acmmp_plan_dataflow.argtypes = [C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float),
                                        C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    lib.acmmp_scan_set_texture_format.argtypes = [vp, C.c_int]
    lib.acmmp_scan_level_is_u8.argtypes = [vp]
    lib.acmmp_scan_num_levels.argtypes = [vp]
    lib.acmmp_scan_begin_level.argtypes = [vp, C.c_int]
    lib.acmmp_scan_run_stage.argtypes = [vp, C.c_int, C.c_uint64]
    lib.acmmp_scan_exchange.argtypes = [vp, C.POINTER(vp), C.POINTER(C.c_int64), C.POINTER(C.c_int)]
    lib.acmmp_scan_view_size.argtypes = [vp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.acmmp_scan_get_view.argtypes = [vp, C.c_int, fp, fp, fp]
    lib.acmmp_scan_timing.argtypes = [vp, C.POINTER(C.c_double * 4), C.POINTER(C.c_int), C.POINTER(C.c_uint64)]
    lib.acmmp_scan_sync.argtypes = [vp]
    lib.acmmp_scan_reset.argtypes = [vp]
    lib.acmmp_scan_host_stats.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.acmmp_fusion_create.argtypes = [C.POINTER(vp), C.c_int, C.c_int, C.POINTER(Camera), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    lib.acmmp_fusion_destroy.argtypes = [vp]
    lib.acmmp_fusion_destroy.restype = None
    lib.acmmp_fusion_last_error.argtypes = [vp]
    lib.acmmp_fusion_last_error.restype = C.c_char_p
    lib.acmmp_fusion_set_view.argtypes = [vp, C.c_int, vp, vp, vp, vp, C.c_int]
    lib.acmmp_fusion_run.argtypes = [vp, C.c_float, C.c_int, C.POINTER(C.c_int64)]
    lib.acmmp_fusion_set_view_second.argtypes = [vp, C.c_int, vp, vp, C.c_int]
    lib.acmmp_fusion_run_prior_aware.argtypes = [vp, C.c_float, C.c_int, C.c_int, C.POINTER(C.c_int64)]
    lib.acmmp_fusion_get_points.argtypes = [vp, vp, C.c_int64, C.POINTER(C.c_int64)]
    lib.acmmp_fusion_get_mask.argtypes = [vp, C.c_int, vp]
    lib.acmmp_fusion_stats.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_int32)]
    lib.acmmp_jpeg_info.argtypes = [vp, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.acmmp_decode_jpeg.argtypes = [C.c_int, vp, C.c_size_t, C.c_int, vp, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.acmmp_jpeg_last_error.restype = C.c_char_p
    dp = C.POINTER(C.c_double)
    lib.acmmp_pair_scores.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.c_int64, dp, dp,
                                      C.c_double, C.c_double, dp, dp]
    lib.acmmp_pairs_last_error.restype = C.c_char_p
    u16p = C.POINTER(C.c_uint16)
    lib.acmmp_scan_set_prior_maps.argtypes = [vp, C.c_int, u16p, u16p, C.c_int, C.c_int]
    lib.acmmp_prior_planes_from_maps.argtypes = [C.c_int, C.POINTER(Camera), u16p, u16p, C.c_int, C.c_int, C.c_int, C.c_int, fp]
    lib.acmmp_seeded_last_error.restype = C.c_char_p
    lib.acmmp_scan_view_device.argtypes = [vp, C.c_int, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    lib.acmmp_scan_count_ncc.argtypes = [vp, C.c_int]
    lib.acmmp_scan_ncc_count.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.acmmp_count_ncc.argtypes = [vp, C.c_int]
    lib.acmmp_ncc_count.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.acmmp_scan_mark.argtypes = [vp, C.c_int]
    lib.acmmp_scan_elapsed.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_float)]
    _lib = lib
    return lib


def fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


def uptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32)) if a is not None else None


def f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)

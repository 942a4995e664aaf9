"""ctypes loader for libgsb200.so (the C ABI declared in include/gsb.h).

The CUDA library is the product: importing a compute entry point without it raises.  There is
no CPU fallback anywhere in this package.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GSB_LIB") or os.path.join(_HERE, "libgsb200.so")   # GSB_LIB: developer A/B builds

ALLOC_FN = C.CFUNCTYPE(C.c_void_p, C.c_size_t, C.c_void_p)
ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p)


class ProjectParams(C.Structure):
    """gsb_project_params: the Project fields Trainer::train reads (reference src/Project.h:26-39)."""
    _fields_ = [(n, C.c_float) for n in (
        "lrLocation", "lrSh", "lrScale", "lrOpacity", "lrRotation", "paramScaleMax",
        "paramCullOpacity", "paramCullSize", "paramDensifyVariance", "paramSplitSize",
        "paramSplitDistance", "paramSplitScale", "paramCloneDistance")]


class CameraC(C.Structure):
    """gsb_camera."""
    _fields_ = [("view", C.c_float * 16), ("projview", C.c_float * 16), ("location", C.c_float * 3),
                ("tan_fovx", C.c_float), ("tan_fovy", C.c_float)]


class StateView(C.Structure):
    """gsb_state_view."""
    _fields_ = [(n, C.c_void_p) for n in (
        "radii", "depths", "tiles_touched", "recs", "cov3D", "clamped", "point_list",
        "ranges", "n_contrib", "final_T", "tile_max", "counters")]


class TrainerStats(C.Structure):
    """gsb_trainer_stats."""
    _fields_ = [("rendered_total", C.c_longlong), ("visible_total", C.c_longlong), ("views", C.c_int),
                ("retries", C.c_int), ("kernel_launches", C.c_longlong), ("sum_sq_residual", C.c_double),
                ("epoch_wraps", C.c_longlong), ("exchange_ms", C.c_double), ("apply_ms", C.c_double),
                ("timed_iterations", C.c_longlong)]


_lib = None


def lib():
    """Load libgsb200.so once; fail loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C gaussian-splatterer_b200`). This package has no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, i, f, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    L.gsb_last_error.restype = C.c_char_p
    L.gsb_version.restype = i
    L.gsb_rasterizer_forward.argtypes = [ALLOC_FN, vp, ALLOC_FN, vp, ALLOC_FN, vp, i, i, i, vp, i, i, vp, vp, vp,
                                         vp, vp, f, vp, vp, vp, vp, vp, f, f, i, vp, vp, i, vp, C.POINTER(i)]
    L.gsb_rasterizer_forward.restype = i
    L.gsb_rasterizer_backward.argtypes = [i, i, i, i, vp, i, i, vp, vp, vp, vp, f, vp, vp, vp, vp, vp, f, f, vp,
                                          vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i, vp]
    L.gsb_rasterizer_backward.restype = i
    L.gsb_rasterizer_mark_visible.argtypes = [i, vp, vp, vp, vp, vp]
    L.gsb_rasterizer_mark_visible.restype = i
    L.gsb_rasterizer_state.argtypes = [i, i, i, i, vp, vp, vp, C.POINTER(StateView)]
    L.gsb_rasterizer_state.restype = i
    L.gsb_rasterizer_sorted_keys.argtypes = [i, i, i, i, vp, vp, vp, vp, vp]
    L.gsb_rasterizer_sorted_keys.restype = i
    L.gsb_rasterizer_tile_order.argtypes = [i, i, i, vp, vp, vp, vp]
    L.gsb_rasterizer_tile_order.restype = i
    L.gsb_trainer_create.argtypes = [C.POINTER(vp), i, i, vp]
    L.gsb_trainer_create.restype = i
    L.gsb_trainer_destroy.argtypes = [vp]
    L.gsb_trainer_destroy.restype = i
    L.gsb_trainer_bind_model.argtypes = [vp, i, i, i, i, vp, vp, vp, vp, vp]
    L.gsb_trainer_bind_model.restype = i
    L.gsb_trainer_set_truths.argtypes = [vp, i, C.POINTER(vp), C.POINTER(vp), C.POINTER(CameraC), i]
    L.gsb_trainer_set_truths.restype = i
    L.gsb_trainer_set_grad_buffer.argtypes = [vp, vp]
    L.gsb_trainer_set_grad_buffer.restype = i
    L.gsb_trainer_grad_floats.argtypes = [vp]
    L.gsb_trainer_grad_floats.restype = sz
    L.gsb_trainer_grad_buffer.argtypes = [vp]
    L.gsb_trainer_grad_buffer.restype = vp
    L.gsb_trainer_set_allreduce.argtypes = [vp, ALLREDUCE_FN, vp]
    L.gsb_trainer_set_allreduce.restype = i
    L.gsb_trainer_accumulate.argtypes = [vp]
    L.gsb_trainer_accumulate.restype = i
    L.gsb_trainer_apply.argtypes = [vp, C.POINTER(ProjectParams)]
    L.gsb_trainer_apply.restype = i
    L.gsb_trainer_train.argtypes = [vp, C.POINTER(ProjectParams), i, C.POINTER(i)]
    L.gsb_trainer_train.restype = i
    L.gsb_trainer_densify_plan.argtypes = [vp, C.POINTER(ProjectParams), C.POINTER(i)]
    L.gsb_trainer_densify_plan.restype = i
    L.gsb_trainer_densify_commit.argtypes = [vp, vp, vp, vp, vp, vp]
    L.gsb_trainer_densify_commit.restype = i
    L.gsb_trainer_render.argtypes = [vp, vp, i, i, f, C.POINTER(CameraC)]
    L.gsb_trainer_render.restype = i
    L.gsb_trainer_get_stats.argtypes = [vp, C.POINTER(TrainerStats)]
    L.gsb_trainer_get_stats.restype = i
    L.gsb_trainer_get_loss.argtypes = [vp, i, C.POINTER(C.c_double)]
    L.gsb_trainer_get_loss.restype = i
    L.gsb_trainer_enable_loss.argtypes = [vp, i]
    L.gsb_trainer_enable_loss.restype = i
    L.gsb_trainer_set_fuse_backgrounds.argtypes = [vp, i]
    L.gsb_trainer_set_fuse_backgrounds.restype = i
    L.gsb_trainer_profile_pass.argtypes = [vp, i, C.POINTER(C.c_float), C.POINTER(C.c_uint)]
    L.gsb_trainer_profile_pass.restype = i
    L.gsb_trainer_set_overlap.argtypes = [vp, i]
    L.gsb_trainer_set_overlap.restype = i
    L.gsb_trainer_set_stage_timing.argtypes = [vp, i]
    L.gsb_trainer_set_stage_timing.restype = i
    L.gsb_trainer_get_stage_times.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_longlong)]
    L.gsb_trainer_get_stage_times.restype = i
    L.gsb_trainer_count_pairs.argtypes = [vp, i, C.POINTER(C.c_ulonglong)]
    L.gsb_trainer_count_pairs.restype = i
    L.gsb_comm_nccl_version.argtypes = [C.POINTER(i)]
    L.gsb_comm_get_unique_id.argtypes = [vp]
    L.gsb_comm_init.argtypes = [C.POINTER(vp), vp, i, i]
    L.gsb_comm_init_all.argtypes = [C.POINTER(vp), i, C.POINTER(i)]
    L.gsb_comm_destroy.argtypes = [vp]
    L.gsb_comm_rank.argtypes = [vp]
    L.gsb_comm_world.argtypes = [vp]
    L.gsb_comm_device.argtypes = [vp]
    L.gsb_comm_allreduce_sum.argtypes = [vp, vp, sz, vp]
    L.gsb_comm_broadcast.argtypes = [vp, vp, sz, i, vp]
    L.gsb_trainer_set_comm.argtypes = [vp, vp]
    for _n in ("nccl_version", "get_unique_id", "init", "init_all", "destroy", "rank", "world", "device", "allreduce_sum",
               "broadcast"):
        getattr(L, "gsb_comm_" + _n).restype = i
    L.gsb_trainer_set_comm.restype = i
    L.gsb_comm_window_create.argtypes = [vp, sz, C.POINTER(vp)]
    L.gsb_comm_window_create.restype = i
    L.gsb_comm_window_destroy.argtypes = [vp]
    L.gsb_comm_window_destroy.restype = i
    L.gsb_window_data.argtypes = [vp]
    L.gsb_window_data.restype = vp
    L.gsb_window_floats.argtypes = [vp]
    L.gsb_window_floats.restype = sz
    L.gsb_window_peer_mapped.argtypes = [vp]
    L.gsb_window_peer_mapped.restype = i
    L.gsb_window_allreduce_sum.argtypes = [vp, sz, vp]
    L.gsb_window_allreduce_sum.restype = i
    L.gsb_trainer_exchange_path.argtypes = [vp]
    L.gsb_trainer_exchange_path.restype = i
    L.gsb_trainer_measure_bwd_skeleton.argtypes = [vp, i, i, C.POINTER(C.c_float)]
    L.gsb_trainer_measure_bwd_skeleton.restype = i
    L.gsb_trainer_measure_staging.argtypes = [vp, i, i, C.POINTER(C.c_float)]
    L.gsb_trainer_measure_staging.restype = i
    L.gsb_trainer_reduction_stats.argtypes = [vp, i, C.POINTER(C.c_ulonglong)]
    L.gsb_trainer_reduction_stats.restype = i
    L.gsb_measure_fp32_peak.argtypes = [C.POINTER(C.c_double), vp]
    L.gsb_measure_fp32_peak.restype = i
    L.gsb_trainer_debug_set_epoch.argtypes = [vp, C.c_uint]
    L.gsb_trainer_debug_set_epoch.restype = i
    L.gsb_image_float_to_int.argtypes = [vp, vp, i, i, vp]
    L.gsb_image_float_to_int.restype = i
    L.gsb_image_int_to_loss.argtypes = [vp, vp, vp, i, i, vp]
    L.gsb_image_int_to_loss.restype = i
    _lib = L
    return L


# Every symbol include/gsb.h declares; the CPU test-suite checks the library exports all of them.
EXPORTED_SYMBOLS = [
    "gsb_last_error", "gsb_version", "gsb_rasterizer_forward", "gsb_rasterizer_backward",
    "gsb_rasterizer_mark_visible", "gsb_rasterizer_state", "gsb_rasterizer_sorted_keys", "gsb_rasterizer_tile_order",
    "gsb_trainer_enable_loss", "gsb_trainer_get_loss", "gsb_trainer_set_fuse_backgrounds", "gsb_trainer_profile_pass",
    "gsb_trainer_set_stage_timing", "gsb_trainer_get_stage_times", "gsb_trainer_count_pairs", "gsb_trainer_set_overlap",
    "gsb_trainer_debug_set_epoch", "gsb_measure_fp32_peak", "gsb_trainer_measure_bwd_skeleton", "gsb_trainer_measure_staging", "gsb_trainer_reduction_stats",
    "gsb_comm_nccl_version", "gsb_comm_get_unique_id", "gsb_comm_init", "gsb_comm_init_all", "gsb_comm_destroy",
    "gsb_comm_rank", "gsb_comm_world", "gsb_comm_device", "gsb_comm_allreduce_sum", "gsb_comm_broadcast",
    "gsb_trainer_set_comm", "gsb_comm_window_create", "gsb_comm_window_destroy", "gsb_window_data", "gsb_window_floats",
    "gsb_window_peer_mapped", "gsb_window_allreduce_sum", "gsb_trainer_exchange_path",
    "gsb_image_float_to_int", "gsb_image_int_to_loss",
    "gsb_trainer_create", "gsb_trainer_destroy", "gsb_trainer_bind_model", "gsb_trainer_set_truths",
    "gsb_trainer_set_grad_buffer", "gsb_trainer_grad_floats", "gsb_trainer_grad_buffer",
    "gsb_trainer_set_allreduce", "gsb_trainer_accumulate", "gsb_trainer_apply", "gsb_trainer_train",
    "gsb_trainer_densify_plan", "gsb_trainer_densify_commit", "gsb_trainer_render", "gsb_trainer_get_stats",
]


class GsbError(RuntimeError):
    """Raised for any non-zero return of the C ABI (the C++ compat layer throws std::runtime_error)."""


def check(code):
    if code != 0:
        raise GsbError(lib().gsb_last_error().decode("utf-8", "replace"))

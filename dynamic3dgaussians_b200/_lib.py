"""ctypes binding of libgsr_b200.so (the C ABI declared in include/gsr_b200.h).

There is deliberately no fallback: if the CUDA library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

HERE = os.path.dirname(os.path.abspath(__file__))
# GSR_LIB_PATH: load another build of the same library (A/B runs of kernel variants; tests/gpu_ab.sh)
LIB_PATH = os.environ.get("GSR_LIB_PATH") or os.path.join(HERE, "libgsr_b200.so")

PEER_ALLREDUCE, PEER_REDUCE_SCATTER, PEER_ALLGATHER = 0, 1, 2
PEER_MAX_WORLD = 8
PEER_MAX_CTAS = 256
PEER_FLAG_WORDS = 4608            # include/gsr_b200.h GSR_PEER_*
PEER_ERROR_WORD = 2 * PEER_MAX_CTAS * PEER_MAX_WORLD + PEER_MAX_CTAS

EXPORTED_SYMBOLS = [
    "gsr_abi_version", "gsr_strerror", "gsr_last_cuda_error", "gsr_set_option", "gsr_get_option",
    "gsr_workspace_sizes", "gsr_workspace_layout",
    "gsr_forward_geometry", "gsr_forward_render", "gsr_read_counters", "gsr_backward", "gsr_mark_visible",
    "gsr_export_view_binning", "gsr_sort_temp_bytes", "gsr_sort_pairs", "gsr_sort_pairs_static_max",
    "gsr_sort_pairs_static", "gsr_peer_allreduce", "gsr_knn_temp_bytes", "gsr_knn",
    "gsr_dyn_losses_temp_bytes", "gsr_dyn_losses",
    "gsr_image_loss_state_bytes", "gsr_image_loss_forward", "gsr_image_loss_backward", "gsr_adam_step",
    "gsr_fp32_peak_probe", "gsr_probe_geometry", "gsr_stage_timing_enable", "gsr_stage_timing_count", "gsr_stage_name",
    "gsr_stage_timing_read", "gsr_debug_blend_trace",
    "gsr_activations_forward", "gsr_activations_backward", "gsr_cam_affine_forward", "gsr_cam_affine_backward",
    "gsr_densify_stats", "gsr_l1_u8_loss_forward", "gsr_l1_u8_loss_backward",
]


class GsrAdamTensor(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("step", C.c_void_p), ("n", C.c_int64), ("lr", C.c_float), ("pad_", C.c_int32)]


ADAM_MAX_TENSORS = 16
ADAM_SCRATCH_BYTES = 256


class GsrCamera(C.Structure):
    _fields_ = [
        ("image_width", C.c_int32), ("image_height", C.c_int32),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("c_x", C.c_float), ("c_y", C.c_float),
        ("scale_modifier", C.c_float), ("sh_degree", C.c_int32), ("prefiltered", C.c_int32),
        ("bg", C.c_float * 8), ("viewmatrix", C.c_float * 16), ("projmatrix", C.c_float * 16),
        ("campos", C.c_float * 3), ("tile_row_begin", C.c_int32), ("tile_row_end", C.c_int32), ("debug", C.c_int32),
    ]


FLAG_DEBUG = 1
ABI_VERSION = 4


class GsrWorkspaceSizes(C.Structure):
    _fields_ = [("geom_bytes", C.c_size_t), ("binning_bytes", C.c_size_t), ("image_bytes", C.c_size_t)]


_LAYOUT_SIZE_T = ["cams", "counters", "radii", "rects", "geomrec", "cov3D", "rgb", "col4", "ext4", "clamped",
                  "dkeys0", "dkeys1", "dvals0", "dvals1", "dhist", "dlookback", "dcounters", "block_sums", "grad_rec",
                  "ikeys0", "ikeys1", "ivals0", "ivals1", "imasks", "ihist", "ilookback", "st_ranges", "chunk_base", "chunk_key",
                  "chunk_counts", "recA", "recB", "recC", "recD",
                  "ranges", "tile_start", "tile_order", "n_contrib", "final_T"]
_LAYOUT_INT = ["tiles_x", "tiles_y", "stiles_x", "stiles_y", "tile_bits", "depth_passes", "tile_passes", "sorted_keys_in",
               "sorted_vals_in", "max_chunks"]


class GsrLayout(C.Structure):
    _fields_ = [(n, C.c_size_t) for n in _LAYOUT_SIZE_T] + [(n, C.c_int32) for n in _LAYOUT_INT]


_lib: Optional[C.CDLL] = None


def load(build_if_missing: bool = True) -> C.CDLL:
    """Loads (building first when the .so is absent and nvcc is available) the CUDA library."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH) and build_if_missing:
        from . import build as _build
        _build.build()
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -m dynamic3dgaussians_b200.build` (no CPU fallback exists)")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, sz = C.c_void_p, C.c_int, C.c_int64, C.c_size_t
    lib.gsr_abi_version.restype = C.c_int
    lib.gsr_strerror.restype = C.c_char_p
    lib.gsr_strerror.argtypes = [C.c_int]
    lib.gsr_last_cuda_error.restype = C.c_char_p
    lib.gsr_workspace_sizes.argtypes = [i32, i32, i32, i32, i32, i32, i64, C.POINTER(GsrWorkspaceSizes)]
    lib.gsr_workspace_layout.argtypes = [i32, i32, i32, i32, i32, i32, i64, C.POINTER(GsrLayout)]
    lib.gsr_forward_geometry.argtypes = [C.POINTER(GsrCamera), i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, sz,
                                         vp, vp, vp]
    lib.gsr_forward_render.argtypes = [i32, i32, i32, i32, i32, i32, i64, vp, vp, vp, i32] + [vp] * 8 + [i32, vp]
    lib.gsr_read_counters.argtypes = [vp, vp, vp]
    lib.gsr_backward.argtypes = [i32, i32, i32, i32, i32, i32, i64] + [vp] * 9 + [i32] + [vp] * 21 + [i32, vp]
    lib.gsr_mark_visible.argtypes = [C.POINTER(GsrCamera), i32, vp, vp, vp]
    lib.gsr_export_view_binning.argtypes = [i32, i32, i32, i32, i32, i32, i32, i64, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.gsr_sort_temp_bytes.restype = sz
    lib.gsr_sort_temp_bytes.argtypes = [i64, i32]
    lib.gsr_sort_pairs.argtypes = [i32, vp, vp, vp, vp, i64, i32, i32, vp, sz, vp]
    lib.gsr_sort_pairs_static_max.restype = i64
    lib.gsr_sort_pairs_static_max.argtypes = []
    lib.gsr_sort_pairs_static.argtypes = [i32, vp, vp, vp, vp, i64, i32, i32, vp]
    lib.gsr_peer_allreduce.argtypes = [i32, i32, i32, vp, vp, vp, i64, i32, vp]
    lib.gsr_knn_temp_bytes.restype = sz
    lib.gsr_knn_temp_bytes.argtypes = [i64]
    lib.gsr_knn.argtypes = [i64, i32, vp, vp, vp, vp, vp, sz, vp]
    lib.gsr_dyn_losses_temp_bytes.restype = sz
    lib.gsr_dyn_losses_temp_bytes.argtypes = [i64]
    lib.gsr_dyn_losses.argtypes = [i64, i32] + [vp] * 12 + [sz, vp]
    lib.gsr_image_loss_state_bytes.restype = sz
    lib.gsr_image_loss_state_bytes.argtypes = [i64, i32, i32]
    lib.gsr_image_loss_forward.argtypes = [i64, i32, i32, vp, vp, C.c_float, C.c_float, vp, sz, vp, vp]
    lib.gsr_image_loss_backward.argtypes = [i64, i32, i32, vp, vp, C.c_float, C.c_float, vp, sz, vp, vp, vp]
    lib.gsr_adam_step.argtypes = [i32, C.POINTER(GsrAdamTensor), C.c_double, C.c_double, C.c_double, vp, vp]
    lib.gsr_fp32_peak_probe.argtypes = [vp, i32, i32, vp]
    lib.gsr_set_option.argtypes = [C.c_char_p, i32]
    lib.gsr_get_option.argtypes = [C.c_char_p, C.POINTER(C.c_int)]
    lib.gsr_stage_timing_enable.argtypes = [i32]
    lib.gsr_stage_name.restype = C.c_char_p
    lib.gsr_stage_name.argtypes = [i32]
    lib.gsr_stage_timing_read.argtypes = [vp, i32]
    lib.gsr_probe_geometry.argtypes = [C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.gsr_debug_blend_trace.argtypes = [vp, vp]
    lib.gsr_activations_forward.argtypes = [i64, vp, vp, vp, vp, vp, vp, vp]
    lib.gsr_activations_backward.argtypes = [i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.gsr_cam_affine_forward.argtypes = [i64, vp, vp, vp, vp, vp]
    lib.gsr_cam_affine_backward.argtypes = [i64, vp, vp, vp, vp, vp, vp, vp]
    lib.gsr_densify_stats.argtypes = [i64, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.gsr_l1_u8_loss_forward.argtypes = [i64, i32, i32, vp, vp, vp, C.c_float, vp, vp]
    lib.gsr_l1_u8_loss_backward.argtypes = [i64, i32, i32, vp, vp, vp, C.c_float, vp, vp, vp]
    for name in EXPORTED_SYMBOLS:
        getattr(lib, name)                                   # every declared entry point must be exported
    if lib.gsr_abi_version() != ABI_VERSION:
        raise RuntimeError(f"{LIB_PATH} has ABI version {lib.gsr_abi_version()}, this package needs {ABI_VERSION}: "
                           f"rebuild with `python -m dynamic3dgaussians_b200.build --force`")
    _lib = lib
    # The library itself never reads the environment (gsr_set_option); the tuning variables the tools of round 1 used are
    # forwarded once, here, by the host binding.
    for env, key in (("GSR_BLEND_ILP", "blend_ilp"), ("GSR_BWD_COST_ORDER", "bwd_cost_order"), ("GSR_NVTX", "nvtx"), ("GSR_PDL", "pdl"), ("GSR_BWD_ILP", "bwd_ilp")):
        v = os.environ.get(env)
        if v not in (None, ""):
            check(lib.gsr_set_option(key.encode(), int(v)), f"set_option({key})")
    return lib


def set_option(key: str, value: int) -> None:
    check(load().gsr_set_option(key.encode(), int(value)), f"set_option({key})")


def get_option(key: str) -> int:
    v = C.c_int(0)
    check(load().gsr_get_option(key.encode(), C.byref(v)), f"get_option({key})")
    return int(v.value)


def check(status: int, what: str = "") -> None:
    if status == 0:
        return
    lib = load()
    msg = lib.gsr_strerror(status).decode()
    if status == -3:
        msg += ": " + lib.gsr_last_cuda_error().decode()
    raise RuntimeError(f"gsr_b200 {what} failed: {msg}")


def ptr(t) -> Optional[int]:
    """Device pointer of a torch tensor (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()


def stage_timing(enable: bool) -> None:
    check(load().gsr_stage_timing_enable(1 if enable else 0), "stage_timing_enable")


def read_stage_timing() -> dict:
    """{stage name: milliseconds} of the most recent forward/backward (waits for their events)."""
    lib = load()
    n = lib.gsr_stage_timing_count()
    buf = (C.c_float * n)()
    check(lib.gsr_stage_timing_read(buf, n), "stage_timing_read")
    return {lib.gsr_stage_name(i).decode(): float(buf[i]) for i in range(n) if buf[i] >= 0}

"""ctypes binding of libdsdf_b200.so (include/dsdf.h).

The library is the product: if it is missing or fails to load, importing the compute entry
points raises -- there is no PyTorch/CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# DSDF_LIB selects a diagnostics build of the same library (build.py DSDF_BUILD_VARIANT); it is
# still the CUDA library -- there is no non-CUDA implementation to select.
LIB_PATH = os.environ.get("DSDF_LIB") or os.path.join(_HERE, "libdsdf_b200.so")

DSDF_NUM_LIN = 9

EXPORTS = [
    "dsdf_arch", "dsdf_create", "dsdf_destroy", "dsdf_last_error", "dsdf_num_sms", "dsdf_set_weights",
    "dsdf_forward_dense", "dsdf_forward_segments", "dsdf_decode_grid", "dsdf_grid_coords",
    "dsdf_train_workspace_bytes", "dsdf_train_fwd_bwd", "dsdf_renorm_rows", "dsdf_clip_adam",
    "dsdf_dense_workspace_bytes", "dsdf_dense_forward_train", "dsdf_dense_backward",
    "dsdf_segments_forward_train", "dsdf_segments_backward", "dsdf_set_dropout", "dsdf_dropout_keep", "dsdf_reconstruct_workspace_bytes", "dsdf_reconstruct",
    "dsdf_selftest_gemm", "dsdf_profile_enable", "dsdf_profile_read", "dsdf_debug_read", "dsdf_selftest_mma_rate", "dsdf_debug_hang_buffer",
    "dsdf_debug_progress_buffer", "dsdf_mc_classify", "dsdf_mc_emit",
    "dsdf_rasterize_views", "dsdf_rasterize_views_static", "dsdf_sample_points", "dsdf_knn_sdf", "dsdf_knn_grid_workspace_bytes", "dsdf_knn_sdf_grid", "dsdf_gather_windows",
]


class DsdfConfig(C.Structure):
    _fields_ = [("latent_size", C.c_int32), ("hidden", C.c_int32), ("num_hidden", C.c_int32),
                ("latent_in", C.c_int32), ("weight_norm_mask", C.c_uint32), ("device", C.c_int32)]


class DsdfParamLayout(C.Structure):
    _fields_ = [("w_off", C.c_int64 * DSDF_NUM_LIN), ("g_off", C.c_int64 * DSDF_NUM_LIN),
                ("b_off", C.c_int64 * DSDF_NUM_LIN), ("total", C.c_int64)]


class DsdfError(RuntimeError):
    pass


_libs = {}
SIREN_LIB_PATH = os.environ.get("DSDF_LIB_SIREN") or LIB_PATH[:-3] + "_siren.so"


def lib(arch: str = "deep_sdf_decoder"):
    """Load (once) and return the C library of network family `arch` with argtypes set.
    One source tree, one build per family (build.py); both export the same C ABI."""
    if arch in _libs:
        return _libs[arch]
    if arch not in ("deep_sdf_decoder", "siren_decoder"):
        raise DsdfError(f"no library for NetworkArch {arch!r}")
    path = LIB_PATH if arch == "deep_sdf_decoder" else SIREN_LIB_PATH
    if not os.path.exists(path):
        raise DsdfError(
            f"{path} not found: build it with `python medical-shape-disentanglement_b200/build.py` "
            "(there is no CPU / PyTorch fallback for this path)")
    L = C.CDLL(path)
    L.dsdf_arch.argtypes = []
    L.dsdf_arch.restype = C.c_char_p
    if L.dsdf_arch().decode() != arch:
        raise DsdfError(f"{path} implements {L.dsdf_arch().decode()}, expected {arch}")
    vp, i64, i32, f32, sz = C.c_void_p, C.c_int64, C.c_int32, C.c_float, C.c_size_t
    L.dsdf_create.argtypes = [C.POINTER(DsdfConfig), C.POINTER(vp)]
    L.dsdf_create.restype = C.c_int
    L.dsdf_destroy.argtypes = [vp]
    L.dsdf_destroy.restype = C.c_int
    L.dsdf_last_error.argtypes = [vp]
    L.dsdf_last_error.restype = C.c_char_p
    L.dsdf_num_sms.argtypes = [vp]
    L.dsdf_num_sms.restype = C.c_int
    L.dsdf_set_weights.argtypes = [vp, vp, C.POINTER(DsdfParamLayout), vp]
    L.dsdf_set_weights.restype = C.c_int
    L.dsdf_forward_dense.argtypes = [vp, vp, i64, vp, vp]
    L.dsdf_forward_dense.restype = C.c_int
    L.dsdf_forward_segments.argtypes = [vp, vp, vp, i64, i64, vp, i64, i64, vp, vp]
    L.dsdf_forward_segments.restype = C.c_int
    L.dsdf_decode_grid.argtypes = [vp, vp, i32, i64, i64, vp, vp]
    L.dsdf_decode_grid.restype = C.c_int
    L.dsdf_grid_coords.argtypes = [vp, i32, i64, i64, vp, vp]
    L.dsdf_grid_coords.restype = C.c_int
    L.dsdf_train_workspace_bytes.argtypes = [vp, i64, i64]
    L.dsdf_train_workspace_bytes.restype = sz
    L.dsdf_train_fwd_bwd.argtypes = [vp, vp, vp, i64, vp, i64, i64, vp, i64, f32, f32, f32, vp, sz,
                                     vp, vp, vp, vp, vp, vp]
    L.dsdf_train_fwd_bwd.restype = C.c_int
    L.dsdf_renorm_rows.argtypes = [vp, vp, i64, vp, i64, f32, vp]
    L.dsdf_renorm_rows.restype = C.c_int
    L.dsdf_clip_adam.argtypes = [vp, vp, vp, vp, vp, i64, f32, f32, f32, f32, f32, i64, vp, vp]
    L.dsdf_clip_adam.restype = C.c_int
    L.dsdf_dense_workspace_bytes.argtypes = [vp, i64]
    L.dsdf_dense_workspace_bytes.restype = sz
    L.dsdf_dense_forward_train.argtypes = [vp, vp, i64, vp, vp, sz, vp]
    L.dsdf_dense_forward_train.restype = C.c_int
    L.dsdf_dense_backward.argtypes = [vp, vp, vp, i64, vp, vp, vp, sz, vp, vp, vp]
    L.dsdf_dense_backward.restype = C.c_int
    L.dsdf_segments_forward_train.argtypes = [vp, vp, vp, i64, i64, vp, i64, vp, vp, sz, vp]
    L.dsdf_segments_forward_train.restype = C.c_int
    L.dsdf_segments_backward.argtypes = [vp, vp, vp, i64, i64, vp, i64, vp, vp, vp, sz, vp, vp, vp]
    L.dsdf_segments_backward.restype = C.c_int
    L.dsdf_set_dropout.argtypes = [vp, C.c_uint32, f32, C.c_uint64, C.c_uint64, i64]
    L.dsdf_set_dropout.restype = C.c_int
    L.dsdf_dropout_keep.argtypes = [vp, i32, i64, i64, vp, vp]
    L.dsdf_dropout_keep.restype = C.c_int
    L.dsdf_reconstruct_workspace_bytes.argtypes = [vp, i64, i64]
    L.dsdf_reconstruct_workspace_bytes.restype = sz
    L.dsdf_reconstruct.argtypes = [vp, vp, vp, i64, vp, vp, vp, i64, i64, vp, vp, vp, vp, sz, vp, vp]
    L.dsdf_reconstruct.restype = C.c_int
    L.dsdf_selftest_gemm.argtypes = [i32, vp, vp, vp, vp]
    L.dsdf_selftest_gemm.restype = C.c_int
    L.dsdf_profile_enable.argtypes = [vp, i32]
    L.dsdf_profile_enable.restype = C.c_int
    L.dsdf_profile_read.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_int32)]
    L.dsdf_profile_read.restype = C.c_int
    L.dsdf_debug_read.argtypes = [vp, C.POINTER(C.c_uint64), i32]
    L.dsdf_debug_read.restype = C.c_int
    L.dsdf_selftest_mma_rate.argtypes = [i32, i32, i32, C.POINTER(C.c_uint64)]
    L.dsdf_selftest_mma_rate.restype = C.c_int
    L.dsdf_debug_hang_buffer.argtypes = [C.POINTER(C.POINTER(C.c_uint32))]
    L.dsdf_debug_hang_buffer.restype = C.c_int
    L.dsdf_debug_progress_buffer.argtypes = [C.POINTER(C.POINTER(C.c_uint32))]
    L.dsdf_debug_progress_buffer.restype = C.c_int
    L.dsdf_mc_classify.argtypes = [vp, i32, i32, i32, f32, vp, vp, vp]
    L.dsdf_mc_classify.restype = C.c_int
    L.dsdf_mc_emit.argtypes = [vp, i32, i32, i32, f32, f32, f32, f32, vp, vp, vp, vp, vp]
    L.dsdf_mc_emit.restype = C.c_int
    L.dsdf_rasterize_views.argtypes = [vp, i64, vp, i64, vp, i32, vp, vp, vp, vp, vp]
    L.dsdf_rasterize_views.restype = C.c_int
    L.dsdf_rasterize_views_static.argtypes = [vp, i64, vp, i64, vp, i32, vp, vp, vp, vp]
    L.dsdf_rasterize_views_static.restype = C.c_int
    L.dsdf_sample_points.argtypes = [vp, vp, vp, i64, i64, i64, f32, f32, f32, C.c_uint64, vp, vp, vp]
    L.dsdf_sample_points.restype = C.c_int
    L.dsdf_knn_sdf.argtypes = [vp, i64, vp, vp, i64, f32, vp, vp, vp]
    L.dsdf_knn_sdf.restype = C.c_int
    L.dsdf_knn_grid_workspace_bytes.argtypes = [i64]
    L.dsdf_knn_grid_workspace_bytes.restype = C.c_size_t
    L.dsdf_knn_sdf_grid.argtypes = [vp, i64, vp, vp, i64, f32, vp, C.c_size_t, vp, vp, vp]
    L.dsdf_knn_sdf_grid.restype = C.c_int
    L.dsdf_gather_windows.argtypes = [vp, vp, vp, i64, i64, vp, vp]
    L.dsdf_gather_windows.restype = C.c_int
    _libs[arch] = L
    return L


def check(rc: int, handle=None, L=None) -> None:
    if rc != 0:
        msg = (L or lib()).dsdf_last_error(handle)
        raise DsdfError(f"libdsdf_b200 error {rc}: {msg.decode() if msg else '?'}")

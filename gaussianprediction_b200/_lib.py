"""ctypes binding of libgp_b200.so (C ABI declared in include/gp_b200.h).

There is NO fallback: if the shared library is missing or a call fails, an exception is raised.
The library is built in-tree by ``python -m gaussianprediction_b200.build`` (or ``__graft_entry__.build()``).
"""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_longlong, c_size_t, c_void_p

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "libgp_b200.so")


class GpError(RuntimeError):
    pass


class RasterCfg(ctypes.Structure):
    """Mirror of ``gp_raster_cfg``."""
    _fields_ = [
        ("P", c_int), ("width", c_int), ("height", c_int), ("sh_degree", c_int), ("sh_coeffs", c_int),
        ("tanfovx", c_float), ("tanfovy", c_float), ("scale_modifier", c_float), ("prefiltered", c_int),
        ("sh_rest", c_void_p), ("dL_dsh_rest", c_void_p), ("dL_dsh_accumulate", c_int),
    ]


class MlpCfg(ctypes.Structure):
    """Mirror of ``gp_mlp_cfg``."""
    _fields_ = [("rows", c_int), ("feature_dim", c_int), ("xyz_freq", c_int), ("time_freq", c_int), ("depth", c_int),
                ("width", c_int), ("out_dim", c_int), ("normalize_q", c_int)]


class AdamTensor(ctypes.Structure):
    """Mirror of ``gp_adam_tensor``."""
    _fields_ = [("param", c_void_p), ("grad", c_void_p), ("exp_avg", c_void_p), ("exp_avg_sq", c_void_p),
                ("n", c_longlong), ("lr", c_float), ("step", c_int)]


HDR_NUM_RENDERED, HDR_OVERFLOW, HDR_NUM_POINTS, HDR_NUM_SORTED, HDR_WORDS = 0, 1, 2, 3, 64

_P = c_void_p  # every device pointer crosses the ABI as a plain address

# name -> (restype, argtypes); must list every symbol include/gp_b200.h declares (tests check this).
SIGNATURES = {
    "gp_abi_version": (c_int, []),
    "gp_error_string": (c_char_p, [c_int]),
    "gp_launch_count": (c_longlong, []),
    "gp_profile_enable": (c_int, [c_int]),
    "gp_profile_report": (c_int, [ctypes.c_char_p, c_size_t]),
    "gp_fma_peak": (c_int, [c_int, _P, POINTER(c_double), _P]),
    "gp_raster_sizes": (c_int, [c_int, c_int, c_int, c_longlong, POINTER(c_size_t), POINTER(c_size_t), POINTER(c_size_t)]),
    "gp_raster_preprocess": (c_int, [POINTER(RasterCfg)] + [_P] * 10 + [_P, _P, _P]),
    "gp_raster_num_rendered": (c_int, [_P, _P, POINTER(c_int)]),
    "gp_raster_render": (c_int, [POINTER(RasterCfg), _P, _P, _P, _P, c_longlong, _P, _P, _P, _P, _P, _P]),
    "gp_raster_backward_ws_bytes": (c_size_t, [c_int]),
    "gp_raster_backward": (c_int, [POINTER(RasterCfg)] + [_P] * 13 + [c_longlong] + [_P] * 12),
    "gp_raster_backward_static_max_gaussians": (c_int, []),
    "gp_raster_backward_static": (c_int, [POINTER(RasterCfg)] + [_P] * 13 + [c_longlong] + [_P] * 11),
    "gp_mark_visible": (c_int, [c_int, _P, _P, _P, _P, _P]),
    "gp_raster_export_state": (c_int, [POINTER(RasterCfg)] + [_P] * 13),
    "gp_knn_points": (c_int, [c_int, c_int, c_int, c_int, c_float, _P, _P, _P, _P, _P]),
    "gp_knn_keypoints": (c_int, [c_int, c_int, c_int, c_int, c_float, c_float, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "gp_spatial_order_ws_bytes": (c_size_t, [c_int]),
    "gp_spatial_order": (c_int, [c_int, _P, _P, _P, _P]),
    "gp_blend_forward": (c_int, [c_int, c_int, c_int] + [_P] * 14),
    "gp_blend_backward": (c_int, [c_int, c_int, c_int] + [_P] * 20),
    "gp_blend_backward_accumulate": (c_int, [c_int, c_int, c_int] + [_P] * 20),
    "gp_blend_dense_weights": (c_int, [c_int, c_int, c_int] + [_P] * 7),
    "gp_dist2_ws_bytes": (c_size_t, [c_int]),
    "gp_dist2_knn3": (c_int, [c_int, _P, _P, _P, _P]),
    "gp_image_loss_ws_bytes": (c_size_t, [c_int, c_int, c_int]),
    "gp_image_loss_forward": (c_int, [c_int, c_int, c_int, _P, _P, _P, _P, _P, _P]),
    "gp_image_loss_backward": (c_int, [c_int, c_int, c_int, _P, _P, _P, c_float, c_float, _P, _P, _P]),
    "gp_kpmlp_sizes": (c_int, [POINTER(MlpCfg), POINTER(c_size_t), POINTER(c_size_t)]),
    "gp_kpmlp_forward": (c_int, [POINTER(MlpCfg), _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "gp_kpmlp_backward": (c_int, [POINTER(MlpCfg), _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "gp_adam_step": (c_int, [c_int, POINTER(AdamTensor), c_double, c_double, c_double, _P]),
    "gp_densification_stats": (c_int, [c_int, _P, _P, _P, _P, _P, _P, _P]),
    "gp_allreduce_flag_bytes": (c_size_t, [c_int]),
    "gp_allreduce_sum_f32": (c_int, [c_int, c_int, _P, POINTER(c_void_p), c_size_t, c_size_t, c_size_t, c_int, _P]),
    "gp_allreduce_max_s32": (c_int, [c_int, c_int, _P, POINTER(c_void_p), c_size_t, c_size_t, c_size_t, c_int, _P]),
    "gp_exchange_sync_words": (c_int, []),
    "gp_exchange_segments": (c_int, [c_int, c_int, _P, POINTER(c_void_p), c_size_t, c_int, c_int, POINTER(c_size_t),
                                     POINTER(c_size_t), POINTER(c_int), _P, ctypes.c_uint, _P]),
    "gp_stream_wait_word": (c_int, [_P, ctypes.c_uint, _P]),
    "gp_peer_barrier_launch": (c_int, [c_int, c_int, POINTER(c_void_p), c_size_t, c_int, _P]),
    "gp_copy_async": (c_int, [_P, _P, c_size_t, _P]),
    "gp_add_inplace_f32": (c_int, [_P, _P, c_size_t, _P]),
    "gp_max_inplace_s32": (c_int, [_P, _P, c_size_t, _P]),
    "gp_linear_forward": (c_int, [c_int, c_int, c_int, _P, c_int, _P, c_int, _P, _P, c_int, _P, c_int, _P]),
    "gp_hashgrid_entries": (c_longlong, [c_int, c_int, c_float, c_int]),
    "gp_hashgrid_forward": (c_int, [c_int, c_int, c_int, c_float, c_int, _P, _P, _P, _P, _P]),
    "gp_hashgrid_backward": (c_int, [c_int, c_int, c_int, c_float, c_int, _P, _P, _P, _P, _P]),
    "gp_raster_export_binning": (c_int, [POINTER(RasterCfg), _P, _P, c_longlong, _P, _P, _P]),
}

_lib = None


def lib():
    """Load (once) and return the ctypes handle; raises GpError if the library has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GpError(
                f"{LIB_PATH} not found: the CUDA library has not been built. Run "
                "`python -m gaussianprediction_b200.build` (there is no CPU fallback).")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(code, what=""):
    if code != 0:
        msg = lib().gp_error_string(int(code))
        raise GpError(f"{what}: error {code}: {msg.decode() if msg else '?'}")


def ptr(t):
    """Device address of a torch tensor (None -> NULL).  The tensor must be contiguous."""
    if t is None:
        return None
    assert t.is_contiguous(), "gp_b200: tensor must be contiguous"
    return t.data_ptr()


def stream_ptr():
    import torch
    return torch.cuda.current_stream().cuda_stream

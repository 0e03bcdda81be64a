"""ctypes binding of libna_b200.so (the C ABI declared in include/na_b200.h).

The product path has no CPU or eager-PyTorch fallback: if the library cannot be
loaded, or a call fails, an exception is raised.
"""
import ctypes
import os

import torch

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG_DIR, "libna_b200.so")

NA_MAX_LEVELS = 32

c_f32p = ctypes.POINTER(ctypes.c_float)


class HashGridDesc(ctypes.Structure):
    """Mirror of `na_hashgrid_desc` (include/na_b200.h)."""
    _fields_ = [
        ("n_levels", ctypes.c_uint32),
        ("n_features", ctypes.c_uint32),
        ("total_entries", ctypes.c_uint32),
        ("reserved", ctypes.c_uint32),
        ("scale", ctypes.c_float * NA_MAX_LEVELS),
        ("resolution", ctypes.c_uint32 * NA_MAX_LEVELS),
        ("entries", ctypes.c_uint32 * NA_MAX_LEVELS),
        ("offset", ctypes.c_uint32 * NA_MAX_LEVELS),
        ("hashed", ctypes.c_uint32 * NA_MAX_LEVELS),
    ]


class UnitMap(ctypes.Structure):
    """Mirror of `na_unit_map`."""
    _fields_ = [("vol_min", ctypes.c_float), ("inv_extent", ctypes.c_float)]


class MlpStage(ctypes.Structure):
    """Mirror of `na_mlp_stage`."""
    _fields_ = [("has_epi", ctypes.c_int32), ("act", ctypes.c_int32), ("n_valid", ctypes.c_int32),
                ("epi_to_a", ctypes.c_int32), ("act_param", ctypes.c_float), ("in_col", ctypes.c_int32),
                ("in_cols", ctypes.c_int32), ("has_gemm", ctypes.c_int32), ("k_chunks", ctypes.c_int32),
                ("n_out", ctypes.c_int32), ("accumulate", ctypes.c_int32), ("reserved", ctypes.c_int32),
                ("bias", ctypes.c_void_p), ("out", ctypes.c_void_p), ("out_ld", ctypes.c_int64),
                ("save", ctypes.c_void_p), ("save_ld", ctypes.c_int64), ("inp", ctypes.c_void_p),
                ("in_ld", ctypes.c_int64), ("w_hi", ctypes.c_void_p), ("w_lo", ctypes.c_void_p)]


class NaError(RuntimeError):
    pass


_lib = None


def _declare(lib):
    vp, i64, u32, f32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_uint32, ctypes.c_float
    descp = ctypes.POINTER(HashGridDesc)
    sig = {
        "na_abi_version": (ctypes.c_int, []),
        "na_last_error": (ctypes.c_char_p, []),
        "na_hashgrid_make_desc": (ctypes.c_int, [u32, u32, u32, u32, ctypes.c_double, descp]),
        "na_hashgrid_fwd": (ctypes.c_int, [descp, vp, vp, i64, UnitMap, u32, vp, i64, vp]),
        "na_hashgrid_bwd": (ctypes.c_int, [descp, vp, i64, UnitMap, u32, vp, i64, vp, vp]),
        "na_hashgrid_taps_fwd": (ctypes.c_int, [descp, vp, vp, i64, UnitMap, u32, c_f32p, u32, vp, i64, i64, vp]),
        "na_hashgrid_taps_bwd": (ctypes.c_int, [descp, vp, i64, UnitMap, u32, c_f32p, u32, vp, i64, i64, vp, vp]),
        "na_hashgrid_taps_fwd_strided": (ctypes.c_int, [descp, vp, vp, i64, UnitMap, u32, c_f32p, u32, vp, i64, i64, i64, vp]),
        "na_hashgrid_taps_bwd_strided": (ctypes.c_int, [descp, vp, i64, UnitMap, u32, c_f32p, u32, vp, i64, i64, i64, vp, vp]),
        "na_hashgrid_lattice_fwd_strided": (ctypes.c_int, [descp, vp, vp, vp, vp, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                           UnitMap, u32, vp, i64, i64, vp, i64, vp]),
        "na_lattice_sdf": (ctypes.c_int, [descp, vp, vp, vp, vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, UnitMap, u32,
                                          vp, vp, ctypes.c_int, ctypes.c_int, vp, vp, ctypes.c_float, vp, vp]),
        "na_sdf_head_ws": (ctypes.c_int, [vp, i64, vp, i64, i64, vp, vp, ctypes.c_int, ctypes.c_int, vp, vp, ctypes.c_float,
                                          vp, vp]),
        "na_pack_split_f16": (ctypes.c_int, [vp, i64, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                             ctypes.c_double, vp, ctypes.c_int, vp, vp, vp]),
        "na_ray_setup": (ctypes.c_int, [vp, vp, vp, i64, ctypes.c_int, ctypes.c_int, vp, vp, vp, vp, vp, vp, vp]),
        "na_alpha_weights_fwd": (ctypes.c_int, [vp, i64, ctypes.c_int, vp, vp]),
        "na_alpha_weights_bwd": (ctypes.c_int, [vp, vp, i64, ctypes.c_int, vp, vp]),
        "na_composite_fwd": (ctypes.c_int, [vp, vp, i64, ctypes.c_int, ctypes.c_int, vp, vp]),
        "na_composite_bwd": (ctypes.c_int, [vp, vp, vp, i64, ctypes.c_int, ctypes.c_int, vp, vp, vp]),
        "na_density_alphas_fwd": (ctypes.c_int, [vp, vp, vp, i64, ctypes.c_int, vp, vp]),
        "na_density_alphas_bwd": (ctypes.c_int, [vp, vp, vp, vp, i64, ctypes.c_int, vp, vp]),
        "na_neus_alphas_fwd": (ctypes.c_int, [vp, vp, vp, vp, vp, vp, vp, i64, ctypes.c_int, vp, vp]),
        "na_neus_alphas_bwd": (ctypes.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, i64, ctypes.c_int, vp, vp, vp, vp]),
        "na_sample_pdf": (ctypes.c_int, [vp, vp, vp, i64, ctypes.c_int, ctypes.c_int, vp, vp]),
        "na_neus_upsample": (ctypes.c_int, [vp, vp, vp, f32, i64, ctypes.c_int, ctypes.c_int, vp, vp]),
        "na_merge_sorted": (ctypes.c_int, [vp, vp, vp, vp, i64, ctypes.c_int, ctypes.c_int, vp, vp, vp]),
        "na_neus_render_fwd": (ctypes.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, ctypes.c_int, vp, vp, vp, ctypes.c_int,
                                              i64, vp, vp, vp, vp]),
        "na_neus_render_bwd": (ctypes.c_int, [vp, vp, vp, vp, vp, vp, vp, vp, ctypes.c_int, vp, vp, vp, ctypes.c_int,
                                              i64, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
        "na_mc_scratch_bytes": (ctypes.c_int64, [ctypes.c_int, ctypes.c_int, ctypes.c_int]),
        "na_mc_count": (ctypes.c_int, [vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_double, vp, i64, vp, vp]),
        "na_mc_emit": (ctypes.c_int, [vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_double, vp, i64,
                                      ctypes.c_double, ctypes.POINTER(ctypes.c_double), vp, i64, vp, i64, vp]),
        "na_mc_static_max_cells": (ctypes.c_int64, []),
        "na_mc_run_static": (ctypes.c_int, [vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                            ctypes.c_double, ctypes.POINTER(ctypes.c_double), vp, i64, vp, i64, vp, vp]),
        "na_adamw_step": (ctypes.c_int, [vp, vp, vp, vp, i64, f32, f32, f32, f32, f32, i64, f32, ctypes.c_int, vp]),
        "na_adamw_step_multimem": (ctypes.c_int, [vp, vp, vp, vp, vp, i64, i64, f32, f32, f32, f32, f32, i64, f32,
                                                  ctypes.c_int, ctypes.c_int, vp]),
        "na_linear_wgrad": (ctypes.c_int, [vp, i64, vp, i64, i64, ctypes.c_int, ctypes.c_int, vp, vp, i64, vp]),
        "na_scale_inplace": (ctypes.c_int, [vp, i64, f32, vp]),
        "na_weight_norm_fwd": (ctypes.c_int, [vp, vp, ctypes.c_int, ctypes.c_int, vp, vp, vp]),
        "na_weight_norm_bwd": (ctypes.c_int, [vp, vp, vp, vp, ctypes.c_int, ctypes.c_int, vp, vp, vp]),
        "na_sdf_mlp_fwd": (ctypes.c_int, [vp, i64, i64, ctypes.c_int, ctypes.c_int, ctypes.c_int, vp, vp, vp, ctypes.c_int,
                                          vp, vp, vp, vp, vp, f32, vp, vp, i64, vp, i64, vp]),
        "na_sdf_head_fwd_planes": (ctypes.c_int, [vp, i64, i64, ctypes.c_int, ctypes.c_int, vp, vp, vp, ctypes.c_int, vp, vp,
                                                  f32, vp, vp, i64, vp]),
        "na_sdf_mlp_bwd": (ctypes.c_int, [i64, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, vp, vp, vp, vp, i64, vp,
                                          vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, i64, vp]),
        "na_mlp_chain_fwd": (ctypes.c_int, [ctypes.POINTER(MlpStage), ctypes.c_int, i64, vp]),
        "na_rgb_input_fwd": (ctypes.c_int, [vp, vp, vp, vp, i64, vp, i64, i64, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                            ctypes.c_int, vp, i64, vp]),
        "na_rgb_input_bwd": (ctypes.c_int, [vp, vp, i64, ctypes.c_int, i64, vp, vp]),
        "na_tap_combine_fwd": (ctypes.c_int, [vp, vp, i64, ctypes.c_int, f32, f32, vp, vp, vp]),
        "na_tap_combine_bwd": (ctypes.c_int, [vp, vp, i64, ctypes.c_int, f32, f32, vp, vp, vp]),
        "na_sdf_losses_fwd": (ctypes.c_int, [vp, vp, vp, i64, ctypes.c_int, vp, vp]),
        "na_sdf_losses_bwd": (ctypes.c_int, [vp, vp, vp, i64, ctypes.c_int, vp, vp, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return sig


def lib():
    """Load (once) and return the ctypes handle; raises if the library is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NaError(
                f"{LIB_PATH} not found: build it with `python -m neuralangelo_b200.build` "
                "(there is no CPU/eager fallback for the CUDA path)")
        handle = ctypes.CDLL(LIB_PATH)
        _declare(handle)
        _lib = handle
    return _lib


def exported_signatures():
    """Names bound from the library (used by the ABI test)."""
    return sorted(_declare(lib()).keys())


def check(rc, what):
    if rc != 0:
        msg = lib().na_last_error().decode("utf-8", "replace")
        raise NaError(f"{what} failed (code {rc}): {msg}")


def current_stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def f32c(t):
    """fp32 contiguous view/copy of a tensor (kernels take dense fp32 rows)."""
    if t is None:
        return None
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise NaError("libna_b200 kernels need CUDA tensors; there is no CPU fallback")

"""ctypes binding of ``include/ivit_b200.h`` (the C ABI of the sm_100a kernels).

There is no CPU fallback: if the library is missing, or a kernel is asked to run on
anything but CUDA tensors, this raises.  torch is used only for device memory and the
current stream.
"""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

_LIB = None

EPI_ACC32, EPI_REQUANT, EPI_RESIDUAL, EPI_LOGITS = 0, 1, 2, 3


class GemmEpilogue(C.Structure):
    _fields_ = [
        ("epilogue", C.c_int),
        ("bias", C.c_void_p),
        ("mult", C.c_void_p),
        ("s_in", C.c_void_p),
        ("qmin", C.c_int), ("qmax", C.c_int),
        ("mult2", C.c_double),
        ("mult_id", C.c_double),
        ("qmin2", C.c_int), ("qmax2", C.c_int),
        ("identity", C.c_void_p),
        ("ld_id", C.c_int64),
        ("scale_out", C.c_void_p),
        ("group", C.c_int), ("out_group_stride", C.c_int), ("out_row_off", C.c_int), ("id_broadcast", C.c_int),
        ("force_slow", C.c_int),
        ("z_bound_ok", C.c_int),
        ("out_bits", C.c_int), ("id_bits", C.c_int),
        ("gelu_lut", C.c_void_p), ("gelu_out", C.c_void_p),
    ]


class AttentionParams(C.Structure):
    _fields_ = [
        ("kind", C.c_int),
        ("output_bit", C.c_int),
        ("mult_attn", C.c_double),
        ("qmin_attn", C.c_int), ("qmax_attn", C.c_int),
        ("s_attn", C.c_float),
        ("mult_out", C.c_double),
        ("qmin_out", C.c_int), ("qmax_out", C.c_int),
        ("mult_exp16", C.c_double),
        ("s_exp16", C.c_float),
        ("exp_table", C.c_void_p),
        ("exp_table_1d", C.c_void_p),
        ("s_pv", C.c_float),
        ("sat_k", C.c_int),
        ("sat_e", C.c_float),
    ]


class WindowAttentionParams(C.Structure):
    _fields_ = [
        ("mult_attn", C.c_double),
        ("qmin_attn", C.c_int), ("qmax_attn", C.c_int),
        ("mult2", C.c_double), ("mult_bias", C.c_double),
        ("qmin2", C.c_int), ("qmax2", C.c_int),
        ("s2", C.c_float),
        ("bias_codes", C.c_void_p),
        ("output_bit", C.c_int),
        ("mult_out", C.c_double),
        ("qmin_out", C.c_int), ("qmax_out", C.c_int),
        ("exp_table", C.c_void_p),
        ("mask_clamps", C.c_int),
        ("bias_terms", C.c_void_p),
        ("windowed", C.c_int),
        ("sat_k", C.c_int), ("sat_e", C.c_float),
    ]


_SIGS = {
    "ivit_abi_version": (C.c_int, []),
    "ivit_set_pdl": (C.c_int, [C.c_int]),
    "ivit_device_sm": (C.c_int, [C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "ivit_last_error": (C.c_char_p, []),
    "ivit_xeff_fma_ok": (C.c_int, [C.c_float, C.c_int]),
    "ivit_quantize_f32_i8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_float, C.c_int64, C.c_void_p]),
    "ivit_quantize_patches_f32_i8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_int,
                                               C.c_int, C.c_int, C.c_void_p]),
    "ivit_quantize_patches_u8_i8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                              C.c_int, C.c_void_p]),
    "ivit_layernorm_requant_i8": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                            C.c_int, C.c_void_p]),
    "ivit_gelu_lut_build": (C.c_int, [C.c_int, C.c_void_p, C.c_float, C.c_int, C.c_double, C.c_int, C.c_int,
                                      C.POINTER(C.c_float), C.c_void_p]),
    "ivit_gelu_ivit_raw_table": (C.c_int, [C.c_void_p, C.c_float, C.c_int, C.c_void_p]),
    "ivit_gelu_lut_apply_i8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "ivit_gelu_i8_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_int,
                                       C.c_double, C.c_int, C.c_int, C.POINTER(C.c_float), C.c_void_p]),
    "ivit_requant_i32_i8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_double,
                                      C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "ivit_requant_i8_i8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_double,
                                     C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "ivit_dequant_i8_f32": (C.c_int, [C.c_void_p, C.c_float, C.c_void_p, C.c_int64, C.c_void_p]),
    "ivit_fill_cls_rows_i8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "ivit_layernorm_requant_i16_i8": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                                C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                                C.c_int, C.c_void_p]),
    "ivit_quantize_f32_codes": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int64, C.c_int, C.c_int,
                                          C.c_void_p]),
    "ivit_layernorm_requant_strided": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_int, C.c_int,
                                                 C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                                 C.c_float, C.c_int, C.c_int, C.c_void_p]),
    "ivit_window_attention_i8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                           C.c_int, C.c_int, C.POINTER(WindowAttentionParams), C.c_void_p]),
    "ivit_window_attention_set_variant": (C.c_int, [C.c_int]),
    "ivit_patch_merge_gather": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "ivit_avgpool_requant_i8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_double,
                                          C.c_int, C.c_int, C.c_void_p]),
    "ivit_fill_cls_rows_i16": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "ivit_requant_codes": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                     C.c_int64, C.c_double, C.c_void_p, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int,
                                     C.c_void_p]),
    "ivit_gemm_i8": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                               C.c_int64, C.POINTER(GemmEpilogue), C.c_void_p]),
    "ivit_gemm_set_layout": (C.c_int, [C.c_int]),
    "ivit_attention_set_threads_per_row": (C.c_int, [C.c_int]),
    "ivit_layernorm_set_variant": (C.c_int, [C.c_int]),
    "ivit_attention_set_variant": (C.c_int, [C.c_int]),
    "ivit_gemm_i8_simple": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                      C.c_int64, C.POINTER(GemmEpilogue), C.c_void_p]),
    "ivit_attention_i8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.POINTER(AttentionParams), C.c_void_p]),
    "ivit_softmax_table_build": (C.c_int, [C.POINTER(AttentionParams), C.c_void_p, C.c_void_p]),
    "ivit_softmax_table_collapse": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "ivit_attention_probs_i8": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                          C.POINTER(AttentionParams), C.c_void_p]),
}

EXPORTS = tuple(_SIGS)


def lib_path() -> str:
    return _build.LIB_PATH


def load(build_if_missing: bool = True):
    """dlopen the in-tree library (building it first if absent and nvcc is available)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.isfile(path):
        if not build_if_missing:
            raise RuntimeError(f"{path} missing: run `python -m ivit_b200.build` (no CPU fallback exists)")
        _build.build()
    lib = C.CDLL(path)
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)   # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    if os.environ.get("IVIT_PDL") is not None:          # A/B switch for programmatic dependent launch
        lib.ivit_set_pdl(int(os.environ["IVIT_PDL"] != "0"))
    _LIB = lib
    return lib


class IvitKernelError(RuntimeError):
    pass


def check(rc: int, what: str):
    if rc != 0:
        msg = load().ivit_last_error()
        raise IvitKernelError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")


def ptr(t) -> int:
    """Device pointer of a CUDA tensor (None -> NULL).  Refuses CPU tensors: no CPU path."""
    if t is None:
        return None
    if not t.is_cuda:
        raise IvitKernelError("ivit_b200 kernels run on CUDA tensors only (there is no CPU fallback)")
    return t.data_ptr()


def stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream

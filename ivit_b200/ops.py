"""Tensor-level wrappers over the C ABI: allocate outputs with torch, pass raw pointers.

Every function enqueues on torch's current CUDA stream and returns device tensors of integer
codes.  No function here has a CPU implementation.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _cabi
from ._cabi import AttentionParams, GemmEpilogue, check, ptr, stream


def dev(a, dtype, device) -> torch.Tensor:
    """numpy host constant -> device tensor (done once per frozen module, then cached by the caller)."""
    return torch.from_numpy(np.ascontiguousarray(a)).to(dtype=dtype).to(device)


def quantize(x: torch.Tensor, inv_scale: float, bits: int = 8) -> torch.Tensor:
    """QuantAct on a float tensor: int8 codes for bits <= 8, int16 codes for 9..16 bits."""
    x = x.contiguous()
    lib = _cabi.load()
    if bits <= 8:
        out = torch.empty(x.shape, dtype=torch.int8, device=x.device)
        if bits == 8:
            check(lib.ivit_quantize_f32_i8(ptr(x), ptr(out), C.c_float(inv_scale), x.numel(), stream()),
                  "ivit_quantize_f32_i8")
            return out
    else:
        out = torch.empty(x.shape, dtype=torch.int16, device=x.device)
    check(lib.ivit_quantize_f32_codes(ptr(x), ptr(out), 16 if bits > 8 else 8, C.c_float(inv_scale), x.numel(),
                                      -(2 ** (bits - 1)), 2 ** (bits - 1) - 1, stream()), "ivit_quantize_f32_codes")
    return out


def quantize_patches(x: torch.Tensor, inv_scale: float, patch: int, out: torch.Tensor | None = None) -> torch.Tensor:
    x = x.contiguous()
    B, Cin, H, W = x.shape
    if out is None:
        out = torch.empty((B * (H // patch) * (W // patch), Cin * patch * patch), dtype=torch.int8, device=x.device)
    assert out.is_contiguous() and out.numel() == x.numel() and out.dtype == torch.int8
    check(_cabi.load().ivit_quantize_patches_f32_i8(ptr(x), ptr(out), C.c_float(inv_scale), B, Cin, H, W, patch,
                                                    stream()), "ivit_quantize_patches_f32_i8")
    return out


def quantize_patches_u8(x: torch.Tensor, lut: torch.Tensor, patch: int, out: torch.Tensor | None = None) -> torch.Tensor:
    """uint8 images [B,Cin,H,W] -> int8 patch matrix through the [Cin][256] code table of `input_code_lut`."""
    x = x.contiguous()
    assert x.dtype == torch.uint8 and lut.dtype == torch.int8
    B, Cin, H, W = x.shape
    assert lut.numel() == Cin * 256
    if out is None:
        out = torch.empty((B * (H // patch) * (W // patch), Cin * patch * patch), dtype=torch.int8, device=x.device)
    assert out.is_contiguous() and out.numel() == x.numel() and out.dtype == torch.int8
    check(_cabi.load().ivit_quantize_patches_u8_i8(ptr(x), ptr(out), ptr(lut), B, Cin, H, W, patch, stream()),
          "ivit_quantize_patches_u8_i8")
    return out


def input_code_lut(inv_scale: float, mean, std) -> np.ndarray:
    """int8 [Cin][256]: qact_input(Normalize(ToTensor(u))) for every uint8 pixel value u, in float32 and in the
    order torchvision applies them: u / 255, then (x - mean[c]) / std[c], then clamp(round(fl32(1/s) * x))
    (quant_utils.py:49,93)."""
    f32 = np.float32
    u = np.arange(256, dtype=f32) / f32(255.0)
    mean, std = np.asarray(mean, dtype=f32), np.asarray(std, dtype=f32)
    xn = ((u[None, :] - mean[:, None]).astype(f32) / std[:, None]).astype(f32)
    q = np.rint((f32(inv_scale) * xn).astype(f32))
    return np.clip(q, -128, 127).astype(np.int8)


def layernorm_requant(kind: int, x: torch.Tensor, s_in: float, bias_int: torch.Tensor, sf: torch.Tensor,
                      mult: torch.Tensor, qmin: int, qmax: int, shift: float = 0.0,
                      use_int_sqrt: bool = False, out: torch.Tensor | None = None) -> torch.Tensor:
    """x: int8 (or int16: 16-bit residual stream) [..., C] (last dim contiguous, uniform row stride), e.g. a
    [B, C] view of the class rows."""
    assert x.dtype in (torch.int8, torch.int16) and x.stride(-1) == 1
    Cdim = x.shape[-1]
    rows = x.numel() // Cdim
    ldx = x.stride(-2) if x.dim() >= 2 else Cdim
    assert x.dim() <= 2 or x.is_contiguous()
    if out is None:
        out = torch.empty(x.shape, dtype=torch.int8, device=x.device)
    lib = _cabi.load()
    fn = lib.ivit_layernorm_requant_i8 if x.dtype == torch.int8 else lib.ivit_layernorm_requant_i16_i8
    check(fn(kind, ptr(x), ldx, ptr(out), rows, Cdim, C.c_float(s_in), ptr(bias_int), ptr(sf), ptr(mult), qmin, qmax,
             C.c_float(shift), int(use_int_sqrt), stream()), "ivit_layernorm_requant")
    return out


def gelu_lut_build(kind: int, s_in: float, output_bit: int, mult_out: float, qmin: int, qmax: int, device):
    lut = torch.empty((256, 256), dtype=torch.int8, device=device)
    s_out = C.c_float(0.0)
    check(_cabi.load().ivit_gelu_lut_build(kind, ptr(lut), C.c_float(s_in), output_bit, C.c_double(mult_out), qmin,
                                           qmax, C.byref(s_out), stream()), "ivit_gelu_lut_build")
    return lut, np.float32(s_out.value)


def gelu_ivit_raw_table(s_in: float, output_bit: int, device) -> torch.Tensor:
    out = torch.empty((256, 256), dtype=torch.float32, device=device)
    check(_cabi.load().ivit_gelu_ivit_raw_table(ptr(out), C.c_float(s_in), output_bit, stream()),
          "ivit_gelu_ivit_raw_table")
    return out


def gelu_lut_apply(x: torch.Tensor, lut: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    assert x.dtype == torch.int8 and x.is_contiguous()
    Cdim = x.shape[-1]
    if out is None:
        out = torch.empty_like(x)
    check(_cabi.load().ivit_gelu_lut_apply_i8(ptr(x), ptr(out), ptr(lut), x.numel() // Cdim, Cdim, stream()),
          "ivit_gelu_lut_apply_i8")
    return out


_BITS = {torch.int8: 8, torch.int16: 16, torch.int32: 32}


def requant(z: torch.Tensor, s_in: torch.Tensor, mult: torch.Tensor, qmin: int, qmax: int,
            identity: torch.Tensor | None = None, mult_id: float = 0.0, out: torch.Tensor | None = None,
            out_bits: int = 8) -> torch.Tensor:
    """Stand-alone QuantAct on int32 / int16 / int8 codes [..., C]; identity int8 or int16 [id_rows, C]
    broadcast over rows; int8 codes out, or int16 for a 16-bit location (out_bits = 16)."""
    assert z.dtype in _BITS and z.is_contiguous()
    Cdim = z.shape[-1]
    rows = z.numel() // Cdim
    odt = torch.int16 if out_bits == 16 else torch.int8
    if out is None:
        out = torch.empty(z.shape, dtype=odt, device=z.device)
    assert out.dtype == odt and out.is_contiguous()
    id_rows = 0
    if identity is not None:
        assert identity.dtype in (torch.int8, torch.int16) and identity.is_contiguous() and identity.shape[-1] == Cdim
        id_rows = identity.numel() // Cdim
    lib = _cabi.load()
    if z.dtype != torch.int16 and out_bits == 8 and (identity is None or identity.dtype == torch.int8):
        fn = lib.ivit_requant_i32_i8 if z.dtype == torch.int32 else lib.ivit_requant_i8_i8
        check(fn(ptr(z), ptr(s_in), ptr(mult), int(mult.numel() > 1), ptr(identity), id_rows, C.c_double(mult_id),
                 ptr(out), rows, Cdim, qmin, qmax, stream()), "ivit_requant")
        return out
    check(lib.ivit_requant_codes(ptr(z), _BITS[z.dtype], ptr(s_in), ptr(mult), int(mult.numel() > 1), ptr(identity),
                                 _BITS[identity.dtype] if identity is not None else 8, id_rows, C.c_double(mult_id),
                                 ptr(out), out_bits, rows, Cdim, qmin, qmax, stream()), "ivit_requant_codes")
    return out


def requant_i32(z, s_in, mult, qmin, qmax, identity=None, mult_id=0.0):
    return requant(z, s_in, mult, qmin, qmax, identity, mult_id)


def dequant(q: torch.Tensor, scale: float) -> torch.Tensor:
    q = q.contiguous()
    out = torch.empty(q.shape, dtype=torch.float32, device=q.device)
    check(_cabi.load().ivit_dequant_i8_f32(ptr(q), C.c_float(scale), ptr(out), q.numel(), stream()),
          "ivit_dequant_i8_f32")
    return out


def fill_cls_rows(out: torch.Tensor, cls_row: torch.Tensor):
    B, N, Cdim = out.shape
    assert out.dtype == cls_row.dtype and out.is_contiguous()
    lib = _cabi.load()
    fn = lib.ivit_fill_cls_rows_i16 if out.dtype == torch.int16 else lib.ivit_fill_cls_rows_i8
    check(fn(ptr(out), ptr(cls_row), B, N, Cdim, stream()), "ivit_fill_cls_rows")


def gemm(A: torch.Tensor, W: torch.Tensor, epilogue: int, *, bias=None, mult=None, s_in=None, qmin=-128, qmax=127,
         mult2=0.0, mult_id=0.0, qmin2=-128, qmax2=127, identity=None, scale_out=None, out=None, out_rows=None,
         group=0, out_group_stride=0, out_row_off=0, id_broadcast=False, force_slow=False,
         z_bound_ok=False, simple=False, out_bits=8, gelu_lut=None, gelu_out=None) -> torch.Tensor:
    """C = A[M,K] @ W[N,K]^T with a fused epilogue (see include/ivit_b200.h).  REQUANT / RESIDUAL store
    int16 codes when `out` is int16 (or out_bits = 16); an int16 `identity` is the 16-bit residual stream."""
    assert A.dtype == torch.int8 and W.dtype == torch.int8 and A.dim() == 2 and W.dim() == 2
    assert A.stride(1) == 1 and W.is_contiguous()
    M, K = A.shape
    N = W.shape[0]
    if out is not None and out.dtype == torch.int16:
        out_bits = 16
    odt = {_cabi.EPI_ACC32: torch.int32, _cabi.EPI_LOGITS: torch.float32}.get(
        epilogue, torch.int16 if out_bits == 16 else torch.int8)
    if out is None:
        out = torch.empty((out_rows if out_rows is not None else M, N), dtype=odt, device=A.device)
    assert out.dtype == odt and out.stride(-1) == 1
    e = GemmEpilogue()
    e.epilogue = epilogue
    e.bias, e.mult, e.s_in = ptr(bias), ptr(mult), ptr(s_in)
    e.qmin, e.qmax, e.mult2, e.mult_id, e.qmin2, e.qmax2 = qmin, qmax, mult2, mult_id, qmin2, qmax2
    e.identity = ptr(identity)
    e.ld_id = identity.stride(-2) if identity is not None else 0
    e.scale_out = ptr(scale_out)
    e.group, e.out_group_stride, e.out_row_off, e.id_broadcast = group, out_group_stride, out_row_off, int(id_broadcast)
    e.force_slow = int(force_slow)
    e.z_bound_ok = int(z_bound_ok)
    e.out_bits = 16 if odt == torch.int16 else 8
    e.id_bits = 16 if (identity is not None and identity.dtype == torch.int16) else 8
    if identity is not None:
        assert identity.dtype in (torch.int8, torch.int16) and identity.stride(-1) == 1
    if gelu_lut is not None:           # REQUANT + the GELU table of the following mlp.act / mlp.qact1, one launch
        assert epilogue == _cabi.EPI_REQUANT and not simple and gelu_out is not None and gelu_out.dtype == torch.int8
        assert gelu_out.shape == out.shape and gelu_out.stride() == out.stride() and gelu_lut.numel() == 65536
        e.gelu_lut, e.gelu_out = ptr(gelu_lut), ptr(gelu_out)
    lib = _cabi.load()
    fn = lib.ivit_gemm_i8_simple if simple else lib.ivit_gemm_i8
    ldo = out.stride(-2) if out.dim() >= 2 else N
    check(fn(ptr(A), A.stride(0), ptr(W), M, N, K, ptr(out), ldo, C.byref(e), stream()),
          "ivit_gemm_i8_simple" if simple else "ivit_gemm_i8")
    return out


def set_pdl(on: bool) -> int:
    """Programmatic dependent launch of the hot kernels.  Returns the old value."""
    return _cabi.load().ivit_set_pdl(int(bool(on)))


def set_layernorm_variant(variant: int) -> int:
    """0: packed-lane I-ViT LayerNorm kernel where C % 128 == 0; 1: strided-lane kernel always."""
    return _cabi.load().ivit_layernorm_set_variant(int(variant))


def set_attention_variant(variant: int) -> int:
    """0 / 1: the thread-per-row attention kernel (default); 2: an 8-bit softmax output runs the row-per-warp-step kernel."""
    return _cabi.load().ivit_attention_set_variant(int(variant))


def set_attention_threads_per_row(tpr: int) -> int:
    """2 or 4 threads per query row in the attention kernel.  Returns the old value."""
    return _cabi.load().ivit_attention_set_threads_per_row(int(tpr))


def set_gemm_layout(layout: int) -> int:
    """0: ivit_gemm_i8 picks its kernel per call; 1: tokens-on-lanes kernel only.  Returns the old value."""
    return _cabi.load().ivit_gemm_set_layout(int(layout))


def _table_saturation(table: torch.Tensor):
    """(sat_k, sat_e) of a [256][256] (row-max code, code) table: every valid entry with row-max code - code >=
    sat_k equals sat_e bit for bit (IntSoftmax clamps its argument, ivit_modules.py:155 / ibert_modules.py:288).
    (0, 0.0) when the table has no such tail."""
    t = table.view(256, 256).view(torch.int32)
    idx = torch.arange(256, device=table.device)
    k = idx[:, None] - idx[None, :]
    e_sat = t[255, 0]
    differs = (k >= 0) & (t != e_sat)
    sat_k = int(k[differs].max().item()) + 1 if bool(differs.any()) else 1
    if sat_k >= 255:
        return 0, 0.0
    return sat_k, float(table.view(256, 256)[255, 0].item())


def softmax_table_build(params: AttentionParams, device, collapse: bool = True, host_table=None) -> torch.Tensor:
    """Tabulate the softmax exp term over (row-max code, code) and attach it to `params`; when the
    table depends on the code difference only (checked on the device, bit for bit) also attach the
    256-entry form the kernel keeps in shared memory.  The returned tensor owns the memory: keep
    it alive as long as `params` is used.  Freeze-time call (reads one flag back)."""
    lib = _cabi.load()
    buf = torch.empty(65536 + 256, dtype=torch.float32, device=device)
    params.exp_table = None
    params.exp_table_1d = None
    if host_table is not None:        # kind 2 (ppoly): the term table comes from the module's stored coefficients
        buf[:65536].copy_(torch.from_numpy(np.ascontiguousarray(host_table, dtype=np.float32)).reshape(-1))
    else:
        check(lib.ivit_softmax_table_build(C.byref(params), ptr(buf), stream()), "ivit_softmax_table_build")
    params.exp_table = buf.data_ptr()
    params.sat_k, params.sat_e = _table_saturation(buf[:65536])
    if collapse:
        flag = torch.zeros(1, dtype=torch.int32, device=device)
        check(lib.ivit_softmax_table_collapse(ptr(buf), buf.data_ptr() + 65536 * 4, ptr(flag), stream()),
              "ivit_softmax_table_collapse")
        if int(flag.item()) == 1:
            params.exp_table_1d = buf.data_ptr() + 65536 * 4
    return buf


def attention(qkv: torch.Tensor, B: int, N: int, H: int, hd: int, params: AttentionParams,
              want_probs: bool = False, out: torch.Tensor | None = None):
    """qkv int8 [B,N,3*H*hd] -> out int8 [B,N,H*hd] (and optionally the softmax codes [B,H,N,N]: uint8, or
    int32 holding the uint16 codes of a 16-bit softmax)."""
    assert qkv.dtype == torch.int8 and qkv.is_contiguous()
    lib = _cabi.load()
    if out is None:
        out = torch.empty((B, N, H * hd), dtype=torch.int8, device=qkv.device)
    assert out.is_contiguous() and out.dtype == torch.int8 and out.numel() == B * N * H * hd
    check(lib.ivit_attention_i8(ptr(qkv), ptr(out), B, N, H, hd, C.byref(params), stream()), "ivit_attention_i8")
    if not want_probs:
        return out
    wide = params.output_bit > 8
    probs = torch.empty((B, H, N, N), dtype=torch.int16 if wide else torch.uint8, device=qkv.device)
    check(lib.ivit_attention_probs_i8(ptr(qkv), ptr(probs), B, N, H, hd, C.byref(params), stream()),
          "ivit_attention_probs_i8")
    if wide:
        probs = probs.to(torch.int32) & 0xffff
    return out, probs


def layernorm_requant_strided(kind: int, x: torch.Tensor, tokens_per_image: int, s_in: float, bias_int: torch.Tensor,
                              sf: torch.Tensor, mult: torch.Tensor, qmin: int, qmax: int, shift: float = 0.0,
                              use_int_sqrt: bool = False, out: torch.Tensor | None = None) -> torch.Tensor:
    """LayerNorm + QuantAct whose reference input is channel-strided (Swin stage 0): x int8 / int16 [rows, C]."""
    assert x.dtype in (torch.int8, torch.int16) and x.dim() == 2 and x.stride(-1) == 1
    rows, Cdim = x.shape
    if out is None:
        out = torch.empty(x.shape, dtype=torch.int8, device=x.device)
    check(_cabi.load().ivit_layernorm_requant_strided(kind, ptr(x), _BITS[x.dtype], x.stride(0), ptr(out), rows, Cdim,
                                                      C.c_float(s_in), ptr(bias_int), ptr(sf), ptr(mult), qmin, qmax,
                                                      C.c_float(shift), int(use_int_sqrt), int(tokens_per_image),
                                                      stream()), "ivit_layernorm_requant_strided")
    return out


def window_attention(qkv: torch.Tensor, B: int, res: int, ws: int, shift: int, nH: int, hd: int, params,
                     out: torch.Tensor | None = None, want_probs: bool = False):
    """Swin window attention on qkv int8 [B, res*res, 3*nH*hd] in image token order (roll / window
    partition / reverse are index maps inside the kernel) -> out int8 [B, res*res, nH*hd]."""
    assert qkv.dtype == torch.int8 and qkv.is_contiguous() and qkv.numel() == B * res * res * 3 * nH * hd
    if out is None:
        out = torch.empty((B, res * res, nH * hd), dtype=torch.int8, device=qkv.device)
    assert out.dtype == torch.int8 and out.is_contiguous() and out.numel() == B * res * res * nH * hd
    probs = None
    if want_probs:
        N = ws * ws
        probs = torch.empty((B * (res // ws) ** 2, nH, N, N), dtype=torch.uint8, device=qkv.device)
    check(_cabi.load().ivit_window_attention_i8(ptr(qkv), ptr(out), ptr(probs), B, res, ws, shift, nH, hd,
                                                C.byref(params), stream()), "ivit_window_attention_i8")
    return (out, probs) if want_probs else out


def set_window_attention_variant(variant: int) -> int:
    """0: tcgen05 products where head_dim = 32; 1: dp4a kernel always.  Returns the old value."""
    return _cabi.load().ivit_window_attention_set_variant(int(variant))


def patch_merge_gather(x: torch.Tensor, B: int, res: int, Cdim: int, out: torch.Tensor | None = None) -> torch.Tensor:
    """[B, res, res, C] -> [B, (res/2)^2, 4C] (PatchMerging's strided slices + cat)."""
    assert x.dtype in (torch.int8, torch.int16) and x.is_contiguous() and x.numel() == B * res * res * Cdim
    if out is None:
        out = torch.empty((B * (res // 2) ** 2, 4 * Cdim), dtype=x.dtype, device=x.device)
    assert out.dtype == x.dtype and out.is_contiguous() and out.numel() == x.numel()
    check(_cabi.load().ivit_patch_merge_gather(ptr(x), ptr(out), _BITS[x.dtype], B, res, Cdim, stream()),
          "ivit_patch_merge_gather")
    return out


def avgpool_requant(x: torch.Tensor, B: int, L: int, Cdim: int, s_in: float, mult: float, qmin: int, qmax: int,
                    out: torch.Tensor | None = None) -> torch.Tensor:
    """Token average pool of the float carrier + QuantAct: x int8 [B, L, C] -> int8 [B, C]."""
    assert x.dtype == torch.int8 and x.is_contiguous() and x.numel() == B * L * Cdim
    if out is None:
        out = torch.empty((B, Cdim), dtype=torch.int8, device=x.device)
    check(_cabi.load().ivit_avgpool_requant_i8(ptr(x), ptr(out), B, L, Cdim, C.c_float(s_in), C.c_double(mult), qmin,
                                               qmax, stream()), "ivit_avgpool_requant_i8")
    return out

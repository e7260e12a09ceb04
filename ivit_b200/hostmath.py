"""Host-side constants of the frozen integer path ("plan building").

After ``freeze_model`` every scale, dyadic multiplier, integer weight and integer bias is a
function of (weights, ``x_min``/``x_max`` buffers) only, but the reference recomputes all of it
on every forward -- including a device->host sync and a python ``decimal`` loop per QuantAct
(quant_utils.py:151-175).  Here it is computed ONCE, on the host, in numpy float32/float64
with the reference's exact operation order, and cached on the module.

This is product code (it never imports ``oracle/``); numpy float32 scalar arithmetic is
IEEE-754, i.e. bit-identical to the torch CPU ops it restates.
"""
from __future__ import annotations

import decimal
from decimal import Decimal

import numpy as np

f32 = np.float32
f64 = np.float64
_EPS = np.finfo(np.float32).eps


def as_f32(a) -> np.ndarray:
    return np.asarray(a, dtype=np.float32)


def sym_scale(num_bits: int, x_min, x_max) -> np.ndarray:
    """quant_utils.py:52-70: max(-min, max) / (2^(k-1)-1), clamped to float32 eps."""
    n = 2 ** (num_bits - 1) - 1
    s = np.maximum(-as_f32(x_min), as_f32(x_max)) / f32(n)
    return np.maximum(s, f32(_EPS)).astype(f32)


def quantize_tensor(x, k: int, scale_bcast) -> np.ndarray:
    """quant_utils.py:49 + :93: clamp(round(fl32(1/s) * x), -2^(k-1), 2^(k-1)-1) as float32."""
    n = 2 ** (k - 1) - 1
    q = np.rint((f32(1.0) / as_f32(scale_bcast)) * as_f32(x))
    return np.clip(q, f32(-n - 1), f32(n)).astype(f32)


def frexp_dyadic(ratio):
    """quant_utils.py:151-175 batch_frexp on float64 ratios.

    Returns (m, e): m = ROUND_HALF_UP(mantissa * 2^31) (away from zero; may equal 2^31 or be
    negative), e = 31 - exponent, so that ratio ~= m / 2^e."""
    r = np.asarray(ratio, dtype=f64).reshape(-1)
    m0, e0 = np.frexp(r)
    m = np.array([int(Decimal(float(v) * (2 ** 31)).quantize(Decimal("1"), rounding=decimal.ROUND_HALF_UP))
                  for v in m0], dtype=np.int64)
    return m, (31 - e0).astype(np.int64)


def dyadic_mult(s_in, s_out) -> np.ndarray:
    """The float64 multiplier m * 2^-e of fixedpoint_mul (quant_utils.py:221-230) for
    ratio = float64(s_in) / float64(float32(s_out)).  Exact: m has <= 32 significant bits."""
    a = as_f32(s_in).astype(f64)
    b = as_f32(s_out).astype(f64)
    m, e = frexp_dyadic(a / b)
    return np.ldexp(m.astype(f64), (-e).astype(np.int64)).reshape(np.shape(a / b))


def qrange(bits: int):
    n = 2 ** (bits - 1) - 1
    return -n - 1, n


def linear_params(weight, bias, s_x, weight_bit: int = 8, bias_bit: int = 32):
    """quant_modules.py:202-220 (QuantLinear) / :486-504 (QuantConv2d).

    Returns dict: s_w [out] f32, w_int int8 [out, K] (K = prod of remaining dims), b_scale [out] f32,
    b_int int32 [out] (saturated to int32), b_int_f32 (the reference's float value)."""
    w = as_f32(weight)
    w2 = w.reshape(w.shape[0], -1)
    s_w = sym_scale(weight_bit, w2.min(axis=1), w2.max(axis=1))
    w_int = quantize_tensor(w2, weight_bit, s_w.reshape(-1, 1))
    b_scale = (s_w * as_f32(s_x).reshape(-1)[0]).astype(f32)
    out = {"s_w": s_w, "w_int": w_int.astype(np.int8), "w_int_f32": w_int, "b_scale": b_scale}
    if bias is not None:
        b = quantize_tensor(as_f32(bias), bias_bit, b_scale)
        out["b_int_f32"] = b
        out["b_int"] = np.clip(b.astype(f64), -2 ** 31, 2 ** 31 - 1).astype(np.int64).astype(np.int32)
    else:
        out["b_int_f32"] = None
        out["b_int"] = np.zeros(w.shape[0], dtype=np.int32)
    return out


def layernorm_params(weight, bias, C: int):
    """ivit_modules.py:53-62 / ibert_modules.py:148-155: per-channel output scale and integer bias."""
    dim_sqrt = np.sqrt(f32(C)).astype(f32)
    sf = (dim_sqrt / f32(2 ** 30)).astype(f32)
    b = as_f32(bias) / as_f32(weight)
    bias_int = np.floor(b / sf).astype(f32)
    sf_c = (sf * as_f32(weight)).astype(f32)
    return {"sf": sf_c, "bias_int": bias_int}


def recover_codes(x, scale) -> np.ndarray:
    """round(x / s): the integer a QuantAct recovers from a float32 carrier (quant_utils.py:220)."""
    return np.rint(as_f32(x) / as_f32(scale))


def requant_host(z, s_in, s_out, bits: int, identity_codes=None, s_id=None) -> np.ndarray:
    """fixedpoint_mul on the host for small, batch-independent tensors (class token row,
    positional embedding); same float64 sequence as quant_utils.py:220-251."""
    lo, hi = qrange(bits)
    z = np.asarray(z, dtype=f64)
    out = np.rint(z * dyadic_mult(s_in, s_out))
    if identity_codes is not None:
        out = out + np.rint(np.asarray(identity_codes, dtype=f64) * dyadic_mult(s_id, s_out))
    return np.clip(out.astype(f32), f32(lo), f32(hi))

"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's integer-only path.

This file is the *checker* for the CUDA kernels.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs import it; the product package (``ivit_b200``) never does and
has no CPU fallback.

Parity status: **pinned by execution**.  The reference ships no golden vectors
or tests for this path, so the oracle is pinned against the
reference *itself*, imported with the shims in ``oracle/ref_shim.py``:
``tests/test_oracle_vs_reference.py`` (build container only) compares every
function here with the reference module it restates on seeded inputs, and
``oracle/make_golden.py`` stores reference outputs under ``tests/golden/`` which
``tests/test_oracle_golden.py`` re-checks anywhere.

Carrier convention = the reference's: a tensor travels as ``x = codes * scale``
in float32 together with its float32 ``scale`` (shape (1,), (C,) or
(1,C,1,1)).  Every function cites the reference lines it follows
(paths relative to /root/reference/models/quantization_utils/).

Deliberate, documented deviations (SURVEY.md section 8a hazards):
  H1/H6  GEMM operands are the *rounded* codes and the accumulation is exact
         (float64 BLAS on integers < 2^53); the reference feeds un-rounded
         ``x / s`` into a float32 GEMM whose noise (<< 0.5) is removed by the
         next ``round``.  Float32 row sums follow torch's CPU summation order
         exactly (``torch_sum_lastdim``).
"""
from __future__ import annotations

import decimal
import math
from decimal import Decimal

import numpy as np

f32 = np.float32
f64 = np.float64
EPS32 = np.finfo(np.float32).eps


def _f32(a):
    return np.asarray(a, dtype=np.float32)


# --------------------------------------------------------------------------
# quant_utils.py
# --------------------------------------------------------------------------
def sym_scale(num_bits: int, min_val, max_val):
    """quant_utils.py:52-70 symmetric_linear_quantization_params."""
    n = 2 ** (num_bits - 1) - 1
    mx = np.maximum(-_f32(min_val), _f32(max_val))
    scale = mx / f32(n)
    return np.maximum(scale, f32(EPS32)).astype(f32)


def _bcast(scale, ndim, is_weight):
    """quant_utils.py:23-46 scale reshape rules of linear_quantize."""
    scale = _f32(scale)
    if is_weight:
        if ndim == 4:
            return scale.reshape(-1, 1, 1, 1)
        if ndim == 2:
            return scale.reshape(-1, 1)
        return scale.reshape(-1)
    if ndim == 2:
        return scale.reshape(1, -1)
    if ndim == 3:
        return scale.reshape(1, 1, -1)
    if ndim == 4:
        return scale.reshape(1, -1, 1, 1)
    raise NotImplementedError


def sym_quantize(x, k: int, scale, is_weight: bool):
    """quant_utils.py:13-49 + :79-97: round(1/scale * x) (reciprocal-multiply), clamp."""
    x = _f32(x)
    s = _bcast(scale, x.ndim, is_weight)
    n = 2 ** (k - 1) - 1
    q = np.rint((f32(1.0) / s) * x + f32(0.0))
    # clamp bounds are python ints -> float32 (2^31-1 becomes 2^31)
    return np.clip(q, f32(-n - 1), f32(n)).astype(f32)


def batch_frexp(r):
    """quant_utils.py:151-175.  r: float64 array.  Returns (m int64, e float64).

    m = ROUND_HALF_UP(mantissa * 2^31) (half away from zero; can equal 2^31), e = 31 - exponent.
    """
    r = np.asarray(r, dtype=f64)
    shape = r.shape
    m0, e0 = np.frexp(r.reshape(-1))
    out = []
    for m in m0:
        out.append(int(Decimal(float(m) * (2 ** 31)).quantize(Decimal("1"), rounding=decimal.ROUND_HALF_UP)))
    return np.array(out, dtype=np.int64).reshape(shape), (31.0 - e0.astype(f64)).reshape(shape)


def _reshape_last(scale, ndim):
    """quant_utils.py:198-205 reshape lambda of fixedpoint_mul."""
    scale = np.asarray(scale)
    if ndim == 2:
        return scale.reshape(1, -1)
    if ndim == 3:
        return scale.reshape(1, 1, -1)
    if ndim == 4:
        return scale.reshape(1, -1, 1, 1)
    raise NotImplementedError


def fixedpoint_mul(pre_act, pre_scale, bit_num: int, z_scale, identity=None, identity_scale=None):
    """quant_utils.py:193-253 (symmetric mode).  Returns float32 codes (clamped)."""
    pre_act = _f32(pre_act)
    nd = pre_act.ndim
    n = 2 ** (bit_num - 1) - 1
    ps = _reshape_last(_f32(pre_scale), nd)
    z_int = np.rint(pre_act / ps)
    A = ps.astype(f64)
    B = _f32(z_scale).astype(f64)
    new_scale = _reshape_last(A / B, nd)
    m, e = batch_frexp(new_scale)
    out = z_int.astype(f64) * m.astype(f64)
    out = np.rint(out / (2.0 ** e))
    if identity is not None:
        ids = _reshape_last(_f32(identity_scale), nd)
        wx_int = np.rint(_f32(identity) / ids)
        A1 = ids.astype(f64)
        new_scale1 = _reshape_last(A1 / B, nd)
        m1, e1 = batch_frexp(new_scale1)
        out1 = wx_int.astype(f64) * m1.astype(f64)
        out1 = np.rint(out1 / (2.0 ** e1))
        out = out1 + out
    out = out.astype(f32)
    if bit_num in (4, 8, 16, 32):
        return np.clip(out, f32(-n - 1), f32(n))
    return out


# --------------------------------------------------------------------------
# quant_modules.py
# --------------------------------------------------------------------------
def quant_act(x, bits: int, x_min, x_max, pre_scale=None, identity=None, identity_scale=None):
    """quant_modules.py:362-387 QuantAct.forward with running_stat=False.

    Returns (x_q float32 = codes*scale, scale float32[1], codes float32)."""
    scale = sym_scale(bits, x_min, x_max)
    if pre_scale is None:
        codes = sym_quantize(x, bits, scale, is_weight=False)
    else:
        codes = fixedpoint_mul(x, pre_scale, bits, scale, identity, identity_scale)
    return (codes * scale.reshape(-1)).astype(f32), scale, codes


def linear_weights(weight, bias, prev_scale, weight_bit=8, bias_bit=32):
    """quant_modules.py:202-220: per-out-channel weight scale, int weights, int32 bias."""
    w = _f32(weight)
    w2 = w.reshape(w.shape[0], -1)
    s_w = sym_scale(weight_bit, w2.min(axis=1), w2.max(axis=1))
    w_int = sym_quantize(w, weight_bit, s_w, is_weight=True)
    b_scale = (s_w * _f32(prev_scale)).astype(f32)
    b_int = None
    if bias is not None:
        b_int = sym_quantize(_f32(bias), bias_bit, b_scale, is_weight=True)
    return s_w, w_int, b_scale, b_int


def _codes_of(x, scale, nd_scale_axis=-1):
    """Rounded integer codes of a carrier tensor (see module docstring, H1)."""
    return np.rint(_f32(x) / _f32(scale))


def quant_linear(x, prev_scale, weight, bias):
    """quant_modules.py:186-226 QuantLinear.forward (symmetric, per-channel).

    Returns (y float32, scale float32[out], acc float32 = integer accumulators incl. bias)."""
    prev_scale = _f32(prev_scale)
    assert prev_scale.shape == (1,)
    s_w, w_int, b_scale, b_int = linear_weights(weight, bias, prev_scale)
    xq = _codes_of(x, prev_scale.reshape(1, -1)).astype(f64)
    acc = xq.reshape(-1, xq.shape[-1]) @ w_int.astype(f64).T
    acc = acc.reshape(*xq.shape[:-1], -1)
    if b_int is not None:
        acc = acc + b_int.astype(f64)
    acc = acc.astype(f32)
    return (acc * b_scale).astype(f32), b_scale, acc


def quant_conv_patch(x, prev_scale, weight, bias, patch: int):
    """quant_modules.py:478-512 QuantConv2d.forward for kernel=stride=patch, padding 0.

    Returns (y float32 [B,C,H/p,W/p], scale float32 (1,C,1,1), acc float32)."""
    prev_scale = _f32(prev_scale)
    s_w, w_int, b_scale, b_int = linear_weights(weight, bias, prev_scale)
    xq = _codes_of(x, prev_scale.reshape(1, -1, 1, 1)).astype(f64)
    B, Cin, H, W = xq.shape
    gh, gw = H // patch, W // patch
    cols = xq.reshape(B, Cin, gh, patch, gw, patch).transpose(0, 2, 4, 1, 3, 5).reshape(B * gh * gw, -1)
    acc = cols @ w_int.reshape(w_int.shape[0], -1).astype(f64).T + b_int.astype(f64)
    acc = acc.reshape(B, gh, gw, -1).transpose(0, 3, 1, 2).astype(f32)
    scale = b_scale.reshape(1, -1, 1, 1)
    return (acc * scale).astype(f32), scale, acc


def quant_matmul(A, sA, B, sB):
    """quant_modules.py:404-409 QuantMatMul.forward."""
    sA, sB = _f32(sA), _f32(sB)
    a = _codes_of(A, sA).astype(f64)
    b = _codes_of(B, sB).astype(f64)
    scale = (sA * sB).astype(f32)
    acc = np.matmul(a, b).astype(f32)
    return (acc * scale).astype(f32), scale, acc


# --------------------------------------------------------------------------
# ivit_modules.py
# --------------------------------------------------------------------------
def _pow2(e):
    """2 ** e for integer-valued float exponents (torch.pow(2, e) is exact there; checked
    against the reference in tests/test_oracle_vs_reference.py)."""
    return np.ldexp(f32(1.0), np.asarray(e).astype(np.int32)).astype(f32)


def torch_sum_lastdim(x):
    """Bit-exact model of ``torch.sum(x, dim=-1, keepdim=True)`` for contiguous float32 on CPU.

    The reference's float32 row reductions (LayerNorm mean of the un-rounded ``x/s``,
    ivit_modules.py:37 / ibert_modules.py:130; softmax denominators, ivit_modules.py:171,
    ibert_modules.py:312; I-BERT variance, ibert_modules.py:134) are order dependent once
    partial sums leave the exactly representable range (hazard H6).  ATen's
    ``vectorized_inner_sum`` (aten/src/ATen/native/cpu/SumKernel.cpp, torch 2.11: 8-float
    vectors even on AVX-512 hosts, 4 interleaved accumulators, cascade every 16 steps)
    fixes the order as: element e of the first 32*floor(n/32) goes to accumulator e%32,
    accumulated sequentially (chunks of 16 steps cascade into a higher level); left-over whole
    vectors go to accumulator row 0; rows 1..3 are added to row 0 in order; the n%8 tail is
    summed first into a scalar, then the 8 lanes are added in order.  Verified bit-for-bit
    against torch in tests/test_oracle_vs_reference.py.
    """
    x = np.ascontiguousarray(x, dtype=np.float32)
    lead, n = x.shape[:-1], x.shape[-1]
    x2 = x.reshape(-1, n)
    R = x2.shape[0]
    V, ilp = 8, 4
    nvec = n // V
    size_ilp = nvec // ilp
    vecs = x2[:, :nvec * V].reshape(R, nvec, V)
    ceil_log2 = (size_ilp - 1).bit_length() if size_ilp > 0 else 0
    level_power = max(4, ceil_log2 // 4)
    level_step = 1 << level_power
    level_mask = level_step - 1
    acc = np.zeros((4, R, ilp, V), dtype=np.float32)
    grp = vecs[:, :size_ilp * ilp].reshape(R, size_ilp, ilp, V)
    i = 0
    while i + level_step <= size_ilp:
        for _ in range(level_step):
            acc[0] = acc[0] + grp[:, i]
            i += 1
        for j in range(1, 4):
            acc[j] = acc[j] + acc[j - 1]
            acc[j - 1] = 0
            if (i & (level_mask << (j * level_power))) != 0:
                break
    while i < size_ilp:
        acc[0] = acc[0] + grp[:, i]
        i += 1
    for j in range(1, 4):
        acc[0] = acc[0] + acc[j]
    ps = acc[0]                                   # [R, ilp, V]
    for v in range(size_ilp * ilp, nvec):
        ps[:, 0] = ps[:, 0] + vecs[:, v]
    for k in range(1, ilp):
        ps[:, 0] = ps[:, 0] + ps[:, k]
    final = np.zeros(R, dtype=np.float32)
    for e in range(nvec * V, n):
        final = final + x2[:, e]
    for l in range(V):
        final = final + ps[:, 0, l]
    return final.reshape(*lead, 1).astype(np.float32)


def _cascade_rows(x2, n_steps, load):
    """ATen multi_row_sum (SumKernel.cpp): `load(i)` -> [R, K] values of step i for K independent
    accumulators; sequential over i, cascading into a higher level every 2^max(4, ceil_log2(n)/4) steps."""
    lp = max(4, max(n_steps - 1, 0).bit_length() // 4)
    step, mask = 1 << lp, (1 << lp) - 1
    first = load(0) if n_steps > 0 else None
    acc = [np.zeros_like(first) if first is not None else 0 for _ in range(4)]
    i = 0
    while i + step <= n_steps:
        for _ in range(step):
            acc[0] = acc[0] + load(i)
            i += 1
        for j in range(1, 4):
            acc[j] = acc[j] + acc[j - 1]
            acc[j - 1] = np.zeros_like(acc[j - 1])
            if (i & (mask << (j * lp))) != 0:
                break
    while i < n_steps:
        acc[0] = acc[0] + load(i)
        i += 1
    for j in range(1, 4):
        acc[0] = acc[0] + acc[j]
    return acc[0]


def torch_outer_sum_lastdim(x):
    """``torch.sum(x, dim=-1, keepdim=True)`` on CPU float32 for x [B, L, C] whose reduced dimension C is
    STRIDED and whose token dimension L is the contiguous one (e.g. ``conv_out.flatten(2).transpose(1, 2)``,
    a layout every elementwise op -- QuantAct included -- preserves).  ATen takes its outer-reduction
    path (SumKernel.cpp ``vectorized_outer_sum``): tokens are processed in blocks of 32 (4 vectors of 8
    floats; threads split the token range on 32-token boundaries), each summing its C elements
    sequentially with a cascade every 16 (``multi_row_sum``); the tokens past the last whole block of 32
    (whole vectors, then scalars) go through ``row_sum`` instead: 4 interleaved accumulators
    (element e -> e % 4) cascaded the same way, left-overs into accumulator 0, then 1..3 added in order.
    Verified against torch in tests/test_swin_oracle.py."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    assert x.ndim == 3
    B, L, C = x.shape
    out = np.zeros((B, L), dtype=np.float32)
    nblk = (L // 32) * 32
    if nblk:
        xb = x[:, :nblk].reshape(-1, C)
        out[:, :nblk] = _cascade_rows(xb, C, lambda i: xb[:, i]).reshape(B, nblk)
    if nblk < L:
        xt = x[:, nblk:].reshape(-1, C)
        size_ilp = C // 4
        ps = _cascade_rows(xt, size_ilp, lambda i: xt[:, 4 * i:4 * i + 4])     # [R, 4]
        if size_ilp == 0:
            ps = np.zeros((xt.shape[0], 4), dtype=np.float32)
        p0 = ps[:, 0].copy()
        for e in range(size_ilp * 4, C):
            p0 = p0 + xt[:, e]
        for k in range(1, 4):
            p0 = p0 + ps[:, k]
        out[:, nblk:] = p0.reshape(B, L - nblk)
    return out[..., None].astype(np.float32)


def ivit_layernorm(x, scale, weight, bias, strided_input: bool = False):
    """ivit_modules.py:30-65 IVITIntLayerNorm.forward.  x: [B,N,C].  strided_input: the reference's
    tensor is a transposed view (channel stride != 1), which changes torch's float32 summation order
    of the mean (Swin's patch_embed.norm, layers_quant.py:198-201).

    Returns (y float32, scale float32[C], y_int float32 = integer output incl. bias_int)."""
    x = _f32(x)
    C = x.shape[2]
    dim_sqrt = np.sqrt(f32(C)).astype(f32)
    x_int = x / _f32(scale)                                   # :36 un-rounded
    row_sum = torch_outer_sum_lastdim(x_int) if strided_input else torch_sum_lastdim(x_int)
    mean_int = np.rint(row_sum / f32(C))                      # :37 (mean = sum / C, torch sum order)
    p = np.trunc(x_int).astype(np.int32)                      # :38 .to(int32) truncates
    mean_i = mean_int.astype(np.int32)
    y_int = p - mean_i                                        # :40
    y_sq = (y_int.astype(np.int64) * y_int.astype(np.int64)).astype(np.int32)   # :41 int32 wrap (H9)
    var_int = y_sq.astype(np.int64).sum(axis=2, keepdims=True)                   # :42 int64 sum
    var_f = var_int.astype(f32)
    k = f32(2 ** 16)
    for _ in range(10):                                       # :45-49
        k = np.floor((k + np.floor(var_f / k)) / f32(2))
    std_int = k
    factor = np.floor(f32(2 ** 31 - 1) / std_int)             # :51
    y = np.floor(y_int.astype(f32) * factor / f32(2))         # :52
    sf = (dim_sqrt / f32(2 ** 30)).astype(f32)                # :53
    b = _f32(bias) / _f32(weight)                             # :56
    bias_int = np.floor(b / sf)                               # :57
    y = (y + bias_int).astype(f32)                            # :61
    sf_c = (sf * _f32(weight)).astype(f32)                    # :62
    return (y * sf_c).astype(f32), sf_c, y


def _ivit_exp_shift(x_int, scaling_factor, n: int):
    """ivit_modules.py:89-103 / :150-162 int_exp_shift (x_int float32, integer- or not)."""
    x_int = _f32(x_int)
    x_int = x_int + np.floor(x_int / f32(2)) - np.floor(x_int / f32(2 ** 4))
    x0_int = np.floor(f32(-1.0) / _f32(scaling_factor))
    x_int = np.maximum(x_int, f32(n) * x0_int)
    q = np.floor(x_int / x0_int)
    r = x_int - x0_int * q
    exp_int = r / f32(2) - x0_int
    exp_int = np.maximum(np.floor(exp_int * _pow2(f32(n) - q)), f32(0))
    return exp_int.astype(f32)


def ivit_gelu(x, scale, output_bit: int = 8, n: int = 23):
    """ivit_modules.py:105-126 IVITIntGELU.forward (ShiftGELU).

    Returns (y float32, scale float32[1], out_int float32)."""
    x = _f32(x)
    scale = _f32(scale)
    pre = np.trunc(x / scale).astype(np.int32)                # :106-107
    s_sig = (scale * f32(1.702)).astype(f32)                  # :108
    mx = pre.max(axis=-1, keepdims=True)                      # :110
    x_int = (pre - mx).astype(f32)                            # :111 (int32), :90 no-op cast
    exp_int = _ivit_exp_shift(x_int, s_sig, n)                # :113
    exp_max = _ivit_exp_shift((-mx).astype(f32), s_sig, n)    # :115
    ssum = np.minimum(exp_int + exp_max, f32(2 ** 31 - 1))    # :116-118
    factor = np.floor(f32(2 ** 31 - 1) / ssum)                # :119
    sig = np.floor(exp_int * factor / f32(2 ** (31 - output_bit + 1)))   # :120
    sig_scale = f32(1 / 2 ** (output_bit - 1))
    out_int = (pre.astype(f32) * sig).astype(f32)             # :123
    out_scale = (scale * sig_scale).astype(f32)               # :124
    return (out_int * out_scale).astype(f32), out_scale, out_int


def ivit_softmax(x, scale, output_bit: int = 8, n: int = 15):
    """ivit_modules.py:164-179 IVITIntSoftmax.forward (Shiftmax)."""
    x = _f32(x)
    scale = _f32(scale)
    x_int = x / scale                                         # :165 un-rounded (:166 discarded)
    x_int = x_int - x_int.max(axis=-1, keepdims=True)         # :167-168
    exp_int = _ivit_exp_shift(x_int, scale, n)                # :170
    ssum = np.minimum(torch_sum_lastdim(exp_int), f32(2 ** 31 - 1))   # :171-173
    factor = np.floor(f32(2 ** 31 - 1) / ssum)                # :174
    out_int = np.floor(exp_int * factor / f32(2 ** (31 - output_bit + 1)))   # :175
    out_scale = np.array([1 / 2 ** (output_bit - 1)], dtype=f32)
    return (out_int * out_scale).astype(f32), out_scale, out_int.astype(f32)


# --------------------------------------------------------------------------
# ibert_modules.py
# --------------------------------------------------------------------------
def ibert_layernorm(x, scale, weight, bias, shift=0.0, use_int_sqrt: bool = False):
    """ibert_modules.py:112-158 IBERTIntLayerNorm.forward (frozen: overflow_handling=False)."""
    x = _f32(x)
    C = x.shape[2]
    dim_sqrt = np.sqrt(f32(C)).astype(f32)
    x_int = x / _f32(scale)                                   # :129
    mean_int = np.rint(torch_sum_lastdim(x_int) / f32(C))     # :130
    y_int = x_int - mean_int                                  # :131 (float, no truncation)
    sh = _pow2(f32(shift))
    y_sh = np.floor(y_int / sh)                               # :132
    var_int = torch_sum_lastdim(y_sh * y_sh)                     # :133-134 (float32 sum; H6)
    if use_int_sqrt:                                          # :85-109, :143
        nn_ = np.maximum(var_int, f32(0))
        bits = np.floor(np.log2(np.maximum(nn_, f32(1)))) + f32(1)
        xk = _pow2(np.ceil(bits / f32(2))).astype(np.int64)
        for _ in range(4):
            inv = np.floor(nn_ / np.maximum(xk, 1).astype(f32))
            xk = np.floor((xk.astype(f32) + inv) / f32(2))
        std = np.where(var_int > 0, xk.astype(np.int32), 0).astype(f32)
        std_int = np.floor(std) * sh
    else:
        std_int = np.floor(np.sqrt(var_int)) * sh             # :145
    factor = np.floor(f32(2 ** 31) / std_int)                 # :146
    y = np.floor(y_int * factor / f32(2))                     # :147
    sf = (dim_sqrt / f32(2 ** 30)).astype(f32)
    b = _f32(bias) / _f32(weight)
    bias_int = np.floor(b / sf)
    y = (y + bias_int).astype(f32)
    sf_c = (sf * _f32(weight)).astype(f32)
    return (y * sf_c).astype(f32), sf_c, y


def ibert_gelu(x, scale):
    """ibert_modules.py:203-235 IBERTIntGELU.forward."""
    x = _f32(x)
    scale = _f32(scale)
    k, n = 1.4142, 6
    coeff = [-0.2888, -1.769, 1.0]
    coeff[2] /= coeff[0]
    x_int = x / scale                                         # :227 un-rounded
    s = (scale / f32(k)).astype(f32)                          # :228
    b_int = np.floor(f32(coeff[1]) / s)                       # :205
    c_int = np.floor(f32(coeff[2]) / (s * s))                 # :206  (s ** 2)
    sign = np.sign(x_int)
    abs_int = np.minimum(np.abs(x_int), -b_int)
    y_int = (abs_int + b_int) * (abs_int + b_int) + c_int     # :212
    y_int = sign * y_int
    s_erf = (s * s) * f32(coeff[0])                           # :214
    y_int = np.floor(y_int / f32(2 ** n))                     # :215
    s_erf = (s_erf * f32(2 ** n)).astype(f32)                 # :216
    shift_int = np.floor(f32(1.0) / s_erf)                    # :230
    out_int = (x_int * (y_int + shift_int)).astype(f32)       # :232
    out_scale = (scale * s_erf / f32(2)).astype(f32)          # :233
    return (out_int * out_scale).astype(f32), out_scale, out_int


def ibert_softmax(x, scale, output_bit: int, act_x_min, act_x_max):
    """ibert_modules.py:297-319 IBERTIntSoftmax.forward (inner QuantAct(16) frozen)."""
    x = _f32(x)
    scale = _f32(scale)
    x0, n = -0.6931, 30
    coef = [0.35815147, 0.96963238, 1.0]
    coef[1] /= coef[0]
    coef[2] /= coef[0]
    x_int = x / scale                                         # :304
    x_int = x_int - x_int.max(axis=-1, keepdims=True)         # :306-307
    x0_int = np.floor(f32(x0) / scale)                        # :287
    x_int = np.maximum(x_int, f32(n) * x0_int)                # :288
    q = np.floor(x_int / x0_int)                              # :290
    r = x_int - x0_int * q                                    # :291
    b_int = np.floor(f32(coef[1]) / scale)                    # :277
    c_int = np.floor(f32(coef[2]) / (scale * scale))          # :278
    z = r + b_int
    z = r * z
    z = z + c_int                                             # :279-281
    exp_scale = (f32(coef[0]) * (scale * scale)).astype(f32)  # :282
    exp_int = np.maximum(np.floor(z * _pow2(f32(n) - q)), f32(0))   # :293
    exp_scale = (exp_scale / f32(2 ** n)).astype(f32)         # :294
    # :310 inner QuantAct(16): input "exp_int" is passed as the value, exp_scale as its scale
    expq, s16, _ = quant_act(exp_int, 16, act_x_min, act_x_max, pre_scale=exp_scale)
    exp_int = expq / s16                                      # :311
    ssum = torch_sum_lastdim(exp_int)                            # :312
    factor = np.floor(f32(2 ** 32) / ssum)                    # :314
    out_int = np.floor(exp_int * factor / f32(2 ** (32 - output_bit + 1)))   # :315
    out_scale = np.array([2 / 2 ** output_bit], dtype=f32)    # :317
    return (out_int * out_scale).astype(f32), out_scale, out_int.astype(f32)


# --------------------------------------------------------------------------
# ppoly_modules.py / ppoly_backend.py (frozen: the stored integer coefficients are inputs)
# --------------------------------------------------------------------------
def ppoly_horner(x_int, bounds, coeffs):
    """ppoly_backend.py:293-351 evaluate_piecewise_polynomial.  x_int integer-valued array; bounds int [S-1];
    coeffs int [S, D+1] (highest power first).  Horner in int64, result WRAPPED to int32 like `.to(torch.int32)`."""
    x = np.asarray(x_int).astype(np.int64)
    bounds = np.asarray(bounds).reshape(-1).astype(np.int64)
    coeffs = np.asarray(coeffs).astype(np.int64).reshape(len(bounds) + 1, -1)
    seg = np.searchsorted(bounds, x, side="right")            # x < b0 -> 0; b_{i-1} <= x < b_i -> i; x >= b_last -> S-1
    r = coeffs[seg, 0]
    for d in range(1, coeffs.shape[1]):
        r = r * x + coeffs[seg, d]
    return r.astype(np.int32)                                 # wraps modulo 2^32


def ppoly_gelu(x, scale, bounds, coeffs, scale_out, scale_bits: int = 22):
    """ppoly_modules.py:110-136 PPolyIntGELU.forward with fixed coefficients.  Returns (y, scale_out)."""
    x = _f32(x)
    scale = _f32(scale)
    scale_out = _f32(scale_out)
    x_int = np.floor(x / scale)                                           # :114
    y_int = ppoly_horner(x_int, bounds, coeffs)                           # :128
    y_float = (y_int.astype(f32) / f32(2 ** scale_bits)).astype(f32)      # :134 (int32 / python int -> float32)
    y = (scale_out * np.floor(y_float / scale_out)).astype(f32)           # :140
    return y, scale_out


def ppoly_softmax(x, scale, bounds, coeffs, output_bit: int = 8, exp_bits: int = 16):
    """ppoly_modules.py:296-342 PPolyIntSoftmax.forward with fixed coefficients (the polynomial's own scale_bits only
    enter through the stored coefficients).  Returns (y, scale_out, codes)."""
    x = _f32(x)
    scale = _f32(scale)
    x_int = np.floor(x / scale).astype(np.int32)                          # :302
    x_int = x_int - x_int.max(axis=-1, keepdims=True) + 127              # :305-306
    exp_int = np.maximum(ppoly_horner(x_int, bounds, coeffs), 0)         # :309, :271
    exp_f = np.floor(exp_int.astype(f32) / f32(2 ** (30 - exp_bits + 1)))   # :312
    ssum = np.maximum(torch_sum_lastdim(exp_f), f32(1.0))                # :315-318
    factor = np.floor(f32(2 ** 32) / ssum)                               # :322
    out_int = np.floor(exp_f * factor / f32(2 ** (32 - output_bit + 1)))   # :323
    s_out = np.array([2 / 2 ** output_bit], dtype=f32)                    # :329
    y = (s_out * np.floor((out_int * s_out).astype(f32) / s_out)).astype(f32)   # :330-336
    return y, s_out, out_int.astype(f32)

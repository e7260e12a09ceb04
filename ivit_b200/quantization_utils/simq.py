"""Float-simulated arithmetic of the I-ViT / I-BERT integer path, as torch functions.

This is the *calibration / unfrozen* mode of the module API (``running_stat=True``: range tracking
before ``freeze_model``; also QAT-style use): tensors travel as ``codes * scale`` in float32 exactly
like the reference, on whatever device they live on.  It is NOT the hot path -- after ``fix()`` the
modules exchange integer carriers and launch the sm_100a kernels (``modules.py``) -- but it must be
value-identical to the reference so that calibrated ``x_min`` / ``x_max`` buffers, and therefore
every frozen scale, agree (tests/test_module_api_cpu.py checks them against the reference fixtures).
Each function cites the reference lines it restates (models/quantization_utils/<file>:<line>).
"""
from __future__ import annotations

import decimal
from decimal import Decimal

import numpy as np
import torch

_EPS = torch.finfo(torch.float32).eps


def sym_scale(bits: int, lo: torch.Tensor, hi: torch.Tensor) -> torch.Tensor:
    """quant_utils.py:52-70: max(-min, max) / (2^(k-1) - 1), clamped to float32 eps."""
    n = 2 ** (bits - 1) - 1
    with torch.no_grad():
        scale = torch.max(-lo, hi) / float(n)
        return scale.clamp(min=_EPS)


def _shape_scale(scale: torch.Tensor, ndim: int, is_weight: bool) -> torch.Tensor:
    """quant_utils.py:23-46."""
    if is_weight:
        return scale.view(-1, *([1] * (ndim - 1))) if ndim in (2, 4) else scale.view(-1)
    if ndim == 2:
        return scale.view(1, -1)
    if ndim == 3:
        return scale.view(1, 1, -1)
    if ndim == 4:
        return scale.view(1, -1, 1, 1)
    raise NotImplementedError(f"{ndim}-d input")


def sym_quantize(x: torch.Tensor, bits: int, scale: torch.Tensor, is_weight: bool) -> torch.Tensor:
    """quant_utils.py:13-49 + :79-97: clamp(round(1/scale * x)) -- reciprocal, then multiply."""
    n = 2 ** (bits - 1) - 1
    s = _shape_scale(scale, x.dim(), is_weight)
    return torch.clamp(torch.round(1.0 / s * x), -n - 1, n)


def batch_frexp(ratio: torch.Tensor):
    """quant_utils.py:151-175: (m, e) with m = ROUND_HALF_UP(mantissa * 2^31), e = 31 - exponent.
    Host-side by nature (numpy frexp + decimal); only used in simulated mode and at freeze time."""
    shape = ratio.size()
    r = ratio.detach().cpu().double().numpy().reshape(-1)
    m0, e0 = np.frexp(r)
    m = np.array([int(Decimal(float(v) * (2 ** 31)).quantize(Decimal("1"), rounding=decimal.ROUND_HALF_UP))
                  for v in m0], dtype=np.float64)
    e = 31.0 - e0.astype(np.float64)
    dev = ratio.device
    return torch.from_numpy(m).to(dev).view(shape), torch.from_numpy(e).to(dev).view(shape)


def _last_dim(t: torch.Tensor, ndim: int) -> torch.Tensor:
    return _shape_scale(t, ndim, False)


def fixedpoint_mul(x, s_in, bits: int, s_out, identity=None, s_id=None) -> torch.Tensor:
    """quant_utils.py:193-253 (forward only): dyadic requantisation, optional identity branch."""
    nd = x.dim()
    n = 2 ** (bits - 1) - 1
    with torch.no_grad():
        s_in = _last_dim(s_in, nd)
        z = torch.round(x / s_in)
        ratio = _last_dim(s_in.double() / s_out.float().double(), nd)
        m, e = batch_frexp(ratio)
        out = torch.round(z.double() * m.double() / (2.0 ** e))
        if identity is not None:
            s_id = _last_dim(s_id, nd)
            zi = torch.round(identity / s_id)
            m1, e1 = batch_frexp(_last_dim(s_id.double() / s_out.float().double(), nd))
            out = torch.round(zi.double() * m1.double() / (2.0 ** e1)) + out
        out = out.float()
        if bits in (4, 8, 16, 32):
            return torch.clamp(out, -n - 1, n)
        return out


# ---------------------------------------------------------------------------- I-ViT nonlinears
def ivit_layernorm(x, s, weight, bias, dim_sqrt):
    """ivit_modules.py:30-65.  Returns (y, per-channel scale, bias_int)."""
    n = torch.tensor(x.shape[2], dtype=torch.float32, device=x.device)
    x_int = x / s
    mean_int = torch.round(x_int.mean(axis=2, keepdim=True)).to(torch.int32)
    y_int = x_int.to(torch.int32) - mean_int
    var_int = torch.sum(y_int ** 2, axis=2, keepdim=True)
    k = 2 ** 16
    for _ in range(10):
        k = torch.floor((k + torch.floor(var_int / k)) / 2)
    factor = torch.floor((2 ** 31 - 1) / k)
    y_int = torch.floor(y_int * factor / 2)
    sf = dim_sqrt / 2 ** 30
    bias_int = torch.floor((bias.detach() / weight.detach()) / sf)
    y_int = y_int + bias_int
    sf = sf * weight
    return y_int * sf, sf, bias_int


def _ivit_exp_shift(x_int, s, n: int):
    """ivit_modules.py:89-103 / :150-162."""
    x_int = x_int + torch.floor(x_int / 2) - torch.floor(x_int / 2 ** 4)
    with torch.no_grad():
        x0 = torch.floor(-1.0 / s)
    x_int = torch.max(x_int, n * x0)
    q = torch.floor(x_int / x0)
    r = x_int - x0 * q
    e = r / 2 - x0
    e = torch.clamp(torch.floor(e * 2 ** (n - q)), min=0)
    return e, s / 2 ** n


def ivit_gelu(x, s, output_bit: int = 8, n: int = 23):
    """ivit_modules.py:105-126 (ShiftGELU)."""
    pre = (x / s).to(torch.int32)
    s_sig = s * 1.702
    mx, _ = pre.max(dim=-1, keepdim=True)
    x_int = (pre - mx).float()
    e, _ = _ivit_exp_shift(x_int, s_sig, n)
    em, _ = _ivit_exp_shift((-mx).float(), s_sig, n)
    ssum = (e + em).clamp_max(2 ** 31 - 1)
    factor = torch.floor((2 ** 31 - 1) / ssum)
    sig = torch.floor(e * factor / 2 ** (31 - output_bit + 1))
    s_out = s * (1 / 2 ** (output_bit - 1))
    return (pre * sig) * s_out, s_out


def ivit_softmax(x, s, output_bit: int = 8, n: int = 15):
    """ivit_modules.py:164-179 (Shiftmax)."""
    x_int = x / s
    x_int = x_int - x_int.max(dim=-1, keepdim=True)[0]
    e, _ = _ivit_exp_shift(x_int, s, n)
    ssum = e.sum(dim=-1, keepdim=True).clamp_max(2 ** 31 - 1)
    factor = torch.floor((2 ** 31 - 1) / ssum)
    out = torch.floor(e * factor / 2 ** (31 - output_bit + 1))
    s_out = torch.tensor([1 / 2 ** (output_bit - 1)], device=x.device)
    return out * s_out, s_out


# ---------------------------------------------------------------------------- I-BERT nonlinears
def ibert_int_sqrt(nv: torch.Tensor) -> torch.Tensor:
    """ibert_modules.py:85-109."""
    mask = nv > 0
    nv = torch.clamp(nv, min=0)
    bits = torch.floor(torch.log2(nv.float().clamp(min=1))) + 1
    xk = (2 ** torch.ceil(bits / 2)).to(torch.int64)
    for _ in range(4):
        inv = torch.floor(nv / torch.clamp(xk, min=1))
        xk = torch.floor((xk + inv) / 2)
    x32 = xk.to(torch.int32)
    return torch.where(mask, x32, torch.zeros_like(x32))


def ibert_layernorm(x, s, weight, bias, dim_sqrt, shift, use_int_sqrt: bool, overflow_cb=None):
    """ibert_modules.py:112-158.  overflow_cb(y_int) -> var_int implements the training-time
    dynamic shift (:137-140)."""
    n = torch.tensor(x.shape[2], dtype=torch.float32, device=x.device)
    x_int = x / s
    mean_int = torch.round(x_int.mean(axis=2, keepdim=True))
    y_int = x_int - mean_int
    y_sh = torch.floor(y_int / 2 ** shift)
    var_int = torch.sum(y_sh ** 2, axis=2, keepdim=True)
    if overflow_cb is not None and var_int.max() >= 2 ** 32:
        var_int = overflow_cb(y_int)
    if use_int_sqrt:
        std_int = torch.floor(ibert_int_sqrt(var_int)) * 2 ** shift
    else:
        std_int = torch.floor(torch.sqrt(var_int)) * 2 ** shift
    factor = torch.floor(2 ** 31 / std_int)
    y_int = torch.floor(y_int * factor / 2)
    sf = dim_sqrt / 2 ** 30
    bias_int = torch.floor((bias.detach() / weight.detach()) / sf)
    y_int = y_int + bias_int
    sf = sf * weight
    return y_int * sf, sf


def ibert_gelu(x, s):
    """ibert_modules.py:203-235."""
    k, n = 1.4142, 6
    c0, c1 = -0.2888, -1.769
    c2 = 1.0 / c0
    x_int = x / s
    sk = s / k
    with torch.no_grad():
        b_int = torch.floor(c1 / sk)
        c_int = torch.floor(c2 / sk ** 2)
    sign = torch.sign(x_int)
    a = torch.min(torch.abs(x_int), -b_int)
    y = sign * ((a + b_int) ** 2 + c_int)
    s_erf = sk ** 2 * c0
    y = torch.floor(y / 2 ** n)
    s_erf = s_erf * 2 ** n
    shift_int = torch.floor(1.0 / s_erf)
    out = x_int * (y + shift_int)
    s_out = s * s_erf / 2
    return out * s_out, s_out


def ibert_int_exp(x_int, s):
    """ibert_modules.py:275-295."""
    x0, n = -0.6931, 30
    c0 = 0.35815147
    c1, c2 = 0.96963238 / c0, 1.0 / c0
    with torch.no_grad():
        x0_int = torch.floor(x0 / s)
    x_int = torch.max(x_int, n * x0_int)
    q = torch.floor(x_int / x0_int)
    r = x_int - x0_int * q
    with torch.no_grad():
        b_int = torch.floor(c1 / s)
        c_int = torch.floor(c2 / s ** 2)
    z = r + b_int
    z = r * z
    z = z + c_int
    s_exp = c0 * s ** 2
    e = torch.clamp(torch.floor(z * 2 ** (n - q)), min=0)
    return e, s_exp / 2 ** n

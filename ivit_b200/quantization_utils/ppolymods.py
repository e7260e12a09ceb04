"""PPolyIntGELU / PPolyIntSoftmax (ppoly_modules.py:14-136, :137-342) on the module API.

Calibration / float-simulation mode runs torch ops in the reference's order (the segment fit is host numpy, as in the
reference).  Frozen mode on integer carriers: the module's stored integer coefficients become a per-layer table
(`ppoly.py`), consumed by the GELU table kernel and the fused attention kernel -- CUDA only, no CPU path.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from .. import _cabi, hostmath as hm, ops
from . import frozen as fz
from . import ppoly as pp
from . import simq
from .carrier import Pending, QTensor
from .modules import _IntGELUBase, _IntSoftmaxBase, _dev

f32 = np.float32


def _horner_torch(x_int, bounds, coeffs, segments, degree):
    """ppoly_backend.py:293-351 on tensors: int64 Horner per segment, int32 result (wrapping)."""
    x = x_int.to(torch.int64)
    b = bounds.reshape(-1).to(torch.int64)
    c = coeffs.reshape(segments, degree + 1).to(torch.int64)
    seg = torch.bucketize(x, b, right=True) if segments > 1 else torch.zeros_like(x)
    acc = c[seg, 0]
    for d in range(1, degree + 1):
        acc = acc * x + c[seg, d]
    return acc.to(torch.int32)


class _PPolyBase:
    def fix(self):
        self.fixed = True

    def unfix(self):
        self.fixed = False
        self.fixed_bounds = None
        self.fixed_coeffs = None
        if hasattr(self, "fixed_scaling_factor_out"):
            self.fixed_scaling_factor_out = None
        self.invalidate()

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs):
        # the coefficient buffers are registered as None until the first fixed forward: accept them from a checkpoint
        for name in ("fixed_bounds", "fixed_coeffs", "fixed_scaling_factor_out"):
            key = prefix + name
            if key in state_dict and hasattr(self, name) and getattr(self, name) is None:
                setattr(self, name, state_dict[key].clone())
        super()._load_from_state_dict(state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs)


class PPolyIntGELU(_PPolyBase, _IntGELUBase, nn.Module):
    """ppoly_modules.py:14-136."""
    _kind = 2

    def __init__(self, output_bit=8, scale_bits=22, seg=16, deg=2, backend="ibert", alpha=0.0, optim_bounds=True):
        super().__init__()
        self.N, self.segments, self.degree = scale_bits, seg, deg
        self.backend, self.alpha, self.optim_bounds = backend, alpha, optim_bounds
        self.fixed = False
        self.gelu_module = nn.GELU()
        self.register_buffer("act_scaling_factor", torch.zeros(1))
        self.register_buffer("fixed_bounds", None)
        self.register_buffer("fixed_coeffs", None)
        self.register_buffer("fixed_scaling_factor_out", None)

    def _fit(self, x, s):
        """ppoly_modules.py:57-108: sample the golden function over [floor(min x), ceil(max x)], fit, integerise."""
        x_lo, x_hi = torch.floor(torch.min(x)).item(), torch.ceil(torch.max(x)).item()
        if self.backend == "ibert":
            xs = torch.linspace(x_lo, x_hi, 10000, device=s.device, dtype=s.dtype)
            ys, _ = simq.ibert_gelu(xs, s)
            xs_np, ys_np = xs.detach().cpu().numpy(), ys.detach().cpu().numpy()
        else:
            from scipy.special import erf
            xs_np = np.linspace(x_lo, x_hi, 10000, dtype=np.float32)
            ys_np = 0.5 * xs_np * (1 + erf(xs_np / np.sqrt(2)))
        pieces = pp.fit_pieces(xs_np, ys_np, x_lo, x_hi, self.segments, self.degree, self.alpha, self.optim_bounds)
        bounds, coeffs = pp.integer_coefficients(pieces, s, self.N, x.device)
        if self.backend == "ibert":
            so = s / 1.4142
            so = so ** 2 * -0.2888
            so = so * (2 ** 6)
            so = s * so / 2
        else:
            so = s / (2 ** self.N)
        self.fixed_bounds, self.fixed_coeffs, self.fixed_scaling_factor_out = bounds, coeffs, so
        return bounds, coeffs, so

    def forward(self, x, scaling_factor=None):
        if isinstance(x, QTensor):
            return self._frozen_forward(x)
        x_int = torch.floor(x / scaling_factor)
        with torch.no_grad():
            if self.fixed and self.fixed_coeffs is not None:
                bounds, coeffs, s_out = self.fixed_bounds, self.fixed_coeffs, self.fixed_scaling_factor_out
            else:
                bounds, coeffs, s_out = self._fit(x, scaling_factor)
            y_int = _horner_torch(x_int, bounds, coeffs, self.segments, self.degree)
        g = self.gelu_module(x)
        y = y_int / (2 ** self.N)
        y = y.detach() + (g - g.detach())
        return s_out * torch.floor(y / s_out), s_out

    # ---- frozen: a [code] table --------------------------------------------------------------------
    def _frozen_forward(self, x: QTensor):
        fz.require_cuda(x, type(self).__name__)
        if self.fixed_coeffs is None:
            raise _cabi.IvitKernelError("frozen PPolyIntGELU: no stored coefficients (run one fixed float forward, or load "
                                        "a checkpoint that holds fixed_bounds / fixed_coeffs)")
        c = self._fz_get()
        if "s_out" not in c:
            h = fz.host_scale(self.fixed_scaling_factor_out).reshape(1).astype(f32)
            c["s_out_h"], c["s_out"] = h, _dev(h, torch.float32, x.device)
            c["s_in"] = float(f32(fz.eff_hscale(x)[0]))
        codes = x.materialize()
        if codes.dtype != torch.int8:
            raise NotImplementedError("frozen IntGELU: 8-bit input codes")
        q = QTensor(pending=Pending("gelu", x=codes, mod=self), canon_shape=tuple(x.shape), scale=c["s_out"],
                    hscale=c["s_out_h"], device=x.device)
        return q, c["s_out"]

    def _frozen_gelu_requant(self, p, qact, device):
        c = self._fz_get()
        key = ("lut", id(qact))
        if key not in c:
            _, s_next_h, (qmin, qmax) = qact._out_scale(device)
            c[key] = gelu_lut_from_coeffs(c["s_in"], self.fixed_bounds.cpu().numpy(), self.fixed_coeffs.cpu().numpy(),
                                          c["s_out_h"], self.N, s_next_h, int(qact.activation_bit), device)
        x = p.x if p.x.is_contiguous() else p.x.contiguous()
        return ops.gelu_lut_apply(x, c[key])


def gelu_lut_from_coeffs(s_in, bounds, coeffs, s_gelu, scale_bits, s_next, bits, device) -> torch.Tensor:
    """int8 [256][256] table for ivit_gelu_lut_apply_i8: PPolyIntGELU on every code, then the following QuantAct
    (quant_utils.py:220-251).  Every row-max row is the same line."""
    y = pp.gelu_code_values(s_in, bounds, coeffs, s_gelu, scale_bits)
    z = hm.recover_codes(y, np.asarray(s_gelu, dtype=f32).reshape(1))
    line = hm.requant_host(z, np.asarray(s_gelu, dtype=f32).reshape(1), np.asarray(s_next, dtype=f32).reshape(1), bits)
    line = np.clip(line, -128, 127).astype(np.int8)
    return _dev(np.ascontiguousarray(np.broadcast_to(line, (256, 256))), torch.int8, device)


class PPolyIntSoftmax(_PPolyBase, _IntSoftmaxBase, nn.Module):
    """ppoly_modules.py:137-342."""
    _kind = 2

    def __init__(self, output_bit=8, scale_bits=28, exp_bits=16, seg=16, deg=2, backend="float", alpha=0.0,
                 optim_bounds=False):
        super().__init__()
        self.output_bit, self.N, self.exp_bitwidth = output_bit, scale_bits, exp_bits
        self.segments, self.degree = seg, deg
        self.backend, self.alpha, self.optim_bounds = backend, alpha, optim_bounds
        self.fixed = False
        self.softmax_module = nn.Softmax(dim=-1)
        self.register_buffer("fixed_bounds", None)
        self.register_buffer("fixed_coeffs", None)
        self.register_buffer("fixed_scaling_factor_out", None)

    def _fit(self, x_int, s):
        """ppoly_modules.py:211-259: fit exp((x_off - 127) * s) over the offset inputs seen in this batch."""
        lo_i, hi_i = torch.floor(torch.min(x_int)).item(), torch.ceil(torch.max(x_int)).item()
        s_np = s.cpu().numpy()
        if self.backend == "ibert":
            xo = torch.linspace(lo_i, hi_i, 10000, device=x_int.device, dtype=x_int.dtype)
            e, se = simq.ibert_int_exp(xo - 127, s)
            xs_np, ys_np = (xo * s).detach().cpu().numpy(), (e * se).detach().cpu().numpy()
        else:
            xo = np.linspace(lo_i, hi_i, 10000)
            ys_np = np.exp((xo - 127) * s_np)
            xs_np = xo * s_np
        pieces = pp.fit_pieces(xs_np, ys_np, lo_i * s_np, hi_i * s_np, self.segments, self.degree, self.alpha,
                               self.optim_bounds)
        self.fixed_bounds, self.fixed_coeffs = pp.integer_coefficients(pieces, s, self.N, x_int.device)
        return self.fixed_bounds, self.fixed_coeffs

    def forward(self, x, scaling_factor):
        if isinstance(x, QTensor):
            if self.fixed_coeffs is None:
                raise _cabi.IvitKernelError("frozen PPolyIntSoftmax: no stored coefficients")
            return self._frozen_forward(x)
        s = scaling_factor.to(x.device)
        x_int = torch.floor(x / s).to(torch.int32)
        x_int = x_int - x_int.max(dim=-1, keepdim=True)[0] + 127
        with torch.no_grad():
            if self.fixed and self.fixed_coeffs is not None:
                bounds, coeffs = self.fixed_bounds, self.fixed_coeffs
            else:
                bounds, coeffs = self._fit(x_int, s)
            e = torch.clamp(_horner_torch(x_int, bounds, coeffs, self.segments, self.degree), min=0)
        e = torch.floor(e / 2 ** (30 - self.exp_bitwidth + 1))
        ssum = torch.clamp(e.sum(dim=-1, keepdim=True), min=1.0)
        with torch.no_grad():
            factor = torch.floor(2 ** 32 / ssum)
            sm_int = torch.floor(e * factor / 2 ** (32 - self.output_bit + 1))
        sf = self.softmax_module(x)
        s_out = torch.tensor([2 / 2 ** self.output_bit], device=x.device)
        out = (sm_int * s_out).detach() + (sf - sf.detach())
        return s_out * torch.floor(out / s_out), s_out

    def host_term_table(self, s_attn) -> np.ndarray:
        return pp.softmax_term_table(s_attn, self.fixed_bounds.cpu().numpy(), self.fixed_coeffs.cpu().numpy(),
                                     self.exp_bitwidth)

"""`float` registry entries (layer_selection.py:11-113): the true float GELU / softmax / LayerNorm
with a quantised OUTPUT, kept by the reference to isolate the error of the integer approximations.

They are floating-point diagnostics, not part of the integer hot path, so they are plain torch ops
in every mode and on any device; a frozen integer carrier coming in is dequantised first, and the
float result goes out with its scale -- the following frozen QuantAct re-enters the integer domain
through ``round(x / scale)`` + the requantisation kernel.  Same constructor signatures, forward
convention and parameter names (``layer_norm.weight`` / ``layer_norm.bias``) as the reference.
"""
from __future__ import annotations

import torch
from torch import nn

from .carrier import QTensor


def _as_float(x):
    return x.dequantize() if isinstance(x, QTensor) else x


def _quantise_output(y, scale, qmin, qmax):
    q = torch.clamp(torch.floor(y / scale), qmin, qmax)         # floor_ste + clamp
    return q * scale


class FloatGELU(nn.Module):
    """layer_selection.py:11-38."""

    def __init__(self, bitwidth=8, **kwargs):
        super().__init__()
        self.gelu = nn.GELU()
        self.bitwidth = bitwidth

    def forward(self, x, scaling_factor):
        out = self.gelu(_as_float(x))
        return _quantise_output(out, scaling_factor, -(2 ** (self.bitwidth - 1)), 2 ** (self.bitwidth - 1) - 1), \
            scaling_factor

    def fix(self):
        pass

    def unfix(self):
        pass


class FloatSoftmax(nn.Module):
    """layer_selection.py:41-68: output scale 2 / 2^bitwidth, codes in [0, 2^(bitwidth-1) - 1]."""

    def __init__(self, bitwidth=8, dim=-1, **kwargs):
        super().__init__()
        self.dim = dim
        self.bitwidth = bitwidth

    def forward(self, x, scaling_factor):
        x = _as_float(x)
        out = torch.softmax(x, dim=self.dim)
        s_out = torch.tensor(2.0 / 2 ** self.bitwidth, device=x.device)
        return _quantise_output(out, s_out, 0, 2 ** (self.bitwidth - 1) - 1), s_out

    def fix(self):
        pass

    def unfix(self):
        pass


class FloatLayerNorm(nn.Module):
    """layer_selection.py:71-113: output scale sqrt(C) / 2^30 * weight (per channel)."""

    def __init__(self, normalized_shape, eps=1e-5, bitwidth=8, **kwargs):
        super().__init__()
        self.layer_norm = nn.LayerNorm(normalized_shape, eps=eps)
        self.bitwidth = bitwidth
        self.normalized_shape = (normalized_shape,) if isinstance(normalized_shape, int) else normalized_shape
        self.dim_sqrt = None

    def forward(self, x, scaling_factor):
        x = _as_float(x)
        if self.dim_sqrt is None:
            self.dim_sqrt = torch.sqrt(torch.tensor(x.shape[-1], dtype=torch.float32, device=x.device))
        out = self.layer_norm(x)
        s_out = self.dim_sqrt / 2 ** 30 * self.layer_norm.weight
        return _quantise_output(out, s_out, -(2 ** (self.bitwidth - 1)), 2 ** (self.bitwidth - 1) - 1), s_out

    def fix(self):
        pass

    def unfix(self):
        pass

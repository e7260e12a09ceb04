"""Frozen-mode plumbing shared by the modules: running pending producers un-fused, dequantising a
carrier for user code, torch.cat of carriers, and host copies of scales.  Every compute call goes to
the C ABI (ivit_b200.ops); there is no CPU path."""
from __future__ import annotations

import numpy as np
import torch

from .. import _cabi, hostmath as hm, ops
from .carrier import Pending, QTensor

f32 = np.float32


def host_scale(t) -> np.ndarray:
    """Device scale tensor -> host float32 array (a synchronising copy: done once per frozen module,
    at its first call; afterwards constants are cached)."""
    if isinstance(t, np.ndarray):
        return t.astype(f32).reshape(-1)
    return t.detach().float().cpu().numpy().reshape(-1).astype(f32)


def require_cuda(t, what: str):
    dev = t.device
    if dev is None or dev.type != "cuda":
        raise _cabi.IvitKernelError(f"{what}: the frozen integer path runs on CUDA tensors only (no CPU fallback); "
                                    f"use running_stat=True / unfix() for the float-simulated calibration mode")


def eff_hscale(x: QTensor) -> np.ndarray:
    """Scale of the value the caller holds: carrier scale times the python-float factors it applied
    (float32 multiply, as `act_scaling_factor * self.scale` in vit_quant.py:75)."""
    h = x.hscale
    if x.factor != 1.0:
        h = (h * f32(x.factor)).astype(f32)
    return h


def run_pending(q: QTensor) -> torch.Tensor:
    """Launch a pending producer without fusing a consumer into it; returns canonical codes."""
    p = q.pending
    if p.kind == "quantize":
        return ops.quantize(p.x.contiguous(), float(p.inv), bits=p.bits)
    if p.kind == "gemm":
        return ops.gemm(p.A, p.lin["w"], _cabi.EPI_ACC32, bias=p.lin["bias"])
    if p.kind == "cat":
        return _cat_codes(q)
    raise NotImplementedError(
        f"a pending '{p.kind}' result was consumed by something other than the QuantAct that follows it in "
        f"vit_quant.py; its integer codes do not fit a storable carrier (DESIGN.md section 1)")


def dequantize(q: QTensor) -> torch.Tensor:
    """codes * scale as float32 -- what the reference's modules return.  For a pending GEMM this is
    the LOGITS epilogue (head: vit_quant.py:311)."""
    if q.codes is None and q.pending.kind == "gemm" and q.is_canonical():
        p = q.pending
        out = ops.gemm(p.A, p.lin["w"], _cabi.EPI_LOGITS, bias=p.lin["bias"], scale_out=p.lin["b_scale"])
        out = out.reshape(tuple(q.shape))
        return out * q.factor if q.factor != 1.0 else out
    codes = q.materialize()
    scale = q.scale
    if scale.numel() > 1 and scale.numel() == codes.shape[-1]:
        scale = scale.view(*([1] * (codes.dim() - 1)), -1)
    elif scale.numel() > 1 and codes.dim() == 4 and scale.numel() == codes.shape[1]:
        scale = scale.view(1, -1, 1, 1)
    out = codes.float() * scale
    return out * q.factor if q.factor != 1.0 else out


def cat(tensors, dim=0, **kw):
    """torch.cat over carriers / raw float tensors (vit_quant.py:293 concatenates the un-quantised
    class token with the patch-embedding carrier; both then share the carrier's scale)."""
    qs = [t for t in tensors if isinstance(t, QTensor)]
    ref = qs[0]
    shape = list(ref.shape)
    d = dim % len(shape)
    shape[d] = sum(int(t.shape[d]) for t in tensors)
    return QTensor(pending=Pending("cat", parts=list(tensors), dim=d), canon_shape=tuple(shape), scale=ref.scale,
                   hscale=ref.hscale, factor=ref.factor, device=ref.device)


def _cat_codes(q: QTensor) -> torch.Tensor:
    """int32 codes of a pending cat: carrier parts contribute their codes; raw float parts the integer
    the next QuantAct would recover, round(x / s) (quant_utils.py:220), un-clamped."""
    p = q.pending
    # carriers only (PatchMerging's four strided slices, swin_quant.py:339-344): the codes keep their storage type
    parts = [t.materialize() if isinstance(t, QTensor) else None for t in p.parts]
    kinds = {c.dtype for c in parts if c is not None}
    dtype = kinds.pop() if all(c is not None for c in parts) and len(kinds) == 1 else torch.int32
    out = torch.empty(q.canon_shape, dtype=dtype, device=q.device)
    off = 0
    s = q.scale
    for t, codes in zip(p.parts, parts):
        n = int(t.shape[p.dim])
        dst = out.narrow(p.dim, off, n)
        if isinstance(t, QTensor):
            if not np.array_equal(t.hscale, q.hscale) or t.factor != q.factor:
                raise NotImplementedError("torch.cat of carriers with different scales")
            dst.copy_(codes)
        else:
            dst.copy_(torch.round(t.detach().float() / s.view(*([1] * (t.dim() - 1)), -1)))
        off += n
    return out


def avgpool1d(x: QTensor, output_size):
    """nn.AdaptiveAvgPool1d(1) on the [B, C, L] transpose of an int8 token carrier (swin_quant.py:554): deferred, the
    QuantAct that follows launches ivit_avgpool_requant_i8 (torch's summation order over the tokens)."""
    one = output_size == 1 or (isinstance(output_size, (tuple, list)) and tuple(output_size) == (1,))
    if not (one and x.dim() == 3 and x.codes is not None and x.codes.dtype == torch.int8
            and x.transpose(1, 2).is_canonical()):
        return torch.nn.functional.adaptive_avg_pool1d(x.dequantize(), output_size)
    B, Cc, L = (int(v) for v in x.shape)
    return QTensor(pending=Pending("avgpool", x=x.codes.reshape(B, L, Cc), s_in_h=eff_hscale(x)), canon_shape=(B, Cc, 1),
                   scale=x.scale, hscale=x.hscale, factor=x.factor, device=x.device)

"""The `models/quantization_utils` module API on top of the sm_100a kernels.

Same class names, constructor signatures, buffer names (= checkpoint keys) and
``forward(x, scale, ...) -> (tensor, scale)`` convention as the reference (quant_modules.py,
ivit_modules.py, ibert_modules.py), so vit_quant.py / layers_quant.py / model_utils.py run unchanged.

Two modes, chosen per call:
  * float carrier in (a plain torch.Tensor = codes*scale)  -> float-simulated arithmetic (simq.py):
    calibration with ``running_stat=True``, CPU runs, anything before ``freeze_model``.
  * frozen QuantAct / QTensor carrier in                   -> integer codes on the GPU, kernels through
    the C ABI, producers fused into the QuantAct that consumes them (carrier.py).  CUDA only.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import _cabi, hostmath as hm, ops
from . import frozen as fz
from . import simq
from .carrier import Pending, QTensor

f32 = np.float32


def _dev(a, dtype, device):
    return ops.dev(np.asarray(a), dtype, device)


class _FrozenCache:
    """Per-module constants of the frozen path, built at the first frozen call (scales are fixed
    once every QuantAct is fixed) and dropped by unfix() / train() / load_state_dict()."""

    def _fz_get(self):
        return self.__dict__.setdefault("_fz_cache", {})

    def invalidate(self):
        self.__dict__["_fz_cache"] = {}

    def _load_from_state_dict(self, *a, **k):
        self.invalidate()
        return super()._load_from_state_dict(*a, **k)


# =============================================================================== QuantAct
class QuantAct(_FrozenCache, nn.Module):
    """quant_modules.py:229-387."""

    def __init__(self, activation_bit=8, act_range_momentum=0.95, running_stat=True, per_channel=False,
                 channel_len=None, quant_mode="symmetric"):
        super().__init__()
        self.activation_bit = activation_bit
        self.act_range_momentum = act_range_momentum
        self.running_stat = running_stat
        self.quant_mode = quant_mode
        self.percentile = None
        self.per_channel = per_channel
        n = 1 if not per_channel else channel_len
        if per_channel:
            assert channel_len is not None
        self.register_buffer("x_min", torch.zeros(n))
        self.register_buffer("x_max", torch.zeros(n))
        self.register_buffer("act_scaling_factor", torch.zeros(n))
        if quant_mode == "asymmetric":
            raise NotImplementedError("unsupported quant mode: asymmetric")
        if quant_mode not in ("none", "symmetric"):
            raise ValueError(f"unknown quant mode: {quant_mode}")

    def __repr__(self):
        return (f"{self.__class__.__name__}(activation_bit={self.activation_bit}, quant_mode: {self.quant_mode}, "
                f"Act_min: {self.x_min.min().item():.2f}, Act_max: {self.x_max.max().item():.2f})")

    def fix(self):
        self.running_stat = False
        self.invalidate()

    def unfix(self):
        self.running_stat = True
        self.invalidate()

    # ------------------------------------------------------------------ calibration (float carrier)
    def _track_range(self, x_act):
        """quant_modules.py:310-360: min/max (or percentile) + EMA with first-call initialisation."""
        d = x_act.data
        if self.percentile is None:
            if not self.per_channel:
                lo, hi = d.min(), d.max()
            else:
                lo = d.min(axis=0).values.min(axis=0).values
                hi = d.max(axis=0).values.max(axis=0).values
        else:
            p_lo = (100 - self.percentile) / 2
            p_hi = 100 - p_lo
            with torch.no_grad():
                if not self.per_channel:
                    flat = d.flatten()
                    lo, hi = torch.quantile(flat, p_lo / 100.0), torch.quantile(flat, p_hi / 100.0)
                else:
                    r = d.reshape(d.shape[0], d.shape[1], -1)
                    lo = torch.stack([torch.quantile(r[:, c, :].flatten(), p_lo / 100.0) for c in range(r.shape[1])])
                    hi = torch.stack([torch.quantile(r[:, c, :].flatten(), p_hi / 100.0) for c in range(r.shape[1])])
        if torch.eq(self.x_min, self.x_max).all():
            self.x_min = self.x_min + lo
            self.x_max = self.x_max + hi
        elif self.act_range_momentum == -1:
            self.x_min = torch.min(self.x_min, lo)
            self.x_max = torch.max(self.x_max, hi)
        else:
            m = self.act_range_momentum
            self.x_min = self.x_min * m + lo * (1 - m)
            self.x_max = self.x_max * m + hi * (1 - m)

    def _forward_sim(self, x, s_pre, identity, s_id, lo, hi):
        x_act = x if identity is None else identity + x
        if self.running_stat:
            self._track_range(x_act)
        if self.quant_mode == "none":
            return x_act, None
        lo = self.x_min if lo is None else lo
        hi = self.x_max if hi is None else hi
        self.act_scaling_factor = simq.sym_scale(self.activation_bit, lo, hi)
        if s_pre is None:
            q = simq.sym_quantize(x, self.activation_bit, self.act_scaling_factor, False)
        else:
            q = simq.fixedpoint_mul(x, s_pre, self.activation_bit, self.act_scaling_factor, identity, s_id)
        return q * self.act_scaling_factor.view(-1), self.act_scaling_factor

    # ------------------------------------------------------------------ frozen (integer carrier)
    def _out_scale(self, device):
        c = self._fz_get()
        if "s_out" not in c:
            if self.per_channel:
                raise NotImplementedError("frozen per-channel QuantAct")
            h = hm.sym_scale(self.activation_bit, fz.host_scale(self.x_min), fz.host_scale(self.x_max)).reshape(1)
            c["s_out_h"] = h
            c["s_out"] = _dev(h, torch.float32, device)
            self.act_scaling_factor = c["s_out"]
            c["q"] = hm.qrange(self.activation_bit)
        return c["s_out"], c["s_out_h"], c["q"]

    def _store_bits(self) -> int:
        """int8 codes up to 8 bits, int16 codes for the 16-bit locations (vit_quant.py:180-187)."""
        if self.activation_bit > 16:
            raise NotImplementedError("frozen QuantAct: at most 16-bit activations")
        return 16 if self.activation_bit > 8 else 8

    def _wrap(self, codes, shape, device, layout=0):
        s, h, _ = self._out_scale(device)
        q = QTensor(codes, scale=s, hscale=h, layout=layout)
        return q.reshape(*shape), s

    def _mult(self, key, s_in_h, device):
        """float64 dyadic multiplier(s) m*2^-e of s_in / s_out (batch_frexp on the host, once)."""
        c = self._fz_get()
        if key not in c:
            _, h, _ = self._out_scale(device)
            m = hm.dyadic_mult(np.asarray(s_in_h, dtype=f32).reshape(-1), h)
            c[key] = (_dev(m.reshape(-1), torch.float64, device), m.reshape(-1))
        return c[key]

    def _forward_frozen(self, x, s_pre, identity, s_id):
        device = x.device
        fz.require_cuda(x, "QuantAct")
        s_out, s_out_h, (qmin, qmax) = self._out_scale(device)
        if not isinstance(x, QTensor):
            # quant_modules.py:374-377: no predecessor scale -> plain symmetric quantiser (deferred:
            # a following QuantConv2d fuses it with its im2col)
            if s_pre is not None:
                # a float module (the `float` registry entries) feeds this QuantAct: quant_utils.py:220
                # z = round(x / s_pre) with torch ops (IEEE division + rint: same codes as on the CPU),
                # then the integer requantisation kernel (identity branch included)
                sp = s_pre.to(device).float()
                z = torch.round(x.detach().float() / (sp.view(1, -1, 1, 1) if x.dim() == 4 and sp.numel() > 1 else sp))
                z = z.to(torch.int32).contiguous()
                s_in_h = fz.host_scale(sp)
                mult, _ = self._mult(("req", s_in_h.tobytes()), s_in_h, device)
                s_in_d = self._fz_get().setdefault(("s_in", s_in_h.tobytes()), _dev(s_in_h, torch.float32, device))
                idc, mult_id = None, 0.0
                if identity is not None:
                    if not isinstance(identity, QTensor):
                        raise _cabi.IvitKernelError("frozen QuantAct: identity must come from a frozen module")
                    idc = identity.materialize().contiguous()
                    _, mh = self._mult(("id", fz.eff_hscale(identity).tobytes()), fz.eff_hscale(identity), device)
                    mult_id = float(mh[0])
                z2 = z.reshape(-1, z.shape[-1])
                out = ops.requant(z2, s_in_d, mult, qmin, qmax, idc.reshape(-1, z.shape[-1]) if idc is not None else None,
                                  mult_id, out_bits=self._store_bits())
                return self._wrap(out.reshape(z.shape), tuple(x.shape), device)
            inv = f32(1.0) / s_out_h[0]
            q = QTensor(pending=Pending("quantize", x=x.detach(), inv=inv, bits=self.activation_bit),
                        canon_shape=tuple(x.shape), scale=s_out, hscale=s_out_h, device=device)
            return q, s_out
        p = x.pending if x.codes is None else None
        kind = p.kind if p is not None else None
        shape = tuple(x.shape)
        # memory layout of the reference's result (hazard H11): elementwise ops keep their input's layout, and
        # fixedpoint_mul adds the identity branch first (quant_utils.py:245), so the identity's layout wins
        lay = identity.layout if isinstance(identity, QTensor) else x.layout
        if identity is None and kind == "gemm" and x.is_canonical() and self.activation_bit <= 16:
            # QuantLinear / QuantConv2d -> QuantAct: requantisation fused into the GEMM epilogue
            mult, mh = self._mult(("gemm", p.lin["b_scale_h"].tobytes()), p.lin["b_scale_h"], device)   # keyed by value, not id()
            out = ops.gemm(p.A, p.lin["w"], _cabi.EPI_REQUANT, bias=p.lin["bias"], mult=mult, s_in=p.lin["b_scale"],
                           qmin=qmin, qmax=qmax, z_bound_ok=p.lin["z_bound_ok"],
                           force_slow=p.lin["force_slow"] or bool(np.abs(mh).max() >= 256.0), out_bits=self._store_bits())
            return self._wrap(out, shape, device, lay)
        if identity is None and kind == "ln":
            return self._wrap(p.mod._frozen_ln_requant(p, x, self, device), shape, device, lay)
        if identity is None and kind == "gelu":
            return self._wrap(p.mod._frozen_gelu_requant(p, self, device), shape, device)
        if identity is None and kind == "mm2":
            return self._wrap(_frozen_attention(p, x, self, device), shape, device)
        if identity is None and kind == "avgpool" and self.activation_bit <= 8:
            # avgpool(x.transpose(1, 2)) -> qact3 (swin_quant.py:554-555): one kernel, torch's summation order
            _, mh = self._mult(("avgpool", p.s_in_h.tobytes()), p.s_in_h, device)
            B, L, C = (int(v) for v in p.x.shape)
            out = ops.avgpool_requant(p.x, B, L, C, float(p.s_in_h[0]), float(mh[0]), qmin, qmax)
            return self._wrap(out, shape, device)
        if kind == "qattn" and isinstance(identity, QTensor):
            # WindowAttention.qact2(attn, s, relative_position_bias, s_table) (swin_quant.py:147): stays symbolic
            q = QTensor(pending=Pending("qattn_bias", qattn=p, s_in_h=fz.eff_hscale(x), qact=self, identity=identity,
                                        mask=None), canon_shape=x.canon_shape, scale=s_out, hscale=s_out_h, meta=x.meta,
                        device=device)
            return q, s_out
        if identity is None and kind == "mm1":
            # qact_attn1: stays symbolic, the softmax / second matmul / qact2 complete the chain
            q = QTensor(pending=Pending("qattn", mm1=p, s_in_h=fz.eff_hscale(x), qact=self), canon_shape=x.canon_shape,
                        scale=s_out, hscale=s_out_h, meta=x.meta, device=device)
            return q, s_out
        # general stand-alone requantisation of materialised codes (+ identity branch)
        z = x.materialize()
        if not z.is_contiguous():
            z = z.contiguous()
        if z.dtype not in (torch.int8, torch.int16, torch.int32):
            z = z.to(torch.int32)
        s_in_h = fz.eff_hscale(x)
        mult, _ = self._mult(("req", s_in_h.tobytes()), s_in_h, device)
        s_in_d = self._fz_get().setdefault(("s_in", s_in_h.tobytes()), _dev(s_in_h, torch.float32, device))
        idc, mult_id = None, 0.0
        if identity is not None:
            if not isinstance(identity, QTensor):
                raise _cabi.IvitKernelError("frozen QuantAct: identity must come from a frozen module")
            idc = identity.materialize()
            if idc.dtype not in (torch.int8, torch.int16):
                raise NotImplementedError("identity branch wider than 16 bits")
            if tuple(idc.shape) != tuple(z.shape):
                if idc.shape[-1] != z.shape[-1] or z.numel() % idc.numel() != 0:
                    raise NotImplementedError("identity broadcast other than over leading dims")
            idc = idc.contiguous()
            _, mh = self._mult(("id", fz.eff_hscale(identity).tobytes()), fz.eff_hscale(identity), device)
            mult_id = float(mh[0])
        out = ops.requant(z, s_in_d, mult, qmin, qmax, idc, mult_id, out_bits=self._store_bits())
        return self._wrap(out, shape, device, lay)

    def forward(self, x, pre_act_scaling_factor=None, identity=None, identity_scaling_factor=None,
                specified_min=None, specified_max=None):
        # fixed + symmetric = the integer path (CUDA only: _forward_frozen refuses CPU tensors)
        frozen_call = isinstance(x, QTensor) or (
            not self.running_stat and self.quant_mode == "symmetric" and specified_min is None
            and specified_max is None)
        if isinstance(x, QTensor) and (self.running_stat or self.quant_mode != "symmetric"):
            x = x.dequantize()
            identity = identity.dequantize() if isinstance(identity, QTensor) else identity
            frozen_call = False
        if frozen_call:
            return self._forward_frozen(x, pre_act_scaling_factor, identity, identity_scaling_factor)
        if isinstance(identity, QTensor):
            identity = identity.dequantize()
        return self._forward_sim(x, pre_act_scaling_factor, identity, identity_scaling_factor, specified_min,
                                 specified_max)


# =============================================================================== QuantLinear / Conv / MatMul
def _linear_constants(mod, weight, bias, s_x_h, device, bias_bit=32):
    """Frozen QuantLinear / QuantConv2d constants (quant_modules.py:202-220 / :486-504), built once:
    int8 weights [N, K], int32 bias, per-channel scale, and the weights-only |acc+bias| bound."""
    p = hm.linear_params(weight.detach().float().cpu().numpy(),
                         None if bias is None else bias.detach().float().cpu().numpy(), s_x_h)
    zb = np.abs(p["w_int_f32"]).sum(axis=1).max() * 128.0 + np.abs(p["b_int"].astype(np.float64)).max()
    lin = {"w": _dev(p["w_int"], torch.int8, device), "bias": _dev(p["b_int"], torch.int32, device),
           "b_scale": _dev(p["b_scale"], torch.float32, device), "b_scale_h": p["b_scale"],
           "z_bound_ok": bool(zb < 2 ** 22), "force_slow": bool(np.abs(p["b_int_f32"]).max() >= 2 ** 30)
           if p["b_int_f32"] is not None else False}
    # keep the reference's derived buffers observable (state_dict: TVM_benchmark/convert_model.py:12-66)
    mod.weight_integer = _dev(p["w_int_f32"].reshape(weight.shape), torch.float32, device)
    if bias is not None:
        mod.bias_integer = _dev(p["b_int_f32"], torch.float32, device)
    return lin, p


class QuantLinear(_FrozenCache, nn.Linear):
    """quant_modules.py:131-226."""

    def __init__(self, in_features, out_features, bias=True, weight_bit=8, bias_bit=32, per_channel=True,
                 quant_mode="symmetric"):
        super().__init__(in_features, out_features, bias)
        self.weight_bit, self.bias_bit = weight_bit, bias_bit
        self.per_channel, self.quant_mode = per_channel, quant_mode
        if quant_mode == "asymmetric":
            raise NotImplementedError("unsupported quant mode: asymmetric")
        if quant_mode not in ("none", "symmetric"):
            raise ValueError(f"unknown quant mode: {quant_mode}")
        self.register_buffer("fc_scaling_factor", torch.zeros(self.out_features))
        self.register_buffer("weight_integer", torch.zeros_like(self.weight))
        if self.bias is not None:
            self.register_buffer("bias_integer", torch.zeros_like(self.bias))

    def __repr__(self):
        return f"({super().__repr__()} weight_bit={self.weight_bit}, quant_mode={self.quant_mode})"

    def fix(self):
        pass

    def unfix(self):
        self.invalidate()

    def forward(self, x, prev_act_scaling_factor=None):
        if self.quant_mode == "none":
            x = x.dequantize() if isinstance(x, QTensor) else x
            return F.linear(x, weight=self.weight, bias=self.bias), None
        assert prev_act_scaling_factor is not None and prev_act_scaling_factor.shape == (1,)
        if isinstance(x, QTensor):
            return self._forward_frozen(x)
        # float-simulated (quant_modules.py:202-226); bias=None handled (the reference crashes)
        w = self.weight.data.detach()
        if self.per_channel:
            w_min, w_max = w.min(dim=1)[0], w.max(dim=1)[0]
        else:
            w_min, w_max = w.min().expand(1), w.max().expand(1)
        self.fc_scaling_factor = simq.sym_scale(self.weight_bit, w_min, w_max)
        self.weight_integer = simq.sym_quantize(self.weight, self.weight_bit, self.fc_scaling_factor, True)
        b_scale = self.fc_scaling_factor * prev_act_scaling_factor
        b_int = None
        if self.bias is not None:
            self.bias_integer = b_int = simq.sym_quantize(self.bias, self.bias_bit, b_scale, True)
        x_int = x / prev_act_scaling_factor.view(1, -1)
        return F.linear(x_int, weight=self.weight_integer, bias=b_int) * b_scale, b_scale

    def _forward_frozen(self, x: QTensor):
        fz.require_cuda(x, "QuantLinear")
        if self.weight_bit != 8 or not self.per_channel:
            raise NotImplementedError("frozen QuantLinear: per-channel 8-bit weights")
        c = self._fz_get()
        s_x = np.asarray(fz.eff_hscale(x), dtype=f32)
        if "lin" in c and not np.array_equal(c["lin_s_x"], s_x):
            c.pop("lin")          # a different producer scale than the constants were built for: rebuild, never reuse
        if "lin" not in c:
            c["lin"], p = _linear_constants(self, self.weight, self.bias, s_x, x.device, self.bias_bit)
            c["lin_s_x"] = s_x.copy()
            self.fc_scaling_factor = _dev(p["s_w"], torch.float32, x.device)
        lin = c["lin"]
        A = x.materialize()
        if A.dtype != torch.int8:
            raise NotImplementedError("QuantLinear input wider than 8 bits")
        lead = tuple(A.shape[:-1])
        A2 = A.reshape(-1, A.shape[-1])
        if A2.stride(-1) != 1 or (A2.stride(0) % 16) != 0:
            A2 = A2.contiguous()
        q = QTensor(pending=Pending("gemm", A=A2, lin=lin), canon_shape=lead + (self.out_features,),
                    scale=lin["b_scale"], hscale=lin["b_scale_h"], device=x.device)
        return q, lin["b_scale"]


class QuantConv2d(_FrozenCache, nn.Conv2d):
    """quant_modules.py:412-512."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, dilation=1, groups=1, bias=True,
                 weight_bit=8, bias_bit=32, quant_mode="symmetric", per_channel=True, weight_percentile=0):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, bias)
        self.weight_bit, self.bias_bit = weight_bit, bias_bit
        self.quant_mode, self.per_channel, self.weight_percentile = quant_mode, per_channel, weight_percentile
        self.register_buffer("conv_scaling_factor", torch.zeros(self.out_channels))
        self.register_buffer("weight_integer", torch.zeros_like(self.weight))
        if self.bias is not None:
            self.register_buffer("bias_integer", torch.zeros_like(self.bias))

    def __repr__(self):
        return f"({super().__repr__()} weight_bit={self.weight_bit}, quant_mode={self.quant_mode})"

    def fix(self):
        pass

    def unfix(self):
        self.invalidate()

    def forward(self, x, pre_act_scaling_factor=None):
        if self.quant_mode == "asymmetric":
            raise NotImplementedError("unsupported quant mode: asymmetric")
        if self.quant_mode != "symmetric":
            raise ValueError(f"unknown quant mode: {self.quant_mode}")
        if not self.per_channel:
            raise Exception("For weight, we only support per_channel quantization.")
        if isinstance(x, QTensor):
            return self._forward_frozen(x)
        with torch.no_grad():
            v = self.weight.reshape(self.weight.shape[0], -1)
            self.conv_scaling_factor = simq.sym_scale(self.weight_bit, v.min(axis=1).values, v.max(axis=1).values)
        self.weight_integer = simq.sym_quantize(self.weight, self.weight_bit, self.conv_scaling_factor, True)
        b_scale = self.conv_scaling_factor * pre_act_scaling_factor
        b_int = None
        if self.bias is not None:
            self.bias_integer = b_int = simq.sym_quantize(self.bias, self.bias_bit, b_scale, True)
        x_int = x / pre_act_scaling_factor.view(1, -1, 1, 1)
        out_scale = b_scale.view(1, -1, 1, 1)
        return (F.conv2d(x_int, self.weight_integer, b_int, self.stride, self.padding, self.dilation, self.groups)
                * out_scale, out_scale)

    def _forward_frozen(self, x: QTensor):
        """Patch embedding: kernel == stride, no padding -> im2col is a pure re-indexing and the
        convolution is a GEMM [B*gh*gw, Cin*P*P] x [Cout, Cin*P*P]^T."""
        fz.require_cuda(x, "QuantConv2d")
        P = self.kernel_size[0]
        if (self.kernel_size != self.stride or self.kernel_size[0] != self.kernel_size[1] or any(self.padding)
                or self.groups != 1 or any(d != 1 for d in self.dilation)):
            raise NotImplementedError("frozen QuantConv2d: kernel == stride, no padding (ViT/Swin patch embedding)")
        c = self._fz_get()
        s_x = np.asarray(fz.eff_hscale(x), dtype=f32)
        if "lin" in c and not np.array_equal(c["lin_s_x"], s_x):
            c.pop("lin")          # constants were built for another input scale: rebuild
        if "lin" not in c:
            c["lin"], p = _linear_constants(self, self.weight, self.bias, s_x, x.device, self.bias_bit)
            c["lin_s_x"] = s_x.copy()
            self.conv_scaling_factor = _dev(p["s_w"], torch.float32, x.device)
        lin = c["lin"]
        B, Cin, H, W = (int(v) for v in x.shape)
        gh, gw = H // P, W // P
        if x.codes is None and x.pending.kind == "quantize" and x.is_canonical() and x.pending.bits == 8 and P % 4 == 0:
            patches = ops.quantize_patches(x.pending.x.contiguous(), float(x.pending.inv), P)   # fused quantiser
        else:
            t = x.materialize()
            patches = t.reshape(B, Cin, gh, P, gw, P).permute(0, 2, 4, 1, 3, 5).reshape(B * gh * gw, Cin * P * P)
            patches = patches.contiguous()
        Cout = self.out_channels
        # canonical storage [B, gh, gw, Cout]; the module's logical output is [B, Cout, gh, gw]
        meta = torch.empty((B, gh, gw, Cout), device="meta", dtype=torch.int8).permute(0, 3, 1, 2)
        scale4 = lin["b_scale"].view(1, -1, 1, 1)
        # layout: the reference's NCHW result read as [B, L, C] tokens is channel-strided (layers_quant.py:198)
        q = QTensor(pending=Pending("gemm", A=patches, lin=lin), canon_shape=(B, gh, gw, Cout), scale=lin["b_scale"],
                    hscale=lin["b_scale_h"], meta=meta, device=x.device, layout=gh * gw)
        return q, scale4


class QuantMatMul(nn.Module):
    """quant_modules.py:390-409."""

    def __init__(self):
        super().__init__()
        self.register_buffer("act_scaling_factor", torch.zeros(1))

    def fix(self):
        pass

    def unfix(self):
        pass

    def forward(self, A, pre_act_scaling_factor_A, B, pre_act_scaling_factor_B):
        if isinstance(A, QTensor) or isinstance(B, QTensor):
            if not (isinstance(A, QTensor) and isinstance(B, QTensor)):
                raise _cabi.IvitKernelError("QuantMatMul: both operands must come from frozen integer modules (the `float` "
                                            "softmax entry runs in calibration mode only: its product is a float GEMM)")
            fz.require_cuda(A, "QuantMatMul")
            sA, sB = fz.eff_hscale(A), fz.eff_hscale(B)
            h = (sA * sB).astype(f32)
            scale = pre_act_scaling_factor_A * pre_act_scaling_factor_B
            self.act_scaling_factor = scale
            shape = tuple(A.shape[:-1]) + (int(B.shape[-1]),)
            if A.codes is None and A.pending is not None and A.pending.kind == "softmax":
                # attn @ v: canonical storage [B, N, H, hd] so that transpose(1,2).reshape(B,N,C) is a view
                Bb, Hh, Nn, hd = shape
                meta = torch.empty((Bb, Nn, Hh, hd), device="meta", dtype=torch.int8).permute(0, 2, 1, 3)
                q = QTensor(pending=Pending("mm2", P=A.pending, v=B, s_h=h), canon_shape=(Bb, Nn, Hh, hd), scale=scale,
                            hscale=h, meta=meta, device=A.device)
            else:
                q = QTensor(pending=Pending("mm1", a=A, b=B), canon_shape=shape, scale=scale, hscale=h,
                            device=A.device)
            return q, scale
        A_int = A / pre_act_scaling_factor_A
        B_int = B / pre_act_scaling_factor_B
        scale = pre_act_scaling_factor_A * pre_act_scaling_factor_B
        self.act_scaling_factor = scale
        return (A_int @ B_int) * scale, scale


# =============================================================================== LayerNorm
class _IntLayerNormBase(_FrozenCache):
    _kind = 0

    def _frozen_forward(self, x: QTensor):
        """IntLayerNorm on an int8 carrier: deferred -- the kernel is the fused LayerNorm + QuantAct
        (the only way the reference uses it: norm -> qact, vit_quant.py:143-144,148-149,301-304)."""
        fz.require_cuda(x, type(self).__name__)
        C = int(x.shape[-1])
        c = self._fz_get()
        if "sf" not in c:
            w = self.weight.detach().float().cpu().numpy()
            b = self.bias.detach().float().cpu().numpy()
            p = hm.layernorm_params(w, b, C)
            c["sf_h"], c["sf"] = p["sf"], _dev(p["sf"], torch.float32, x.device)
            c["bias_int"] = _dev(p["bias_int"], torch.float32, x.device)
            if hasattr(self, "norm_scaling_factor"):
                self.norm_scaling_factor = c["sf"]
                self.bias_integer = c["bias_int"]
        s_in = fz.eff_hscale(x)
        q = QTensor(pending=Pending("ln", x=x, mod=self, s_in=float(s_in[0])), canon_shape=tuple(x.shape),
                    scale=c["sf"], hscale=c["sf_h"], device=x.device, layout=x.layout)
        return q, c["sf"]

    def _frozen_ln_requant(self, p, view: QTensor, qact: "QuantAct", device):
        c = self._fz_get()
        codes = p.x.materialize()
        if codes.dtype not in (torch.int8, torch.int16):
            raise NotImplementedError("frozen IntLayerNorm: int8 or int16 input codes")
        if not codes.is_contiguous():
            codes = codes.contiguous()
        C = codes.shape[-1]
        rows = view.strided(codes)                       # the caller may have sliced rows (x[:, 0])
        if rows.stride(-1) != 1 or rows.shape[-1] != C:
            raise NotImplementedError("IntLayerNorm output sliced along the channel dimension")
        rows2 = rows.reshape(-1, C) if rows.is_contiguous() else rows
        if rows2.dim() != 2:
            rows2 = rows.contiguous().reshape(-1, C)
        mult, _ = qact._mult(("ln", id(self)), c["sf_h"], device)
        _, _, (qmin, qmax) = qact._out_scale(device)
        if p.x.layout:
            # the reference's input is channel-strided (Swin stage 0, hazard H11): outer-reduction summation order
            if self._kind != 0 or not view.is_canonical() or rows.dim() != 3 or int(rows.shape[-2]) != p.x.layout:
                raise NotImplementedError("IntLayerNorm on a channel-strided carrier: ivit LayerNorm over whole [B, L, C]")
            return ops.layernorm_requant_strided(0, rows.reshape(-1, C), p.x.layout, p.s_in, c["bias_int"], c["sf"], mult,
                                                 qmin, qmax)
        shift = float(self.shift.item()) if self._kind == 1 and "shift_h" not in c else c.get("shift_h", 0.0)
        c["shift_h"] = shift
        if rows2.dtype == torch.int8 and C not in (96, 128, 192, 384, 768, 1024):
            # the int8 row kernel is specialised per width; other widths (Swin-B stage 1: 256, a merged 512) take
            # the 16-bit-code kernel, same integer arithmetic on widened codes (what FrozenSwinPlan's stream does)
            rows2 = rows2.to(torch.int16)
        return ops.layernorm_requant(self._kind, rows2, p.s_in, c["bias_int"], c["sf"], mult, qmin, qmax, shift=shift,
                                     use_int_sqrt=bool(getattr(self, "use_int_sqrt", False)))


class IVITIntLayerNorm(_IntLayerNormBase, nn.LayerNorm):
    """ivit_modules.py:10-65."""
    _kind = 0

    def __init__(self, normalized_shape, eps=1e-6, elementwise_affine=True):
        super().__init__(normalized_shape, eps, elementwise_affine)
        self.dim_sqrt = None
        self.register_buffer("norm_scaling_factor", torch.zeros(normalized_shape))
        self.register_buffer("bias_integer", torch.zeros_like(self.bias))

    def fix(self):
        pass

    def unfix(self):
        self.invalidate()

    def forward(self, x, scaling_factor=None):
        if isinstance(x, QTensor):
            return self._frozen_forward(x)
        if self.dim_sqrt is None or self.dim_sqrt.device != x.device:
            self.dim_sqrt = torch.sqrt(torch.tensor(x.shape[2], dtype=torch.float, device=x.device))
        y, sf, bias_int = simq.ivit_layernorm(x, scaling_factor, self.weight, self.bias, self.dim_sqrt)
        self.bias_integer = bias_int
        self.norm_scaling_factor = sf
        return y, sf


class IBERTIntLayerNorm(_IntLayerNormBase, nn.Module):
    """ibert_modules.py:12-158."""
    _kind = 1

    def __init__(self, normalized_shape, output_bit=8, overflow_handling=True, quant_mode="symmetric",
                 force_dequant="none", elementwise_affine=True, eps=1e-5, use_int_sqrt=False):
        super().__init__()
        self.quant_mode = "none" if force_dequant in ("nonlinear", "layernorm") else quant_mode
        if quant_mode == "asymmetric":
            raise NotImplementedError("unsupported quant mode: asymmetric")
        if quant_mode not in ("none", "symmetric"):
            raise ValueError(f"unknown quant mode: {quant_mode}")
        self.overflow_handling = overflow_handling
        self.use_int_sqrt = use_int_sqrt
        self.register_buffer("shift", torch.zeros(1))
        self.output_bit = output_bit
        self.dim_sqrt = None
        if isinstance(normalized_shape, int):
            normalized_shape = (normalized_shape,)
        self.normalized_shape = normalized_shape
        self.weight = nn.Parameter(torch.ones(self.normalized_shape))
        self.bias = nn.Parameter(torch.zeros(self.normalized_shape))
        self.eps = eps

    def fix(self):
        self.overflow_handling = False

    def unfix(self):
        self.overflow_handling = True
        self.invalidate()

    def _overflow_fallback(self, y_int):
        """ibert_modules.py:69-83: raise `shift` until the sum of squares fits 32 bits."""
        with torch.no_grad():
            var = torch.sum(y_int ** 2, axis=2, keepdim=True)
            shift = (torch.log2(torch.sqrt(var / 2 ** 32)).ceil()).max()
            self.shift = torch.max(self.shift, shift)
        y_sh = torch.floor(y_int / 2 ** self.shift)
        return torch.sum(y_sh ** 2, axis=2, keepdim=True)

    def forward(self, x, scaling_factor=None):
        if self.quant_mode == "none":
            x = x.dequantize() if isinstance(x, QTensor) else x
            mean = x.mean(axis=2, keepdim=True)
            y = x - mean
            var = torch.mean(y ** 2, axis=2, keepdim=True)
            return (y / torch.sqrt(self.eps + var)) * self.weight + self.bias, None
        if isinstance(x, QTensor):
            return self._frozen_forward(x)
        if self.dim_sqrt is None or self.dim_sqrt.device != x.device:
            self.dim_sqrt = torch.sqrt(torch.tensor(x.shape[2], dtype=torch.float, device=x.device))
        cb = self._overflow_fallback if (self.overflow_handling and self.training) else None
        return simq.ibert_layernorm(x, scaling_factor, self.weight, self.bias, self.dim_sqrt, self.shift,
                                    self.use_int_sqrt, cb)


# =============================================================================== GELU
class _IntGELUBase(_FrozenCache):
    _kind = 0

    def _frozen_forward(self, x: QTensor):
        fz.require_cuda(x, type(self).__name__)
        s = fz.eff_hscale(x)
        c = self._fz_get()
        if "s_out" not in c:
            s0 = f32(s[0])
            if self._kind == 0:
                h = np.array([s0 * f32(1 / 2 ** (self.output_bit - 1))], dtype=f32)      # ivit_modules.py:121-124
            else:
                k = f32(s0 / f32(1.4142))
                s_erf = f32(f32(f32(k * k) * f32(-0.2888)) * f32(64.0))
                h = np.array([f32(f32(s0 * s_erf) / f32(2.0))], dtype=f32)               # ibert_modules.py:233
            c["s_out_h"], c["s_out"] = h, _dev(h, torch.float32, x.device)
            c["s_in"] = float(s0)
        codes = x.materialize()
        if codes.dtype != torch.int8:
            raise NotImplementedError("frozen IntGELU: 8-bit input codes")
        q = QTensor(pending=Pending("gelu", x=codes, mod=self), canon_shape=tuple(x.shape), scale=c["s_out"],
                    hscale=c["s_out_h"], device=x.device)
        return q, c["s_out"]

    def _frozen_gelu_requant(self, p, qact: "QuantAct", device):
        """mlp.act + mlp.qact1 as one table lookup per element; the (row max, code) table is built
        on the device by the exact arithmetic when this (GELU, QuantAct) pair first meets."""
        c = self._fz_get()
        key = ("lut", id(qact))
        if key not in c:
            _, mh = qact._mult(("gelu", id(self)), c["s_out_h"], device)
            _, _, (qmin, qmax) = qact._out_scale(device)
            lut, s_dev = ops.gelu_lut_build(self._kind, c["s_in"], getattr(self, "output_bit", 8), float(mh[0]), qmin,
                                            qmax, device)
            if f32(s_dev) != c["s_out_h"][0]:
                raise _cabi.IvitKernelError("GELU output scale: host and library disagree")
            c[key] = lut
        x = p.x if p.x.is_contiguous() else p.x.contiguous()
        return ops.gelu_lut_apply(x, c[key])


class IVITIntGELU(_IntGELUBase, nn.Module):
    """ivit_modules.py:68-126 (ShiftGELU)."""
    _kind = 0

    def __init__(self, output_bit=8, n=23):
        super().__init__()
        self.output_bit, self.n = output_bit, n
        self.register_buffer("act_scaling_factor", torch.zeros(1))

    def fix(self):
        pass

    def unfix(self):
        self.invalidate()

    def forward(self, x, scaling_factor=None):
        if isinstance(x, QTensor):
            if self.n != 23:
                raise NotImplementedError("frozen ShiftGELU: n = 23")
            return self._frozen_forward(x)
        y, s = simq.ivit_gelu(x, scaling_factor, self.output_bit, self.n)
        self.act_scaling_factor = s
        return y, s


class IBERTIntGELU(_IntGELUBase, nn.Module):
    """ibert_modules.py:161-235."""
    _kind = 1

    def __init__(self, quant_mode="symmetric", force_dequant="none"):
        super().__init__()
        self.register_buffer("input_scaling_factor", torch.ones(1))
        self.quant_mode = "none" if force_dequant in ("nonlinear", "gelu") else quant_mode
        if quant_mode == "asymmetric":
            raise NotImplementedError("unsupported quant mode: asymmetric")
        if quant_mode not in ("none", "symmetric"):
            raise ValueError(f"unknown quant mode: {quant_mode}")
        if self.quant_mode == "none":
            self.activation_fn = nn.GELU()
        self.k, self.n = 1.4142, 6

    def fix(self):
        pass

    def unfix(self):
        self.invalidate()

    def forward(self, x, scaling_factor=None):
        if self.quant_mode == "none":
            x = x.dequantize() if isinstance(x, QTensor) else x
            return self.activation_fn(x), None
        if isinstance(x, QTensor):
            return self._frozen_forward(x)
        return simq.ibert_gelu(x, scaling_factor)


# =============================================================================== Softmax
class _IntSoftmaxBase(_FrozenCache):
    _kind = 0

    def _frozen_forward(self, x: QTensor):
        """Deferred: the softmax is computed on the score tile inside the attention kernel."""
        fz.require_cuda(x, type(self).__name__)
        if x.codes is not None or x.pending.kind not in ("qattn", "qattn_bias"):
            raise NotImplementedError("frozen IntSoftmax outside the attention chains of vit_quant.py:72-82 / "
                                      "swin_quant.py:131-156")
        if not 2 <= self.output_bit <= 16:
            raise NotImplementedError("frozen IntSoftmax: output_bit in [2, 16]")
        c = self._fz_get()
        if "s_out" not in c:
            v = 1 / 2 ** (self.output_bit - 1) if self._kind == 0 else 2 / 2 ** self.output_bit
            c["s_out_h"] = np.array([v], dtype=f32)
            c["s_out"] = _dev(c["s_out_h"], torch.float32, x.device)
        q = QTensor(pending=Pending("softmax", qattn=x.pending, mod=self), canon_shape=x.canon_shape, scale=c["s_out"],
                    hscale=c["s_out_h"], meta=x.meta, device=x.device)
        return q, c["s_out"]


class IVITIntSoftmax(_IntSoftmaxBase, nn.Module):
    """ivit_modules.py:129-179 (Shiftmax)."""
    _kind = 0

    def __init__(self, output_bit=8):
        super().__init__()
        self.output_bit, self.n = output_bit, 15
        self.register_buffer("act_scaling_factor", torch.zeros(1))

    def fix(self):
        pass

    def unfix(self):
        self.invalidate()

    def forward(self, x, scaling_factor):
        if isinstance(x, QTensor):
            return self._frozen_forward(x)
        y, s = simq.ivit_softmax(x, scaling_factor, self.output_bit, self.n)
        self.act_scaling_factor = s
        return y, s


class IBERTIntSoftmax(_IntSoftmaxBase, nn.Module):
    """ibert_modules.py:237-319."""
    _kind = 1

    def __init__(self, output_bit, quant_mode="symmetric", force_dequant="none"):
        super().__init__()
        self.output_bit = output_bit
        self.quant_mode = "none" if force_dequant in ("nonlinear", "softmax") else quant_mode
        self.act = QuantAct(16, quant_mode=self.quant_mode)
        self.x0, self.n = -0.6931, 30

    def fix(self):
        pass

    def unfix(self):
        self.invalidate()

    def forward(self, x, scaling_factor):
        if self.quant_mode == "none":
            x = x.dequantize() if isinstance(x, QTensor) else x
            return F.softmax(x, dim=-1, dtype=torch.float32), None
        if isinstance(x, QTensor):
            return self._frozen_forward(x)
        x_int = x / scaling_factor
        x_int = x_int - x_int.max(dim=-1, keepdim=True)[0]
        e, s_e = simq.ibert_int_exp(x_int, scaling_factor)
        e, s_e = self.act(e, s_e)                                   # inner QuantAct(16), :310
        e = e / s_e
        ssum = e.sum(dim=-1, keepdim=True)
        factor = torch.floor(2 ** 32 / ssum)
        out = torch.floor(e * factor / 2 ** (32 - self.output_bit + 1))
        s_out = torch.tensor([2 / 2 ** self.output_bit], device=x.device)
        return out * s_out, s_out


# =============================================================================== fused attention
def _frozen_attention(p, view: QTensor, qact2: QuantAct, device):
    """vit_quant.py:72-85 recognised as one unit: matmul_1(q, k^T) * scale -> qact_attn1 -> int_softmax
    -> matmul_2(attn, v) -> transpose/reshape -> qact2  ==>  ivit_attention_i8 on the qkv buffer."""
    sm, v = p.P, p.v
    qattn = sm.qattn
    if qattn.kind == "qattn_bias":
        return _frozen_window_attention(p, view, qact2, device)
    mm1 = qattn.mm1
    q, kT = mm1.a, mm1.b
    base = q.codes
    if base is None or kT.codes is not base or v.codes is not base or base.dtype != torch.int8:
        raise NotImplementedError("QuantMatMul operands must be slices of one int8 qkv carrier (vit_quant.py:65-71)")
    B, H, N, hd = (int(s) for s in q.shape)
    C = H * hd
    want = torch.empty((B, N, 3, H, hd), device="meta", dtype=torch.int8).permute(2, 0, 3, 1, 4)

    def same(mt, ref):
        return tuple(mt.shape) == tuple(ref.shape) and tuple(mt.stride()) == tuple(ref.stride()) and \
            mt.storage_offset() == ref.storage_offset()
    if not (same(q.meta, want[0]) and same(kT.meta, want[1].transpose(-2, -1)) and same(v.meta, want[2])):
        raise NotImplementedError("QuantMatMul operand layout differs from qkv.reshape(B,N,3,H,hd).permute(2,0,3,1,4)")
    if not view.is_canonical():
        raise NotImplementedError("attention output must be consumed as x.transpose(1,2).reshape(B,N,C)")
    cache = qact2._fz_get()
    key = ("attn", id(sm.mod))
    if key not in cache:
        a = _cabi.AttentionParams()
        a.kind, a.output_bit = sm.mod._kind, int(sm.mod.output_bit)
        a.s_pv = float(np.asarray(p.s_h, dtype=f32).reshape(-1)[0])
        qa = qattn.qact
        s_attn, s_attn_h, (lo, hi) = qa._out_scale(device)
        a.mult_attn = float(hm.dyadic_mult(qattn.s_in_h, s_attn_h).reshape(-1)[0])
        a.qmin_attn, a.qmax_attn = lo, hi
        a.s_attn = float(s_attn_h[0])
        if sm.mod._kind == 1:
            inner = sm.mod.act
            s16 = hm.sym_scale(16, fz.host_scale(inner.x_min), fz.host_scale(inner.x_max)).reshape(1)
            exp_scale = ((f32(0.35815147) * (s_attn_h * s_attn_h)) / f32(2 ** 30)).astype(f32)
            a.mult_exp16 = float(hm.dyadic_mult(exp_scale, s16).reshape(-1)[0])
            a.s_exp16 = float(s16[0])
        _, s_out_h, (lo2, hi2) = qact2._out_scale(device)
        a.mult_out = float(hm.dyadic_mult(p.s_h, s_out_h).reshape(-1)[0])
        a.qmin_out, a.qmax_out = lo2, hi2
        table = ops.softmax_table_build(a, device, host_table=sm.mod.host_term_table(s_attn_h) if sm.mod._kind == 2 else None)
        cache[key] = (a, table)
    a, _ = cache[key]
    return ops.attention(base.reshape(B, N, 3 * C), B, N, H, hd, a)


def _frozen_window_attention(p, view: QTensor, qact3: QuantAct, device):
    """swin_quant.py:131-160 recognised as one unit: matmul_1(q, k^T) * scale -> qact_attn1 -> qact2 with the
    gathered relative-position-bias codes as identity -> (+ shifted-window mask) -> IntSoftmax -> matmul_2(attn, v)
    -> transpose/reshape -> qact3  ==>  ivit_window_attention_i8 in window order (the caller has already rolled and
    partitioned the tokens: swin_quant.py:258-272)."""
    from ..swin_plan import mask_always_clamps, shift_window_mask
    sm, v = p.P, p.v
    chain = sm.qattn                       # qact2 + identity (+ mask)
    qattn = chain.qattn                    # qact_attn1
    mm1 = qattn.mm1
    q, kT = mm1.a, mm1.b
    base = q.codes
    if base is None or kT.codes is not base or v.codes is not base or base.dtype != torch.int8:
        raise NotImplementedError("QuantMatMul operands must be slices of one int8 qkv carrier (swin_quant.py:127-129)")
    Bw, H, N, hd = (int(s) for s in q.shape)
    C = H * hd
    want = torch.empty((Bw, N, 3, H, hd), device="meta", dtype=torch.int8).permute(2, 0, 3, 1, 4)

    def same(mt, ref):
        return tuple(mt.shape) == tuple(ref.shape) and tuple(mt.stride()) == tuple(ref.stride()) and \
            mt.storage_offset() == ref.storage_offset()
    if not (same(q.meta, want[0]) and same(kT.meta, want[1].transpose(-2, -1)) and same(v.meta, want[2])):
        raise NotImplementedError("QuantMatMul operand layout differs from qkv.reshape(B_,N,3,H,hd).permute(2,0,3,1,4)")
    if not view.is_canonical():
        raise NotImplementedError("window attention output must be consumed as x.transpose(1,2).reshape(B_,N,C)")
    if sm.mod._kind != 0:
        raise NotImplementedError("frozen window attention: the ivit softmax only (swin_quant.py:11 imports no other)")
    ws = int(round(N ** 0.5))
    if ws * ws != N:
        raise NotImplementedError("window attention: square windows")
    # geometry: without a mask every window is independent (one window per 'image'); with one, the mask must be the
    # shifted-window mask of (res, ws, ws // 2) -- checked once per mask tensor against the reference construction
    B, res, shift = Bw, ws, 0
    mask = chain.mask
    cache = qact3._fz_get()
    if mask is not None:
        nW = mask.numel() // (N * N) if tuple(mask.shape[-2:]) == (N, N) else 0    # [1, nW, 1, N, N] at the add (:150)
        n = int(round(nW ** 0.5))
        if n * n != nW or nW == 0 or Bw % nW:
            raise NotImplementedError("window attention mask: [nW, N, N] with a square number of windows")
        res, shift, B = n * ws, ws // 2, Bw // nW
        mkey = ("mask_ok", mask.data_ptr(), tuple(mask.shape), mask._version)
        if mkey not in cache:
            want_mask = torch.from_numpy(shift_window_mask(res, ws, shift)).to(mask.device)
            cache[mkey] = bool(torch.equal(mask.detach().float().reshape(nW, N, N), want_mask))
        if not cache[mkey]:
            raise NotImplementedError("window attention mask differs from SwinTransformerBlock's 0 / -100 shifted-window "
                                      "mask for shift = window_size // 2 (swin_quant.py:223-247)")
    ident = chain.identity
    bias = ident.materialize()
    if bias.dtype != torch.int8 or tuple(bias.shape)[-3:] != (H, N, N) or bias.numel() != H * N * N:
        raise NotImplementedError("qact2 identity must be the int8 bias codes [1, nH, N, N] (swin_quant.py:142-147)")
    key = ("wattn", id(sm.mod), id(chain.qact), id(qattn.qact))
    s_tab_h = fz.eff_hscale(ident)
    if key in cache and not np.array_equal(cache[key][2], s_tab_h):
        cache.pop(key)
    if key not in cache:
        a = _cabi.WindowAttentionParams()
        _, s_a1_h, (lo, hi) = qattn.qact._out_scale(device)
        a.mult_attn = float(hm.dyadic_mult(qattn.s_in_h, s_a1_h).reshape(-1)[0])
        a.qmin_attn, a.qmax_attn = lo, hi
        _, s2_h, (lo2, hi2) = chain.qact._out_scale(device)
        a.mult2 = float(hm.dyadic_mult(s_a1_h, s2_h).reshape(-1)[0])
        a.mult_bias = float(hm.dyadic_mult(s_tab_h, s2_h).reshape(-1)[0])
        a.qmin2, a.qmax2 = lo2, hi2
        a.s2 = float(s2_h[0])
        a.output_bit = int(sm.mod.output_bit)
        a.mask_clamps = int(mask_always_clamps(s2_h[0], lo2, hi2))
        keep = []
        if -128 <= lo2 and hi2 <= 127:
            tp = _cabi.AttentionParams()
            tp.kind, tp.output_bit, tp.s_attn = 0, int(sm.mod.output_bit), float(s2_h[0])
            keep.append(ops.softmax_table_build(tp, device, collapse=False))
            a.exp_table = tp.exp_table
            a.sat_k, a.sat_e = tp.sat_k, tp.sat_e
        _, s3_h, (lo3, hi3) = qact3._out_scale(device)
        a.mult_out = float(hm.dyadic_mult(p.s_h, s3_h).reshape(-1)[0])
        a.qmin_out, a.qmax_out = lo3, hi3
        a.windowed = 1
        cache[key] = (a, keep, s_tab_h.copy())
    a = cache[key][0]
    if qact3.activation_bit > 8 or not 2 <= int(sm.mod.output_bit) <= 8:
        raise NotImplementedError("frozen window attention: 8-bit qact3 and softmax output (swin_quant.py:108-109)")
    bias = bias.contiguous()
    a.bias_codes = bias.data_ptr()                       # re-gathered by the caller every forward: pass this call's codes
    a.bias_terms = None
    out = ops.window_attention(base.reshape(B, res * res, 3 * C), B, res, ws, shift, H, hd, a)
    return out.reshape(Bw, N, C)

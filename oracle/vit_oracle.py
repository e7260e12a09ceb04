"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's frozen ViT/DeiT forward.

Composes the per-module functions of ``oracle/ivit_oracle.py`` in the order of
``vit_quant.py:285-312`` (VisionTransformer.forward_features/forward),
``vit_quant.py:142-155`` (Block.forward), ``vit_quant.py:61-90``
(Attention.forward), ``layers_quant.py:145-154`` (Mlp.forward) and
``layers_quant.py:191-203`` (PatchEmbed.forward).  Activation ranges
(``x_min``/``x_max`` of every QuantAct) come from a calibrated checkpoint, the
frozen-model lifecycle of SURVEY.md section 3.3.

``forward`` optionally records the integer codes at every module boundary under
the reference's module names, so that the same names index the reference's
forward-hook dumps, the golden fixtures, and the CUDA path's taps.
"""
from __future__ import annotations

import numpy as np

from . import ivit_oracle as O

f32 = np.float32


def _np(t):
    return t.detach().cpu().numpy() if hasattr(t, "detach") else np.asarray(t)


def _layer_kwargs(name: str) -> dict:
    """'ppoly_scale-bits_20_seg_16' -> {'scale_bits': 20, 'seg': 16} (layer_selection.py:137-179)."""
    parts = name.lower().split("_")
    out = {}
    if len(parts) >= 3:
        for k, v in zip(parts[1:-1:2], parts[2::2]):
            out[k.replace("-", "_")] = int(v) if v.isdigit() else v
    return out


class VitOracle:
    def __init__(self, cfg, state: dict):
        """cfg: ivit_b200.synthetic.VitConfig; state: name -> array (weights and x_min/x_max buffers)."""
        self.cfg = cfg
        self.sd = {k: (_np(v) if ".fixed_" in k else _np(v).astype(np.float32)) for k, v in state.items()
                   if not k.endswith(("weight_integer", "bias_integer"))}       # ppoly coefficients stay integers
        self.taps: dict = {}
        self.record = False

    # -- helpers -------------------------------------------------------------
    def _tap(self, name, y, scale):
        """Record the float32 carrier (codes*scale) and its scale, exactly what a
        forward hook on the reference module of the same name sees."""
        if self.record:
            self.taps[name] = (np.asarray(y), np.asarray(scale))

    def _qact(self, name, x, bits, pre_scale=None, identity=None, identity_scale=None):
        y, s, codes = O.quant_act(x, bits, self.sd[name + ".x_min"], self.sd[name + ".x_max"],
                                  pre_scale, identity, identity_scale)
        self._tap(name, y, s)
        return y, s

    def _linear(self, name, x, s):
        y, sc, acc = O.quant_linear(x, s, self.sd[name + ".weight"], self.sd.get(name + ".bias"))
        self._tap(name, y, sc)
        return y, sc

    def _ln(self, name, x, s):
        w, b = self.sd[name + ".weight"], self.sd[name + ".bias"]
        if self.cfg.layernorm_type.startswith("ivit"):
            y, sc, yi = O.ivit_layernorm(x, s, w, b)
        else:
            use_int = "use-int-sqrt_true" in self.cfg.layernorm_type
            y, sc, yi = O.ibert_layernorm(x, s, w, b, shift=float(self.sd.get(name + ".shift", np.zeros(1))[0]),
                                          use_int_sqrt=use_int)
        self._tap(name, y, sc)
        return y, sc

    def _gelu(self, name, x, s):
        if self.cfg.gelu_type.startswith("ivit"):
            y, sc, yi = O.ivit_gelu(x, s)
        elif self.cfg.gelu_type.startswith("ppoly"):
            kw = _layer_kwargs(self.cfg.gelu_type)
            y, sc = O.ppoly_gelu(x, s, self.sd[name + ".fixed_bounds"], self.sd[name + ".fixed_coeffs"],
                                 self.sd[name + ".fixed_scaling_factor_out"], int(kw.get("scale_bits", 22)))
        else:
            y, sc, yi = O.ibert_gelu(x, s)
        self._tap(name, y, sc)
        return y, sc

    def _softmax(self, name, x, s, bits):
        if self.cfg.softmax_type.startswith("ivit"):
            y, sc, yi = O.ivit_softmax(x, s, bits)
        elif self.cfg.softmax_type.startswith("ppoly"):
            kw = _layer_kwargs(self.cfg.softmax_type)
            y, sc, yi = O.ppoly_softmax(x, s, self.sd[name + ".fixed_bounds"], self.sd[name + ".fixed_coeffs"], bits,
                                        int(kw.get("exp_bits", 16)))
        else:
            y, sc, yi = O.ibert_softmax(x, s, bits, self.sd[name + ".act.x_min"], self.sd[name + ".act.x_max"])
        self._tap(name, y, sc)
        return y, sc

    # -- modules ---------------------------------------------------------------
    def _attention(self, p, x, s, bw_out, bw_softmax):
        cfg = self.cfg
        B, N, C = x.shape
        H = cfg.num_heads
        x, s = self._linear(p + "qkv", x, s)
        x, s1 = self._qact(p + "qact1", x, 8, s)
        qkv = x.reshape(B, N, 3, H, C // H).transpose(2, 0, 3, 1, 4)
        q, k, v = qkv[0], qkv[1], qkv[2]
        attn, s, acc = O.quant_matmul(q, s1, np.swapaxes(k, -2, -1), s1)
        self._tap(p + "matmul_1", attn, s)
        sc = f32((C // H) ** -0.5)
        attn = (attn * sc).astype(f32)                 # vit_quant.py:74
        s = (s * sc).astype(f32)                       # vit_quant.py:75
        attn, s = self._qact(p + "qact_attn1", attn, 8, s)
        attn, s = self._softmax(p + "int_softmax", attn, s, bw_softmax)
        x, s, acc = O.quant_matmul(attn, s, v, s1)
        self._tap(p + "matmul_2", x, s)
        x = x.transpose(0, 2, 1, 3).reshape(B, N, C)
        x, s = self._qact(p + "qact2", x, 8, s)
        x, s = self._linear(p + "proj", x, s)
        x, s = self._qact(p + "qact3", x, bw_out, s)
        return x, s

    def _mlp(self, p, x, s, bw_out):
        x, s = self._linear(p + "fc1", x, s)
        x, s = self._qact(p + "qact_gelu", x, 8, s)
        x, s = self._gelu(p + "act", x, s)
        x, s = self._qact(p + "qact1", x, 8, s)
        x, s = self._linear(p + "fc2", x, s)
        x, s = self._qact(p + "qact2", x, bw_out, s)
        return x, s

    def _block(self, i, x1, s1):
        (_, _, _, attention_out_bw, softmax_bw, mlp_out_bw, norm2_in_bw, att_block_out_bw) = self.cfg.bitwidths
        p = f"blocks.{i}."
        x, s = self._ln(p + "norm1", x1, s1)
        x, s = self._qact(p + "qact1", x, 8, s)
        x, s = self._attention(p + "attn.", x, s, attention_out_bw, softmax_bw)
        x2, s2 = self._qact(p + "qact2", x, norm2_in_bw, s, x1, s1)
        x, s = self._ln(p + "norm2", x2, s2)
        x, s = self._qact(p + "qact3", x, 8, s)
        x, s = self._mlp(p + "mlp.", x, s, mlp_out_bw)
        x, s = self._qact(p + "qact4", x, att_block_out_bw, s, x2, s2)
        return x, s

    def forward(self, images, record: bool = False, blocks: int | None = None):
        """images: float32 [B,3,H,W].  Returns float32 logits [B,num_classes]."""
        cfg = self.cfg
        self.record = record
        self.taps = {}
        patch_embed_bw, pos_bw, block_in_bw = cfg.bitwidths[:3]
        x = np.asarray(images, dtype=f32)
        B = x.shape[0]
        x, s = self._qact("qact_input", x, 8)
        y, s, acc = O.quant_conv_patch(x, s, self.sd["patch_embed.proj.weight"],
                                       self.sd["patch_embed.proj.bias"], cfg.patch_size)
        self._tap("patch_embed.proj", y, s)
        C = y.shape[1]
        x = y.reshape(B, C, -1).transpose(0, 2, 1)               # flatten(2).transpose(1,2)
        x, s = self._qact("patch_embed.qact", x, patch_embed_bw, s.reshape(-1))
        cls = np.broadcast_to(self.sd["cls_token"], (B, 1, C))
        x = np.concatenate([cls, x], axis=1).astype(f32)         # vit_quant.py:293
        xp, sp = self._qact("qact_pos", self.sd["pos_embed"], pos_bw)
        x, s = self._qact("qact1", x, block_in_bw, s, xp, sp)
        for i in range(cfg.depth if blocks is None else blocks):
            x, s = self._block(i, x, s)
        x, s = self._ln("norm", x, s)
        x = x[:, 0]
        x, s = self._qact("qact2", x, 8, s)
        y, s = self._linear("head", x, s)
        return y


def codes_of(tap):
    """Integer codes of a recorded (carrier, scale) pair."""
    y, s = tap
    return np.rint(np.asarray(y, dtype=f32) / np.asarray(s, dtype=f32))

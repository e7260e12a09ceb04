"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's frozen Swin forward.

Composes ``oracle/ivit_oracle.py`` in the order of ``swin_quant.py``: ``SwinTransformer.forward``
(:536-563), ``BasicLayer.forward`` (:404-412), ``SwinTransformerBlock.forward`` (:251-301),
``WindowAttention.forward`` (:121-169), ``PatchMerging.forward`` (:327-349), ``window_partition`` /
``window_reverse`` (:18-50), the shifted-window mask (:223-247) and ``PatchEmbed.forward`` with its
norm (layers_quant.py:191-203).  Layer types are the I-ViT ones (the ``Int*`` names swin_quant.py
imports), bit-widths as hard-coded there (16-bit residual stream).

Pinned against the reference itself (shimmed import, oracle/ref_shim.load_reference_swin) by
tests/golden/swin_*.npz and tests/test_swin_oracle.py.
"""
from __future__ import annotations

import numpy as np

from . import ivit_oracle as O

f32 = np.float32


def _np(t):
    return t.detach().cpu().numpy() if hasattr(t, "detach") else np.asarray(t)


def window_partition(x, ws):
    """swin_quant.py:18-32: [B,H,W,C] -> [B*nW, ws, ws, C]."""
    B, H, W, C = x.shape
    x = x.reshape(B, H // ws, ws, W // ws, ws, C)
    return x.transpose(0, 1, 3, 2, 4, 5).reshape(-1, ws, ws, C)


def window_reverse(windows, ws, H, W):
    """swin_quant.py:35-50."""
    B = windows.shape[0] // ((H // ws) * (W // ws))
    x = windows.reshape(B, H // ws, W // ws, ws, ws, -1)
    return x.transpose(0, 1, 3, 2, 4, 5).reshape(B, H, W, -1)


def relative_position_index(ws):
    """swin_quant.py:80-95."""
    coords = np.stack(np.meshgrid(np.arange(ws), np.arange(ws), indexing="ij"))
    cf = coords.reshape(2, -1)
    rel = (cf[:, :, None] - cf[:, None, :]).transpose(1, 2, 0).copy()
    rel[:, :, 0] += ws - 1
    rel[:, :, 1] += ws - 1
    rel[:, :, 0] *= 2 * ws - 1
    return rel.sum(-1)


def shift_mask(res, ws, shift):
    """swin_quant.py:223-247: float 0 / -100 mask [nW, ws*ws, ws*ws] of the shifted-window blocks."""
    if shift == 0:
        return None
    img = np.zeros((1, res, res, 1), dtype=f32)
    cnt = 0
    for h in (slice(0, -ws), slice(-ws, -shift), slice(-shift, None)):
        for w in (slice(0, -ws), slice(-ws, -shift), slice(-shift, None)):
            img[:, h, w, :] = cnt
            cnt += 1
    mw = window_partition(img, ws).reshape(-1, ws * ws)
    m = mw[:, None, :] - mw[:, :, None]
    return np.where(m != 0, f32(-100.0), f32(0.0)).astype(f32)


def avgpool_tokens(x):
    """``nn.AdaptiveAvgPool1d(1)(x.transpose(1, 2))`` (swin_quant.py:554) on CPU float32: torch lowers it
    to ``sum(dim) / L`` over a strided dimension, i.e. ATen's ``multi_row_sum``: every output sums
    its L tokens sequentially, cascading into a higher-level accumulator every 16 tokens
    (SumKernel.cpp).  x [B, L, C] -> [B, C].  Verified against torch in tests/test_swin_oracle.py."""
    x = np.asarray(x, dtype=f32)
    B, L, C = x.shape
    lp = max(4, max(L - 1, 0).bit_length() // 4)
    step, mask = 1 << lp, (1 << lp) - 1
    acc = np.zeros((4, B, C), dtype=f32)
    i = 0
    while i + step <= L:
        for _ in range(step):
            acc[0] = acc[0] + x[:, i, :]
            i += 1
        for j in range(1, 4):
            acc[j] = acc[j] + acc[j - 1]
            acc[j - 1] = 0
            if (i & (mask << (j * lp))) != 0:
                break
    while i < L:
        acc[0] = acc[0] + x[:, i, :]
        i += 1
    for j in range(1, 4):
        acc[0] = acc[0] + acc[j]
    return (acc[0] / f32(L)).astype(f32)


class SwinOracle:
    def __init__(self, cfg, state: dict):
        """cfg: ivit_b200.synthetic.SwinConfig; state: name -> array (weights and x_min / x_max buffers)."""
        self.cfg = cfg
        self.sd = {k: _np(v).astype(np.float32) for k, v in state.items()
                   if not k.endswith(("weight_integer", "bias_integer"))}
        self.taps: dict = {}
        self.record = False

    def _tap(self, name, y, scale):
        if self.record:
            self.taps[name] = (np.asarray(y), np.asarray(scale))

    def _qact(self, name, x, bits, pre_scale=None, identity=None, identity_scale=None):
        y, s, _ = O.quant_act(x, bits, self.sd[name + ".x_min"], self.sd[name + ".x_max"], pre_scale, identity,
                              identity_scale)
        self._tap(name, y, s)
        return y, s

    def _linear(self, name, x, s):
        y, sc, _ = O.quant_linear(x, s, self.sd[name + ".weight"], self.sd.get(name + ".bias"))
        self._tap(name, y, sc)
        return y, sc

    def _ln(self, name, x, s, strided_input=False):
        y, sc, _ = O.ivit_layernorm(x, s, self.sd[name + ".weight"], self.sd[name + ".bias"], strided_input)
        self._tap(name, y, sc)
        return y, sc

    # ------------------------------------------------------------------ WindowAttention.forward
    def _attention(self, p, x, s, nH, ws, mask):
        B_, N, C = x.shape
        x, s = self._linear(p + "qkv", x, s)
        x, s1 = self._qact(p + "qact1", x, 8, s)
        qkv = x.reshape(B_, N, 3, nH, C // nH).transpose(2, 0, 3, 1, 4)
        q, k, v = qkv[0], qkv[1], qkv[2]
        attn, s, _ = O.quant_matmul(q, s1, np.swapaxes(k, -2, -1), s1)
        self._tap(p + "matmul_1", attn, s)
        sc = f32((C // nH) ** -0.5)
        attn = (attn * sc).astype(f32)                                   # swin_quant.py:135
        s = (s * sc).astype(f32)                                         # :136
        attn, s = self._qact(p + "qact_attn1", attn, 8, s)
        tab, s_tab = self._qact(p + "qact_table", self.sd[p + "relative_position_bias_table"], 8)   # :140-141
        idx = relative_position_index(ws).reshape(-1)
        bias = tab[idx].reshape(N, N, -1).transpose(2, 0, 1)             # :142-145  [nH, N, N]
        attn, s = self._qact(p + "qact2", attn, 8, s, np.broadcast_to(bias[None], attn.shape), s_tab)   # :147
        if mask is not None:                                             # :149-153 float add on the carrier
            nW = mask.shape[0]
            attn = (attn.reshape(B_ // nW, nW, nH, N, N) + mask[None, :, None]).astype(f32).reshape(-1, nH, N, N)
        attn_p, s_p, _ = O.ivit_softmax(attn, s, 8)
        self._tap(p + "log_int_softmax", attn_p, s_p)
        x, s, _ = O.quant_matmul(attn_p, s_p, v, s1)
        self._tap(p + "matmul_2", x, s)
        x = x.transpose(0, 2, 1, 3).reshape(B_, N, C)
        x, s = self._qact(p + "qact3", x, 8, s)
        x, s = self._linear(p + "proj", x, s)
        x, s = self._qact(p + "qact4", x, 16, s)
        return x, s

    def _mlp(self, p, x, s):
        x, s = self._linear(p + "fc1", x, s)
        x, s = self._qact(p + "qact_gelu", x, 8, s)
        y, sc, _ = O.ivit_gelu(x, s)
        self._tap(p + "act", y, sc)
        x, s = self._qact(p + "qact1", y, 8, sc)
        x, s = self._linear(p + "fc2", x, s)
        x, s = self._qact(p + "qact2", x, 8, s)
        return x, s

    # ------------------------------------------------------------------ SwinTransformerBlock.forward
    def _block(self, i, j, x1, s1):
        cfg = self.cfg
        p = f"layers.{i}.blocks.{j}."
        res, C, nH = cfg.stage_res(i), cfg.stage_dim(i), cfg.num_heads[i]
        ws, shift = cfg.block_window_shift(i, j)
        B = x1.shape[0]
        # Stage 0 only: the residual stream still has the memory layout of the patch-embedding conv output
        # (flatten(2).transpose(1, 2): channel stride = L), which every elementwise op up to the first
        # PatchMerging `cat` preserves -- the identity branch comes first in `output1 + output`
        # (quant_utils.py:245) -- so torch sums the LayerNorm mean in its outer-reduction order there
        # [traced on the reference: x.stride() at every norm1 / norm2 of layers.0].
        strided = i == 0
        x, s = self._ln(p + "norm1", x1, s1, strided)
        x, s = self._qact(p + "qact1", x, 8, s)
        x = x.reshape(B, res, res, C)
        if shift > 0:
            x = np.roll(x, (-shift, -shift), axis=(1, 2))
        xw = window_partition(x, ws).reshape(-1, ws * ws, C)
        aw, s = self._attention(p + "attn.", xw, s, nH, ws, shift_mask(res, ws, shift))
        x = window_reverse(aw.reshape(-1, ws, ws, C), ws, res, res)
        if shift > 0:
            x = np.roll(x, (shift, shift), axis=(1, 2))
        x = x.reshape(B, res * res, C)
        x2, s2 = self._qact(p + "qact2", x, 16, s, x1, s1)
        x, s = self._ln(p + "norm2", x2, s2, strided)
        x, s = self._qact(p + "qact3", x, 8, s)
        x, s = self._mlp(p + "mlp.", x, s)
        return self._qact(p + "qact4", x, 16, s, x2, s2)

    def _merge(self, i, x, s):
        """PatchMerging.forward (swin_quant.py:327-349)."""
        p = f"layers.{i}.downsample."
        res, C = self.cfg.stage_res(i), self.cfg.stage_dim(i)
        B = x.shape[0]
        x = x.reshape(B, res, res, C)
        x = np.concatenate([x[:, 0::2, 0::2], x[:, 1::2, 0::2], x[:, 0::2, 1::2], x[:, 1::2, 1::2]], axis=-1)
        x = x.reshape(B, -1, 4 * C)
        x, s = self._ln(p + "norm", x, s)
        x, s = self._qact(p + "qact1", x, 8, s)
        x, s = self._linear(p + "reduction", x, s)
        return self._qact(p + "qact2", x, 8, s)

    def forward(self, images, record: bool = False):
        cfg = self.cfg
        self.record, self.taps = record, {}
        x = np.asarray(images, dtype=f32)
        B = x.shape[0]
        x, s = self._qact("qact_input", x, 8)
        y, s, _ = O.quant_conv_patch(x, s, self.sd["patch_embed.proj.weight"], self.sd["patch_embed.proj.bias"],
                                     cfg.patch_size)
        self._tap("patch_embed.proj", y, s)
        C = y.shape[1]
        x = y.reshape(B, C, -1).transpose(0, 2, 1)
        x, s = self._qact("patch_embed.qact_before_norm", x, 8, s.reshape(-1))
        # the conv output reaches the norm as flatten(2).transpose(1, 2): channel stride != 1 (layers_quant.py:198)
        x, s = self._ln("patch_embed.norm", x, s, strided_input=True)
        x, s = self._qact("patch_embed.qact", x, 8, s)
        x, s = self._qact("qact1", x, 16, s)
        for i, depth in enumerate(cfg.depths):
            for j in range(depth):
                x, s = self._block(i, j, x, s)
            if i < len(cfg.depths) - 1:
                x, s = self._merge(i, x, s)
        x, s = self._ln("norm", x, s)
        x, s = self._qact("qact2", x, 8, s)
        x = avgpool_tokens(x)                                            # :554
        x, s = self._qact("qact3", x[:, :, None], 8, s)                  # [B, C, 1] with a scalar scale
        x = x[:, :, 0]
        y, s = self._linear("head", x, s)
        return y


def codes_of(tap):
    y, s = tap
    return np.rint(np.asarray(y, dtype=f32) / np.asarray(s, dtype=f32))

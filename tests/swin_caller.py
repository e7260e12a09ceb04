"""A Swin caller written against the `quantization_utils` module API: the same module names (= state_dict
keys), constructor arguments and call order as the reference's swin_quant.py:53-169 (WindowAttention),
:172-301 (SwinTransformerBlock), :304-349 (PatchMerging), :362-412 (BasicLayer), :419-563
(SwinTransformer) and layers_quant.py:116-203 (Mlp, PatchEmbed) -- which cannot travel to the GPU box --
but written here from the oracle's statement of that order (oracle/swin_oracle.py).  Test infrastructure
only; tests/test_module_api_cpu.py checks that it calibrates to the reference's own ranges."""
import torch
import torch.nn as nn

from ivit_b200.quantization_utils import (IntGELU, IntLayerNorm, IntSoftmax, QuantAct, QuantConv2d, QuantLinear,
                                          QuantMatMul)


def window_partition(x, ws):
    B, H, W, C = x.shape
    x = x.view(B, H // ws, ws, W // ws, ws, C)
    return x.permute(0, 1, 3, 2, 4, 5).contiguous().view(-1, ws, ws, C)


def window_reverse(windows, ws, H, W):
    B = int(windows.shape[0] / (H * W / ws / ws))
    x = windows.view(B, H // ws, W // ws, ws, ws, -1)
    return x.permute(0, 1, 3, 2, 4, 5).contiguous().view(B, H, W, -1)


class Mlp(nn.Module):
    def __init__(self, dim, hidden):
        super().__init__()
        self.fc1 = QuantLinear(dim, hidden)
        self.qact_gelu = QuantAct()
        self.act = IntGELU()
        self.qact1 = QuantAct()
        self.fc2 = QuantLinear(hidden, dim)
        self.qact2 = QuantAct(8)

    def forward(self, x, s):
        x, s = self.fc1(x, s)
        x, s = self.qact_gelu(x, s)
        x, s = self.act(x, s)
        x, s = self.qact1(x, s)
        x, s = self.fc2(x, s)
        return self.qact2(x, s)


class WindowAttention(nn.Module):
    def __init__(self, dim, ws, num_heads):
        super().__init__()
        self.dim, self.window_size, self.num_heads = dim, (ws, ws), num_heads
        self.scale = (dim // num_heads) ** -0.5
        self.relative_position_bias_table = nn.Parameter(torch.zeros((2 * ws - 1) * (2 * ws - 1), num_heads))
        coords = torch.stack(torch.meshgrid([torch.arange(ws), torch.arange(ws)], indexing="ij")).flatten(1)
        rel = (coords[:, :, None] - coords[:, None, :]).permute(1, 2, 0).contiguous()
        rel[:, :, 0] += ws - 1
        rel[:, :, 1] += ws - 1
        rel[:, :, 0] *= 2 * ws - 1
        self.register_buffer("relative_position_index", rel.sum(-1))
        self.qkv = QuantLinear(dim, dim * 3, bias=True)
        self.qact1, self.qact_attn1, self.qact_table, self.qact2 = QuantAct(), QuantAct(), QuantAct(), QuantAct()
        self.log_int_softmax = IntSoftmax()
        self.qact3 = QuantAct()
        self.qact4 = QuantAct(16)
        self.proj = QuantLinear(dim, dim)
        self.matmul_1, self.matmul_2 = QuantMatMul(), QuantMatMul()

    def forward(self, x, s, mask=None):
        B_, N, C = x.shape
        x, s = self.qkv(x, s)
        x, s1 = self.qact1(x, s)
        qkv = x.reshape(B_, N, 3, self.num_heads, C // self.num_heads).permute(2, 0, 3, 1, 4)
        q, k, v = qkv[0], qkv[1], qkv[2]
        attn, s = self.matmul_1(q, s1, k.transpose(-2, -1), s1)
        attn = attn * self.scale
        s = s * self.scale
        attn, s = self.qact_attn1(attn, s)
        tab, s_tab = self.qact_table(self.relative_position_bias_table)
        bias = tab[self.relative_position_index.view(-1)].view(N, N, -1).permute(2, 0, 1).contiguous()
        attn, s = self.qact2(attn, s, bias.unsqueeze(0), s_tab)
        if mask is not None:
            nW = mask.shape[0]
            attn = attn.view(B_ // nW, nW, self.num_heads, N, N) + mask.unsqueeze(1).unsqueeze(0)
            attn = attn.view(-1, self.num_heads, N, N)
        attn, s = self.log_int_softmax(attn, s)
        x, s = self.matmul_2(attn, s, v, s1)
        x = x.transpose(1, 2).reshape(B_, N, C)
        x, s = self.qact3(x, s)
        x, s = self.proj(x, s)
        return self.qact4(x, s)


class SwinTransformerBlock(nn.Module):
    def __init__(self, dim, res, num_heads, window_size, shift_size, mlp_ratio):
        super().__init__()
        self.input_resolution, self.window_size, self.shift_size = (res, res), window_size, shift_size
        if res <= window_size:
            self.shift_size, self.window_size = 0, res
        self.norm1 = IntLayerNorm(dim, eps=1e-6)
        self.qact1 = QuantAct()
        self.attn = WindowAttention(dim, self.window_size, num_heads)
        self.qact2 = QuantAct(16)
        self.norm2 = IntLayerNorm(dim, eps=1e-6)
        self.qact3 = QuantAct()
        self.mlp = Mlp(dim, int(dim * mlp_ratio))
        self.qact4 = QuantAct(16)
        mask = None
        if self.shift_size > 0:
            ws, sh = self.window_size, self.shift_size
            img = torch.zeros((1, res, res, 1))
            cnt = 0
            for h in (slice(0, -ws), slice(-ws, -sh), slice(-sh, None)):
                for w in (slice(0, -ws), slice(-ws, -sh), slice(-sh, None)):
                    img[:, h, w, :] = cnt
                    cnt += 1
            mw = window_partition(img, ws).view(-1, ws * ws)
            mask = mw.unsqueeze(1) - mw.unsqueeze(2)
            mask = mask.masked_fill(mask != 0, float(-100.0)).masked_fill(mask == 0, float(0.0))
        self.register_buffer("attn_mask", mask)

    def forward(self, x1, s1):
        H, W = self.input_resolution
        B, L, C = x1.shape
        x, s = self.norm1(x1, s1)
        x, s = self.qact1(x, s)
        x = x.view(B, H, W, C)
        if self.shift_size > 0:
            x = torch.roll(x, shifts=(-self.shift_size, -self.shift_size), dims=(1, 2))
        xw = window_partition(x, self.window_size).view(-1, self.window_size * self.window_size, C)
        aw, s = self.attn(xw, s, mask=self.attn_mask)
        x = window_reverse(aw.view(-1, self.window_size, self.window_size, C), self.window_size, H, W)
        if self.shift_size > 0:
            x = torch.roll(x, shifts=(self.shift_size, self.shift_size), dims=(1, 2))
        x = x.view(B, H * W, C)
        x2, s2 = self.qact2(x, s, x1, s1)
        x, s = self.norm2(x2, s2)
        x, s = self.qact3(x, s)
        x, s = self.mlp(x, s)
        return self.qact4(x, s, x2, s2)


class PatchMerging(nn.Module):
    def __init__(self, res, dim):
        super().__init__()
        self.input_resolution = (res, res)
        self.norm = IntLayerNorm(4 * dim, eps=1e-6)
        self.qact1 = QuantAct()
        self.reduction = QuantLinear(4 * dim, 2 * dim, bias=False)
        self.qact2 = QuantAct()

    def forward(self, x, s):
        H, W = self.input_resolution
        B, L, C = x.shape
        x = x.view(B, H, W, C)
        x = torch.cat([x[:, 0::2, 0::2, :], x[:, 1::2, 0::2, :], x[:, 0::2, 1::2, :], x[:, 1::2, 1::2, :]], -1)
        x = x.view(B, -1, 4 * C)
        x, s = self.norm(x, s)
        x, s = self.qact1(x, s)
        x, s = self.reduction(x, s)
        return self.qact2(x, s)


class BasicLayer(nn.Module):
    def __init__(self, dim, res, depth, num_heads, window_size, mlp_ratio, downsample):
        super().__init__()
        self.blocks = nn.ModuleList([SwinTransformerBlock(dim, res, num_heads, window_size,
                                                          0 if i % 2 == 0 else window_size // 2, mlp_ratio)
                                     for i in range(depth)])
        self.downsample = PatchMerging(res, dim) if downsample else None

    def forward(self, x, s):
        for blk in self.blocks:
            x, s = blk(x, s)
        if self.downsample is not None:
            x, s = self.downsample(x, s)
        return x, s


class PatchEmbed(nn.Module):
    def __init__(self, img_size, patch, chans, dim):
        super().__init__()
        self.img_size, self.patch_size = (img_size, img_size), (patch, patch)
        self.grid_size = (img_size // patch, img_size // patch)
        self.proj = QuantConv2d(chans, dim, kernel_size=(patch, patch), stride=(patch, patch))
        self.qact_before_norm = QuantAct()
        self.norm = IntLayerNorm(dim, eps=1e-6)
        self.qact = QuantAct(8)

    def forward(self, x, s):
        x, s = self.proj(x, s)
        x = x.flatten(2).transpose(1, 2)
        x, s = self.qact_before_norm(x, s)
        x, s = self.norm(x, s)
        return self.qact(x, s)


class SwinTransformer(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.cfg, self.embed_dim = cfg, cfg.embed_dim
        n = len(cfg.depths)
        self.qact_input = QuantAct()
        self.patch_embed = PatchEmbed(cfg.img_size, cfg.patch_size, cfg.in_chans, cfg.embed_dim)
        self.qact1 = QuantAct(16)
        self.layers = nn.Sequential(*[BasicLayer(cfg.stage_dim(i), cfg.stage_res(i), cfg.depths[i], cfg.num_heads[i],
                                                 cfg.window_size, cfg.mlp_ratio, i < n - 1) for i in range(n)])
        self.norm = IntLayerNorm(cfg.stage_dim(n - 1), eps=1e-6)
        self.qact2 = QuantAct()
        self.avgpool = nn.AdaptiveAvgPool1d(1)
        self.qact3 = QuantAct()
        self.head = QuantLinear(cfg.stage_dim(n - 1), cfg.num_classes)
        self.act_out = QuantAct()

    def forward(self, x):
        x, s = self.qact_input(x)
        x, s = self.patch_embed(x, s)
        x, s = self.qact1(x, s)
        for layer in self.layers:
            x, s = layer(x, s)
        x, s = self.norm(x, s)
        x, s = self.qact2(x, s)
        x = self.avgpool(x.transpose(1, 2))
        x, s = self.qact3(x, s)
        x = torch.flatten(x, 1)
        x, s = self.head(x, s)
        return x

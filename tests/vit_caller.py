"""A ViT/DeiT caller written against the `quantization_utils` module API: the same module names
(= state_dict keys) and the same call order as the reference's vit_quant.py:61-90,142-155,285-312 and
layers_quant.py:145-154,191-203 -- which cannot travel to the GPU box -- but written here from the
oracle's statement of that order (oracle/vit_oracle.py).  Test infrastructure only."""
import torch
import torch.nn as nn

from ivit_b200.quantization_utils import (QuantAct, QuantConv2d, QuantLinear, QuantMatMul, get_gelu, get_layernorm,
                                          get_softmax)


class Attention(nn.Module):
    def __init__(self, dim, heads, softmax_cls, bw_out, bw_softmax):
        super().__init__()
        self.num_heads, self.scale = heads, (dim // heads) ** -0.5
        self.qkv = QuantLinear(dim, dim * 3)
        self.qact1 = QuantAct()
        self.qact_attn1 = QuantAct()
        self.qact2 = QuantAct()
        self.proj = QuantLinear(dim, dim)
        self.qact3 = QuantAct(bw_out)
        self.int_softmax = softmax_cls(bw_softmax)
        self.attn_drop = nn.Dropout(0.0)
        self.matmul_1, self.matmul_2 = QuantMatMul(), QuantMatMul()

    def forward(self, x, s):
        B, N, C = x.shape
        x, s = self.qkv(x, s)
        x, s1 = self.qact1(x, s)
        qkv = x.reshape(B, N, 3, self.num_heads, C // self.num_heads).permute(2, 0, 3, 1, 4)
        q, k, v = qkv[0], qkv[1], qkv[2]
        a, s = self.matmul_1(q, s1, k.transpose(-2, -1), s1)
        a = a * self.scale
        s = s * self.scale
        a, s = self.qact_attn1(a, s)
        a, s = self.int_softmax(a, s)
        a = self.attn_drop(a)
        x, s = self.matmul_2(a, s, v, s1)
        x = x.transpose(1, 2).reshape(B, N, C)
        x, s = self.qact2(x, s)
        x, s = self.proj(x, s)
        return self.qact3(x, s)


class Mlp(nn.Module):
    def __init__(self, dim, hidden, act_layer, bw_out):
        super().__init__()
        self.fc1 = QuantLinear(dim, hidden)
        self.qact_gelu = QuantAct()
        self.act = act_layer()
        self.qact1 = QuantAct()
        self.fc2 = QuantLinear(hidden, dim)
        self.qact2 = QuantAct(bw_out)

    def forward(self, x, s):
        x, s = self.fc1(x, s)
        x, s = self.qact_gelu(x, s)
        x, s = self.act(x, s)
        x, s = self.qact1(x, s)
        x, s = self.fc2(x, s)
        return self.qact2(x, s)


class Block(nn.Module):
    def __init__(self, dim, heads, ratio, act_layer, norm_layer, softmax_cls, bw):
        super().__init__()
        self.norm1 = norm_layer(dim)
        self.qact1 = QuantAct()
        self.attn = Attention(dim, heads, softmax_cls, bw[3], bw[4])
        self.qact2 = QuantAct(bw[6])
        self.norm2 = norm_layer(dim)
        self.qact3 = QuantAct()
        self.mlp = Mlp(dim, int(dim * ratio), act_layer, bw[5])
        self.qact4 = QuantAct(bw[7])

    def forward(self, x1, s1):
        x, s = self.norm1(x1, s1)
        x, s = self.qact1(x, s)
        x, s = self.attn(x, s)
        x2, s2 = self.qact2(x, s, x1, s1)
        x, s = self.norm2(x2, s2)
        x, s = self.qact3(x, s)
        x, s = self.mlp(x, s)
        return self.qact4(x, s, x2, s2)


class PatchEmbed(nn.Module):
    def __init__(self, patch, chans, dim, bw_out):
        super().__init__()
        self.proj = QuantConv2d(chans, dim, kernel_size=patch, stride=patch)
        self.qact = QuantAct(bw_out)

    def forward(self, x, s):
        x, s = self.proj(x, s)
        x = x.flatten(2).transpose(1, 2)
        return self.qact(x, s)


class VisionTransformer(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        bw = cfg.bitwidths
        C = cfg.embed_dim
        self.qact_input = QuantAct()
        self.patch_embed = PatchEmbed(cfg.patch_size, cfg.in_chans, C, bw[0])
        self.cls_token = nn.Parameter(torch.zeros(1, 1, C))
        self.pos_embed = nn.Parameter(torch.zeros(1, cfg.tokens, C))
        self.qact_pos = QuantAct(bw[1])
        self.qact1 = QuantAct(bw[2])
        self.blocks = nn.ModuleList([Block(C, cfg.num_heads, cfg.mlp_ratio, get_gelu(cfg.gelu_type),
                                           get_layernorm(cfg.layernorm_type), get_softmax(cfg.softmax_type), bw)
                                     for _ in range(cfg.depth)])
        self.norm = get_layernorm(cfg.layernorm_type)(C)
        self.qact2 = QuantAct()
        self.head = QuantLinear(C, cfg.num_classes)

    def forward(self, x):
        B = x.shape[0]
        x, s = self.qact_input(x)
        x, s = self.patch_embed(x, s)
        x = torch.cat((self.cls_token.expand(B, -1, -1), x), dim=1)
        xp, sp = self.qact_pos(self.pos_embed)
        x, s = self.qact1(x, s, xp, sp)
        for blk in self.blocks:
            x, s = blk(x, s)
        x, s = self.norm(x, s)
        x = x[:, 0]
        x, s = self.qact2(x, s)
        x, s = self.head(x, s)
        return x


def freeze(model):
    """model_utils.py:5-16."""
    model.eval()
    for m in model.modules():
        if hasattr(m, "fix") and callable(m.fix):
            m.fix()

"""Self-consistency stress on the GPU: the restructured kernels of round 2 against the kernels they replaced, over many
random scales / shapes / value spreads (both families are separately tested against the oracle; this hunts for rare-path
differences: near-tie replays, int32 square wraps, zero-variance rows, Newton fixed points, masked windows).
  python tools/stress_variants.py [iterations]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from ivit_b200 import _cabi, ops, hostmath as hm
from ivit_b200.swin_plan import mask_always_clamps, relative_position_index

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 100
rng = np.random.default_rng(12345)
f32 = np.float32
dev = lambda a, d: torch.from_numpy(np.ascontiguousarray(a)).to(d).cuda()
bad = 0

def ln_case(it):
    global bad
    kind = int(rng.integers(0, 2))
    bits = int(rng.choice([8, 16]))
    strided = kind == 0 and rng.random() < 0.3
    C = int(rng.choice([96, 128])) if strided else int(rng.choice([192, 384, 768] if kind == 1 or bits == 8 else [192, 256, 384, 512, 768]))
    tokens = int(rng.choice([32, 64, 96])) if strided else 0
    rows = tokens * int(rng.integers(1, 9)) if strided else int(rng.integers(1, 700))
    lim = 2 ** (bits - 1)
    spread = lim / float(rng.choice([2.5, 4, 8, 40, 400]))
    q = np.clip(np.rint(rng.standard_normal((rows, C)) * spread + rng.integers(-lim // 4, lim // 4)), -lim, lim - 1)
    if rows > 2:
        q[0, :] = q[0, 0]                                  # zero variance
        q[1, ::2], q[1, 1::2] = -lim, lim - 1              # int32 square wrap at 16 bit
    s_in = f32(np.exp(rng.uniform(np.log(2e-5), np.log(0.2))))
    w = (rng.random(C) + 0.5).astype(f32) * np.where(rng.random(C) < 0.05, -1, 1).astype(f32)
    b = (rng.standard_normal(C) * 0.1).astype(f32)
    p = hm.layernorm_params(w, b, C)
    mult = hm.dyadic_mult(p["sf"], np.array([f32(np.exp(rng.uniform(np.log(0.005), np.log(0.08))))], dtype=f32))
    x = dev(q, torch.int16 if bits == 16 else torch.int8)
    bi, sf, mu = dev(p["bias_int"], torch.float32), dev(p["sf"], torch.float32), dev(mult, torch.float64)
    shift = float(rng.integers(0, 7)) if kind == 1 else 0.0
    outs = []
    for variant in (0, 1):
        ops.set_layernorm_variant(variant)
        if strided:
            outs.append(ops.layernorm_requant_strided(0, x, tokens, float(s_in), bi, sf, mu, -128, 127).clone())
        else:
            outs.append(ops.layernorm_requant(kind, x, float(s_in), bi, sf, mu, -128, 127, shift=shift).clone())
    ops.set_layernorm_variant(0)
    ok = outs[0][2:] if rows > 2 else outs[0]              # rows 0 (inf / nan in the reference) may differ by design? no: compare all
    if not torch.equal(outs[0], outs[1]):
        n = int((outs[0] != outs[1]).sum())
        rows_bad = torch.unique((outs[0] != outs[1]).nonzero()[:, 0])[:5].tolist()
        print(f"LN MISMATCH it={it} kind={kind} bits={bits} C={C} rows={rows} tokens={tokens} s_in={s_in} shift={shift}: {n} codes, rows {rows_bad}")
        bad += 1

def win_case(it):
    global bad
    res = 7 * int(rng.choice([1, 2, 4]))
    shift = int(rng.choice([0, 3])) if res > 7 else 0
    nH, hd, ws, N = int(rng.choice([1, 2, 3, 6])), 32, 7, 49
    B = int(rng.integers(1, 6))
    C = nH * hd
    qkv = np.clip(np.rint(rng.standard_normal((B, res * res, 3 * C)) * rng.choice([10, 40, 90])), -128, 127).astype(np.int8)
    p = _cabi.WindowAttentionParams()
    p.mult_attn = float(np.exp(rng.uniform(np.log(1e-4), np.log(3e-3))))
    p.qmin_attn, p.qmax_attn, p.qmin2, p.qmax2, p.qmin_out, p.qmax_out = -128, 127, -128, 127, -128, 127
    p.mult2 = float(np.exp(rng.uniform(np.log(0.2), np.log(3.0))))
    p.mult_bias = float(np.exp(rng.uniform(np.log(0.05), np.log(1.5))))
    s2 = f32(np.exp(rng.uniform(np.log(0.02), np.log(0.35))))
    p.s2, p.output_bit = float(s2), 8
    p.mult_out = float(np.exp(rng.uniform(np.log(2e-4), np.log(2e-3))))
    tab = rng.integers(-128, 128, size=((2 * ws - 1) ** 2, nH)).astype(np.int8)
    bias = np.ascontiguousarray(tab[relative_position_index(ws).reshape(-1)].reshape(N, N, nH).transpose(2, 0, 1))
    bias_d = dev(bias, torch.int8)
    p.bias_codes = bias_d.data_ptr()
    terms = None
    if rng.random() < 0.5:
        terms = dev(np.rint(bias.astype(np.float64) * p.mult_bias).astype(np.int32), torch.int32)
        p.bias_terms = terms.data_ptr()
    tp = _cabi.AttentionParams()
    tp.kind, tp.output_bit, tp.s_attn = 0, 8, float(s2)
    keep = ops.softmax_table_build(tp, "cuda", collapse=False)
    p.exp_table, p.mask_clamps = tp.exp_table, int(mask_always_clamps(s2, -128, 127))
    if rng.random() < 0.7:
        p.sat_k, p.sat_e = tp.sat_k, tp.sat_e
    p.windowed = int(rng.integers(0, 2))
    x = torch.from_numpy(qkv).cuda()
    outs = []
    for variant in (0, 2, 1):
        prev = ops.set_window_attention_variant(variant)
        o, pr = ops.window_attention(x, B, res, ws, shift, nH, hd, p, want_probs=True)
        outs.append((o.clone(), pr.clone()))
        ops.set_window_attention_variant(prev)
    for k in (1, 2):
        if not (torch.equal(outs[0][0], outs[k][0]) and torch.equal(outs[0][1], outs[k][1])):
            print(f"WINDOW MISMATCH it={it} vs variant {(0, 2, 1)[k]}: res={res} shift={shift} nH={nH} B={B} s2={s2} windowed={p.windowed} "
                  f"out {int((outs[0][0] != outs[k][0]).sum())} probs {int((outs[0][1] != outs[k][1]).sum())}")
            bad += 1
    del keep, terms, bias_d

for it in range(iters):
    ln_case(it)
    if it % 2 == 0:
        win_case(it)
torch.cuda.synchronize()
print(f"stress: {iters} LayerNorm cases, {(iters + 1) // 2} window cases, mismatching cases: {bad}")
sys.exit(1 if bad else 0)

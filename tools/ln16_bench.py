"""16-bit LayerNorm + QuantAct timing, CUDA events: contiguous order (DeiT / Swin stages 1..3 widths) and the
channel-strided order of Swin stage 0, restructured kernels (variant 0) against the division kernels (variant 1)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from ivit_b200 import ops, hostmath as hm

def run(rows, C, tokens, bits=16, kind=0):
    rng = np.random.default_rng(0)
    lim = 2 ** (bits - 1)
    dt = torch.int16 if bits == 16 else torch.int8
    xs = [torch.from_numpy(np.clip(np.rint(rng.standard_normal((rows, C)) * lim / 4), -lim, lim - 1).astype(np.int16)).to(dt).cuda()
          for _ in range(3)]
    w = (rng.random(C) + 0.5).astype(np.float32); b = (rng.standard_normal(C) * 0.1).astype(np.float32)
    p = hm.layernorm_params(w, b, C)
    mult = hm.dyadic_mult(p["sf"], np.array([0.02], dtype=np.float32))
    bi, sf, mu = (torch.from_numpy(a).cuda() for a in (p["bias_int"], p["sf"], mult))
    out = torch.empty((rows, C), dtype=torch.int8, device="cuda")
    s_in = 1.37e-4 if bits == 16 else 0.0371
    outs = []
    for variant in (0, 1):
        ops.set_layernorm_variant(variant)
        def call(i):
            if tokens:
                ops.layernorm_requant_strided(0, xs[i % 3], tokens, s_in, bi, sf, mu, -128, 127, out=out)
            else:
                ops.layernorm_requant(kind, xs[i % 3], s_in, bi, sf, mu, -128, 127, shift=5.0, out=out)
        for i in range(3):
            call(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(30):
            call(i)
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 30 * 1e3
        call(0); outs.append(out.clone())
        print(f"LN16 kind={kind} bits={bits} rows={rows} C={C} tokens={tokens} variant={variant}: {us:.1f} us  "
              f"{(bits // 8 + 1) * rows * C / us / 1e3:.0f} GB/s", flush=True)
    ops.set_layernorm_variant(0)
    print("   variants equal:", bool(torch.equal(outs[0], outs[1])), flush=True)

run(256 * 3136, 96, 3136)          # Swin-T stage 0
run(256 * 3136, 96, 3136, bits=8)  # patch_embed.norm
run(256 * 784, 192, 0)             # stage 1
run(256 * 196, 384, 0)             # stage 2 / DeiT-S w16
run(256 * 49, 768, 0)              # stage 3
run(64 * 197, 768, 0)              # DeiT-B w16 (ivit kind)
run(64 * 197, 768, 0, kind=1)      # DeiT-B ibert w16 (configs[4])
run(256 * 197, 768, 0, kind=1)

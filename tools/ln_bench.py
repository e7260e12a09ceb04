"""LayerNorm+QuantAct kernel timing (DeiT-S shape), CUDA events."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from ivit_b200 import ops, hostmath as hm
rows, C = 50432, int(sys.argv[1]) if len(sys.argv) > 1 else 384
rng = np.random.default_rng(0)
xs = [torch.from_numpy(np.clip(np.rint(rng.standard_normal((rows, C)) * 35), -128, 127).astype(np.int8)).cuda() for _ in range(4)]
w = (rng.random(C) + 0.5).astype(np.float32); b = (rng.standard_normal(C) * 0.1).astype(np.float32)
p = hm.layernorm_params(w, b, C)
mult = hm.dyadic_mult(p["sf"], np.array([0.02], dtype=np.float32))
bi, sf, mu = (torch.from_numpy(a).cuda() for a in (p["bias_int"], p["sf"], mult))
out = torch.empty_like(xs[0])
for kind, variant in ((0, 0), (0, 1), (1, 0)):
    ops.set_layernorm_variant(variant)
    for i in range(4):
        ops.layernorm_requant(kind, xs[i % 4], 0.0371, bi, sf, mu, -128, 127, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(40):
        ops.layernorm_requant(kind, xs[i % 4], 0.0371, bi, sf, mu, -128, 127, out=out)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 40 * 1e3
    print(f"LN kind={kind} variant={variant} rows={rows} C={C}: {us:.1f} us  {2*rows*C/us/1e3:.0f} GB/s", flush=True)

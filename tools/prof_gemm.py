"""A few launches of ivit_gemm_i8 per shape (fast epilogue): the command ncu wraps.
usage: prof_gemm.py N,K,epi[,z_bound_ok] [N,K,epi ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ivit_b200 import ops
m = 50432
g = torch.Generator(device="cuda").manual_seed(0)
for spec in (sys.argv[1:] or ["1152,384,1"]):
    f = [int(v) for v in spec.split(",")]
    n, k, epi, zb = f[0], f[1], f[2], (f[3] if len(f) > 3 else 1)
    A = torch.randint(-128, 128, (m, k), device="cuda", dtype=torch.int8, generator=g)
    W = torch.randint(-20, 21, (n, k), device="cuda", dtype=torch.int8, generator=g)
    bias = torch.randint(-20000, 20000, (n,), device="cuda", dtype=torch.int32, generator=g)
    mult = (torch.rand(n, device="cuda", dtype=torch.float64, generator=g) * 1e-3 + 1e-3)
    s_in = torch.full((n,), 1e-5, device="cuda", dtype=torch.float32)
    ident = torch.randint(-128, 128, (m, n), device="cuda", dtype=torch.int8, generator=g)
    out = torch.empty((m, n), device="cuda", dtype=torch.int8)
    kw = dict(bias=bias, mult=mult, s_in=s_in, out=out, z_bound_ok=bool(zb))
    if epi == 2:
        kw.update(identity=ident, mult2=0.7, mult_id=0.4)
    for i in range(3):
        ops.gemm(A, W, epi, **kw)
    torch.cuda.synchronize()
print("ok")

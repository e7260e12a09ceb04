"""Per-shape timing of ivit_gemm_i8 (CUDA events, 20 launches each, inputs > L2 rotate)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from ivit_b200 import _cabi, ops
M = 50432
shapes = {"qkv": (M, 1152, 384, 1), "proj": (M, 384, 384, 2), "fc1": (M, 1536, 384, 1), "fc2": (M, 384, 1536, 2),
          "patch": (50176, 384, 768, 2)}
g = torch.Generator(device="cuda").manual_seed(0)
for name, (m, n, k, epi) in shapes.items():
    A = [torch.randint(-128, 128, (m, k), device="cuda", dtype=torch.int8, generator=g) for _ in range(3)]
    W = torch.randint(-20, 21, (n, k), device="cuda", dtype=torch.int8, generator=g)
    bias = torch.randint(-20000, 20000, (n,), device="cuda", dtype=torch.int32, generator=g)
    mult = (torch.rand(n, device="cuda", dtype=torch.float64, generator=g) * 1e-3 + 1e-3)
    s_in = torch.full((n,), 1e-5, device="cuda", dtype=torch.float32)
    ident = torch.randint(-128, 128, (m, n), device="cuda", dtype=torch.int8, generator=g)
    out = torch.empty((m, n), device="cuda", dtype=torch.int8)
    for layout, zb in ((1, False), (1, True), (0, False), (0, True)):
        ops.set_gemm_layout(layout)
        kw = dict(bias=bias, mult=mult, s_in=s_in, out=out, z_bound_ok=zb)
        if epi == 2:
            kw.update(identity=ident, mult2=0.7, mult_id=0.4)
        for i in range(3):
            ops.gemm(A[i % 3], W, epi, **kw)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(20):
            ops.gemm(A[i % 3], W, epi, **kw)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 20 * 1e3
        print(f"{name:6s} layout={"rows" if layout else "auto"} M={m} N={n} K={k} epi={epi} z_bound_ok={zb}: {us:7.1f} us  {2.0*m*n*k/us/1e6:7.1f} TOP/s", flush=True)

"""Per-phase clock probe of attention_i8_kernel (build with IVIT_NVCC_EXTRA=-DIVIT_ATTN_TIMING): average cycles thread 0
of a CTA spends between the phase marks, over one DeiT-S forward at batch 256."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ivit_b200 import _cabi
from ivit_b200.plan import FrozenVitPlan
from ivit_b200.synthetic import load_case, synth_images
cfg, state, _ = load_case("deit_small_ivit")
plan = FrozenVitPlan(cfg, state, "cuda")
x = synth_images(256, cfg, seed=0).cuda()
plan.forward(x); torch.cuda.synchronize()
lib = _cabi.load()
buf = (C.c_ulonglong * 16)()
lib.ivit_debug_attn_timing(buf, 1)
plan.forward(x); torch.cuda.synchronize()
lib.ivit_debug_attn_timing(buf, 1)
ctas = 12 * 256 * cfg.num_heads * 2
names = ["init+alloc", "TMA Q/K + S MMA", "pass 1 + exchange", "pass 2", "sums", "pass 3 + barrier", "PV MMA", "epilogue"]
tot = sum(buf[i] for i in range(8))
for i, n in enumerate(names):
    print(f"{n:22s} {buf[i] / ctas:9.0f} cycles/CTA  {buf[i] / tot:6.1%}")
print("total", tot / ctas, "cycles per CTA")

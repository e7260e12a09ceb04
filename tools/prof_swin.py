"""Two eager forwards of Swin-T (batch 64): the command ncu wraps for the Swin / 16-bit kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ivit_b200.swin_plan import FrozenSwinPlan
from ivit_b200.synthetic import load_swin_case, synth_images
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
cfg, state, _ = load_swin_case("swin_tiny")
plan = FrozenSwinPlan(cfg, state, "cuda")
x = synth_images(B, cfg, seed=0).cuda()
for _ in range(2):
    y = plan.forward(x)
torch.cuda.synchronize()
print("ok", float(y.abs().sum()))

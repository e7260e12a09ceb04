"""Two eager forwards of the bench workload (DeiT-S ivit, batch 256): the command ncu wraps."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ivit_b200.plan import FrozenVitPlan
from ivit_b200.synthetic import load_case, synth_images
case = sys.argv[1] if len(sys.argv) > 1 else "deit_small_ivit"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
cfg, state, _ = load_case(case)
plan = FrozenVitPlan(cfg, state, "cuda")
x = synth_images(B, cfg, seed=0).cuda()
for _ in range(2):
    y = plan.forward(x)
torch.cuda.synchronize()
print("ok", float(y.abs().sum()))

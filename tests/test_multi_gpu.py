"""Two devices in one process (GPU, skipped on a single-GPU box): a plan built and run on cuda:1 while cuda:0 is
the current device, and the batch-sharded forward of SURVEY 8e -- shards on two devices, concatenated logits
bit-equal to one device running the whole batch."""
import numpy as np
import pytest
import torch

from ivit_b200.plan import FrozenVitPlan
from ivit_b200.sharding import shard_range
from ivit_b200.synthetic import load_case, synth_images

pytestmark = pytest.mark.gpu


def _two_devices():
    return torch.cuda.is_available() and torch.cuda.device_count() >= 2


@pytest.mark.skipif(not _two_devices(), reason="needs two CUDA devices")
def test_plan_on_second_device_while_first_is_current():
    cfg, state, _ = load_case("toy_ivit")
    x = synth_images(5, cfg, seed=3)
    torch.cuda.set_device(0)
    ref = FrozenVitPlan(cfg, state, "cuda:0").forward(x.to("cuda:0")).cpu()
    plan1 = FrozenVitPlan(cfg, state, "cuda:1")                 # built with cuda:0 current
    assert torch.cuda.current_device() == 0
    out1 = plan1.forward(x.to("cuda:1")).cpu()
    assert torch.equal(out1, ref)
    g = plan1.capture(5)
    assert torch.equal(g(x.to("cuda:1")).cpu(), ref)
    assert torch.cuda.current_device() == 0


@pytest.mark.skipif(not _two_devices(), reason="needs two CUDA devices")
@pytest.mark.parametrize("case,batch", [("deit_tiny_ivit", 16), ("toy_ibert_w16", 7)])
def test_sharded_logits_equal_single_device(case, batch):
    cfg, state, _ = load_case(case)
    x = synth_images(batch, cfg, seed=1)
    whole = FrozenVitPlan(cfg, state, "cuda:0").forward(x.to("cuda:0")).cpu()
    parts = []
    for rank in range(2):
        lo, hi = shard_range(batch, rank, 2)
        dev = f"cuda:{rank}"
        parts.append(FrozenVitPlan(cfg, state, dev).forward(x[lo:hi].to(dev)).cpu())
    assert torch.equal(torch.cat(parts), whole)

"""CPU: the frozen plan's host constants (scale wiring, dyadic multipliers, integer weights)
against the oracle's module-boundary scales.  No kernel runs here."""
import numpy as np
import pytest
import torch

from ivit_b200 import hostmath as hm
from ivit_b200.plan import FrozenVitPlan
from ivit_b200.synthetic import load_case, synth_images
from oracle import ivit_oracle as O
from oracle import vit_oracle as VO

f32 = np.float32


@pytest.mark.parametrize("case", ["toy_ivit", "toy_ibert", "toy_mixed"])
def test_plan_scales_match_oracle(case):
    cfg, state, _ = load_case(case)
    plan = FrozenVitPlan(cfg, state, host_only=True)
    o = VO.VitOracle(cfg, state)
    o.forward(synth_images(1, cfg, seed=0).numpy(), record=True)
    sc = {k: np.asarray(v[1], dtype=f32).reshape(-1) for k, v in o.taps.items()}
    assert plan.s_input[0] == sc["qact_input"][0]
    assert plan.patch.s_out == sc["patch_embed.qact"][0] and plan.patch.s_out2 == sc["qact1"][0]
    assert np.array_equal(plan.patch.b_scale.numpy(), sc["patch_embed.proj"])
    for i, b in enumerate(plan.blocks):
        p = f"blocks.{i}."
        assert np.array_equal(b.norm1.sf.numpy(), sc[p + "norm1"])
        assert b.norm1.s_out == sc[p + "qact1"][0]
        assert np.array_equal(b.qkv.b_scale.numpy(), sc[p + "attn.qkv"])
        assert b.qkv.s_out == sc[p + "attn.qact1"][0]
        assert f32(b.attn.s_attn) == sc[p + "attn.qact_attn1"][0]
        assert b.proj.s_out == sc[p + "attn.qact3"][0] and b.proj.s_out2 == sc[p + "qact2"][0]
        assert b.fc1.s_out == sc[p + "mlp.qact_gelu"][0]
        assert b.scales["gelu_out"] == sc[p + "mlp.act"][0]
        assert b.fc2.s_out == sc[p + "mlp.qact2"][0] and b.fc2.s_out2 == sc[p + "qact4"][0]
        assert not b.qkv.force_slow and not b.fc2.force_slow
    assert plan.norm.s_out == sc["qact2"][0]
    assert np.array_equal(plan.head.b_scale.numpy(), sc["head"])
    # class-token row and positional codes = row 0 / identity of the oracle's qact1
    codes = VO.codes_of(o.taps["qact1"])
    assert np.array_equal(plan.cls_row.numpy().astype(np.int64), codes[0, 0].astype(np.int64))
    assert np.array_equal(plan.pos_codes.numpy().astype(np.int64),
                          VO.codes_of(o.taps["qact_pos"])[0].astype(np.int64))


def test_dyadic_matches_oracle_frexp():
    rng = np.random.default_rng(3)
    a = (rng.random(300) * 1e-2 + 1e-6).astype(f32)
    a[::9] *= -1
    b = np.array([0.0173], dtype=f32)
    m, e = O.batch_frexp(a.astype(np.float64) / b.astype(np.float64))
    assert np.array_equal(hm.dyadic_mult(a, b), m.astype(np.float64) / 2.0 ** e)


def test_plan_refuses_cpu():
    cfg, state, _ = load_case("toy_ivit")
    from ivit_b200._cabi import IvitKernelError
    with pytest.raises(IvitKernelError):
        FrozenVitPlan(cfg, state, device="cpu")
    plan = FrozenVitPlan(cfg, state, host_only=True)
    with pytest.raises(IvitKernelError):
        plan.forward(torch.zeros(1, 3, 224, 224))

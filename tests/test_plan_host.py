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


@pytest.mark.parametrize("case", ["toy_ivit", "toy_ibert", "toy_mixed", "toy_ivit_w16", "toy_ibert_w16", "toy_mixed_bw"])
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


def test_plan_dtypes_follow_the_bit_widths():
    cfg, state, _ = load_case("toy_mixed_bw")            # (16, 8, 8, 16, 16, 8, 16, 8): norm2_in is 16-bit
    plan = FrozenVitPlan(cfg, state, host_only=True)
    assert plan.stream_dtype == torch.int16 and plan.pos_dtype == torch.int8
    assert plan.blocks[0].attn.output_bit == 16 and plan.blocks[0].attn.s_pv > 0
    assert (plan.blocks[0].proj.qmin, plan.blocks[0].proj.qmax) == (-32768, 32767)        # attention_out_bw = 16
    assert (plan.blocks[0].fc2.qmin2, plan.blocks[0].fc2.qmax2) == (-128, 127)            # att_block_out_bw = 8
    cfg8, state8, _ = load_case("toy_ivit")
    assert FrozenVitPlan(cfg8, state8, host_only=True).stream_dtype == torch.int8
    import dataclasses
    with pytest.raises(NotImplementedError):
        FrozenVitPlan(dataclasses.replace(cfg8, bitwidths=(32,) * 8), state8, host_only=True)


def test_swin_plan_scales_match_oracle():
    from ivit_b200.swin_plan import FrozenSwinPlan, mask_always_clamps
    from ivit_b200.synthetic import load_swin_case
    from oracle import swin_oracle as SO
    cfg, state, _ = load_swin_case("swin_toy")
    plan = FrozenSwinPlan(cfg, state, host_only=True)
    o = SO.SwinOracle(cfg, state)
    o.forward(synth_images(1, cfg, seed=0).numpy(), record=True)
    sc = {k: np.asarray(v[1], dtype=f32).reshape(-1) for k, v in o.taps.items()}
    assert plan.patch.s_out == sc["patch_embed.qact_before_norm"][0]
    assert plan.patch_norm.s_out == sc["patch_embed.qact"][0]
    for i, st in enumerate(plan.stages):
        for j, b in enumerate(st["blocks"]):
            p = f"layers.{i}.blocks.{j}."
            assert b["norm1"].s_out == sc[p + "qact1"][0] and np.array_equal(b["norm1"].sf.numpy(), sc[p + "norm1"])
            assert b["qkv"].s_out == sc[p + "attn.qact1"][0]
            assert f32(b["attn"].s2) == sc[p + "attn.qact2"][0]
            assert b["proj"].s_out == sc[p + "attn.qact4"][0] and b["proj"].s_out2 == sc[p + "qact2"][0]
            assert b["fc2"].s_out == sc[p + "mlp.qact2"][0] and b["fc2"].s_out2 == sc[p + "qact4"][0]
            assert (b["proj"].qmin2, b["proj"].qmax2) == (-32768, 32767)
        if "reduction" in st:
            assert st["reduction"].s_out == sc[f"layers.{i}.downsample.qact2"][0]
            assert int(st["reduction"].bias.abs().max()) == 0          # QuantLinear(bias=False)
    assert plan.norm.s_out == sc["qact2"][0] and np.array_equal(plan.head.b_scale.numpy(), sc["head"])
    # the mask-clamp proof: true at ordinary softmax scales, false once a masked score could win the row maximum
    assert mask_always_clamps(0.03, -128, 127) and mask_always_clamps(0.25, -128, 127)
    assert not mask_always_clamps(0.6, -128, 127) and not mask_always_clamps(1.7, -128, 127)


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

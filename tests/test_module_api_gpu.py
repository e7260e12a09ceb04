"""GPU: the frozen module path (QTensor carriers, producers fused into QuantActs) under a
vit_quant-style caller, against the oracle.  Ranges are loaded from the reference fixtures
(scripts/inference.py --strict-model-load behaviour), then freeze, then forward on CUDA."""
import numpy as np
import pytest
import torch

from ivit_b200.quantization_utils import QTensor
from ivit_b200.synthetic import load_case, synth_images
from oracle import vit_oracle as VO

from vit_caller import VisionTransformer, freeze

pytestmark = pytest.mark.gpu


def _build(case):
    cfg, state, fix = load_case(case)
    model = VisionTransformer(cfg)
    sd = {k: torch.as_tensor(v).reshape(model.state_dict()[k].shape) for k, v in state.items()
          if k in model.state_dict()}
    sd.update({k: torch.as_tensor(v) for k, v in state.items() if ".fixed_" in k})    # ppoly coefficient buffers
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not unexpected
    assert not [k for k in missing if k.endswith((".weight", ".bias", ".x_min", ".x_max"))], missing
    model.cuda()
    freeze(model)
    return cfg, state, fix, model


@pytest.mark.parametrize("case", ["toy_ivit", "toy_ibert", "toy_mixed", "deit_tiny_ivit", "toy_ivit_w16", "toy_ibert_w16",
                                  "toy_mixed_bw", "toy_ivit_sm16", "toy_ppoly", "toy_ppoly_args"])
def test_frozen_modules_match_oracle(case):
    cfg, state, fix, model = _build(case)
    x = synth_images(2, cfg, seed=0)
    taps = {}

    def hook(name):
        def f(mod, inp, out):
            if isinstance(out, tuple) and isinstance(out[0], QTensor) and out[0].codes is not None:
                taps[name] = out[0].materialize().cpu().numpy().astype(np.int64)
        return f
    for name, mod in model.named_modules():
        if type(mod).__name__ == "QuantAct":
            mod.register_forward_hook(hook(name))
    with torch.no_grad():
        out = model(x.cuda())
        assert isinstance(out, QTensor)
        logits = out.dequantize().cpu().numpy()
        again = model(x.cuda()).float().cpu().numpy()        # cached constants, tensor-method fallback
    o = VO.VitOracle(cfg, state)
    exp = o.forward(x.numpy(), record=True)
    bad = []
    for name, got in taps.items():
        want = VO.codes_of(o.taps[name]).astype(np.int64).reshape(got.shape)
        if not np.array_equal(got, want):
            bad.append(f"{name}: {(got != want).sum()}/{got.size}")
    assert len(taps) >= 20 and not bad, bad[:10]
    assert np.array_equal(logits, exp)
    assert np.array_equal(again, exp)
    assert np.array_equal(logits.argmax(1), fix["logits"][:2].argmax(1))


def test_compile_frozen_16bit_equals_module_path():
    from ivit_b200.api import compile_frozen, config_from_model
    cfg, state, fix, model = _build("toy_ibert_w16")
    assert tuple(config_from_model(model).bitwidths) == (16,) * 8
    x = synth_images(4, cfg, seed=3)
    with torch.no_grad():
        ref = model(x.cuda()).float().cpu()
    sess = compile_frozen(model, batch=4)
    assert sess.plan.stream_dtype == torch.int16
    assert torch.equal(sess.infer(x.pin_memory()), ref)


def test_state_dict_exposes_integer_weights():
    """TVM_benchmark/convert_model.py:12-66 reads weight_integer / bias_integer / *_scaling_factor."""
    cfg, state, fix, model = _build("toy_ivit")
    with torch.no_grad():
        model(synth_images(1, cfg, seed=0).cuda())
    sd = model.state_dict()
    w = sd["blocks.0.attn.qkv.weight_integer"].cpu().numpy()
    assert np.abs(w).max() == 127 and np.array_equal(w, np.rint(w))
    assert float(sd["blocks.0.attn.qkv.fc_scaling_factor"].min()) > 0
    assert float(sd["blocks.0.qact1.act_scaling_factor"]) > 0


def test_compile_frozen_equals_module_path():
    from ivit_b200.api import compile_frozen, config_from_model
    cfg, state, fix, model = _build("toy_mixed")
    got = config_from_model(model)
    assert (got.embed_dim, got.depth, got.num_heads, got.patch_size, got.img_size) == \
        (cfg.embed_dim, cfg.depth, cfg.num_heads, cfg.patch_size, cfg.img_size)
    assert (got.gelu_type, got.softmax_type, got.layernorm_type) == (cfg.gelu_type, cfg.softmax_type, cfg.layernorm_type)
    x = synth_images(4, cfg, seed=3)
    with torch.no_grad():
        ref = model(x.cuda()).float().cpu()
    sess = compile_frozen(model, batch=4)
    out = sess.infer(x.pin_memory()).clone()
    assert torch.equal(out, ref)
    sess2 = compile_frozen(model, batch=4, chunks=2)
    assert torch.equal(sess2.infer(x.pin_memory()), ref)


def test_float_gelu_and_layernorm_under_frozen_quantacts():
    """The `float` registry entries (diagnostics: true float op, quantised output) inside a frozen
    model: their float outputs re-enter the integer path through round(x / scale) + the requant
    kernel.  Against the same model in float-simulated mode on the same device (floating-point
    modules: logits within 2 % of the logit range, same argmax)."""
    import dataclasses
    cfg, state, fix = load_case("toy_ivit")
    cfg = dataclasses.replace(cfg, gelu_type="float", layernorm_type="float")
    torch.manual_seed(1)
    model = VisionTransformer(cfg)
    sd = {k: torch.as_tensor(v).reshape(model.state_dict()[k].shape) for k, v in state.items()
          if k in model.state_dict() and not k.endswith((".x_min", ".x_max"))}
    model.load_state_dict(sd, strict=False)
    model.cuda().eval()
    x = synth_images(4, cfg, seed=0).cuda()
    with torch.no_grad():
        sim = model(synth_images(4, cfg, seed=1).cuda())       # calibration pass (running statistics)
        for m in model.modules():
            if hasattr(m, "running_stat"):
                m.running_stat = False
        sim = model(x).float()                                 # float-simulated, ranges fixed
        freeze(model)
        out = model(x)
    out = (out.dequantize() if isinstance(out, QTensor) else out).float()
    assert out.shape == sim.shape and torch.isfinite(out).all()
    span = float(sim.max() - sim.min())
    assert float((out - sim).abs().max()) <= 0.02 * span
    assert torch.equal(out.argmax(-1), sim.argmax(-1))

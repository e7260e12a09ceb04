"""GPU parity of the frozen Swin forward (FrozenSwinPlan -> C ABI -> sm_100a kernels) against the numpy
oracle at every materialised module boundary, and against the committed logits of the unmodified
reference (tests/golden/swin_*.npz).  Bar: bit-exact integer codes, logits == oracle exactly and within
1e-6 relative of the reference's float32-carried logits."""
import numpy as np
import pytest
import torch

from ivit_b200 import _cabi, hostmath as hm, ops
from ivit_b200.swin_plan import FrozenSwinPlan
from ivit_b200.synthetic import load_swin_case, synth_images
from oracle import ivit_oracle as O
from oracle import swin_oracle as SO

pytestmark = pytest.mark.gpu
f32 = np.float32


def _to_image_order(codes, cfg, i, j, B):
    """oracle tap of attn.qact3 [B*nW, N, C] (window order) -> [B, L, C] (image token order)."""
    res, C = cfg.stage_res(i), cfg.stage_dim(i)
    ws, shift = cfg.block_window_shift(i, j)
    x = SO.window_reverse(codes.reshape(-1, ws, ws, C), ws, res, res)
    if shift > 0:
        x = np.roll(x, (shift, shift), axis=(1, 2))
    return x.reshape(B, res * res, C)


def _run_case(case, batch, seed=0):
    cfg, state, fix = load_swin_case(case)
    x = synth_images(batch, cfg, seed=seed)
    plan = FrozenSwinPlan(cfg, state, "cuda")
    taps = {}
    logits = plan.forward(x.cuda(), taps=taps).cpu().numpy()
    o = SO.SwinOracle(cfg, state)
    exp_logits = o.forward(x.numpy(), record=True)
    bad = []
    for name, t in taps.items():
        got = t.cpu().numpy().astype(np.int64)
        if name.endswith(".imgorder"):
            base = name[:-len(".imgorder")]
            i, j = int(base.split(".")[1]), int(base.split(".")[3])
            exp = _to_image_order(SO.codes_of(o.taps[base]), cfg, i, j, batch).astype(np.int64)
        else:
            exp = SO.codes_of(o.taps[name]).astype(np.int64).reshape(got.shape)
        n = int((got != exp).sum())
        if n:
            idx = tuple(np.argwhere(got != exp)[0])
            bad.append(f"{name}: {n}/{got.size} (first {idx}: got {got[idx]} exp {exp[idx]})")
    assert not bad, f"{case}: {len(bad)}/{len(taps)} boundaries differ:\n" + "\n".join(bad[:12])
    assert np.array_equal(logits, exp_logits), np.abs(logits - exp_logits).max()
    return cfg, plan, logits, fix, len(taps)


def test_swin_toy_all_boundaries():
    """Two stages: 14x14 tokens in four 7x7 windows (block 1 shifted, masked), PatchMerging, one 7x7 window."""
    _, _, logits, fix, ntaps = _run_case("swin_toy", 2)
    assert ntaps >= 30
    ref = fix["logits"]
    assert np.array_equal(logits.argmax(1), ref.argmax(1))
    assert np.allclose(logits, ref, rtol=1e-6, atol=1e-6 * np.abs(ref).max())


@pytest.mark.parametrize("case", ["swin_toy_b", "swin_toy_w4"])
def test_swin_other_shapes(case):
    """Swin-B widths (C = 128 / 256, 4 and 8 heads, LayerNorm over 512 channels) and 4 x 4 windows (16 tokens)."""
    _, _, logits, fix, ntaps = _run_case(case, 2)
    ref = fix["logits"]
    assert np.array_equal(logits.argmax(1), ref.argmax(1))
    assert np.allclose(logits, ref, rtol=1e-6, atol=1e-6 * np.abs(ref).max())


def test_swin_tiny_graph_session_and_permutation_invariance():
    """Swin-T at batch 32 through InferenceSession (CUDA graph, host in / host out): every image's logits are
    independent of its position in the batch."""
    from ivit_b200.api import InferenceSession
    cfg, state, _ = load_swin_case("swin_tiny")
    sess = InferenceSession(cfg, state, batch=32)
    x = synth_images(32, cfg, seed=9)
    full = sess.infer(x.pin_memory()).clone()
    perm = torch.randperm(32, generator=torch.Generator().manual_seed(0))
    assert torch.equal(sess.infer(x[perm].contiguous().pin_memory()), full[perm])
    assert torch.equal(sess.plan.forward(x[:5].cuda()).cpu(), full[:5])


def test_swin_toy_fresh_inputs_batch3():
    _run_case("swin_toy", 3, seed=5)


def test_swin_single_image_and_empty_batch():
    _run_case("swin_toy", 1, seed=6)
    cfg, state, _ = load_swin_case("swin_toy")
    plan = FrozenSwinPlan(cfg, state, "cuda")
    out = plan.forward(torch.zeros((0, 3, cfg.img_size, cfg.img_size), device="cuda"))
    torch.cuda.synchronize()
    assert tuple(out.shape) == (0, cfg.num_classes)


def test_swin_tiny_configs3():
    """BASELINE.json configs[3]: Swin-Tiny INT8 window attention (shifted windows), 224x224."""
    _, _, logits, fix, ntaps = _run_case("swin_tiny", 2)
    assert ntaps >= 90
    ref = fix["logits"]
    assert np.array_equal(logits.argmax(1), ref.argmax(1))
    assert np.allclose(logits, ref, rtol=1e-6, atol=1e-6 * np.abs(ref).max())


def test_swin_tiny_batch_invariance():
    cfg, state, _ = load_swin_case("swin_tiny")
    plan = FrozenSwinPlan(cfg, state, "cuda")
    x = synth_images(8, cfg, seed=3).cuda()
    full = plan.forward(x).clone()
    assert torch.equal(plan.forward(x[2:5].contiguous()), full[2:5])


def test_compile_frozen_from_a_module_api_swin_model():
    """The drop-in route for Swin: a model built on ivit_b200.quantization_utils with swin_quant.py's module
    names, calibrated ranges loaded (scripts/inference.py --strict-model-load behaviour), frozen, then
    compile_frozen -> FrozenSwinPlan -> InferenceSession: host images in, host logits out == oracle."""
    from ivit_b200.api import compile_frozen
    from swin_caller import SwinTransformer
    cfg, state, fix = load_swin_case("swin_toy")
    model = SwinTransformer(cfg)
    sd = {k: torch.as_tensor(v).reshape(model.state_dict()[k].shape) for k, v in state.items() if k in model.state_dict()}
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not unexpected
    assert not [k for k in missing if k.endswith((".weight", ".bias", ".x_min", ".x_max"))], missing
    model.cuda().eval()
    for m in model.modules():
        if hasattr(m, "fix"):
            m.fix()
    sess = compile_frozen(model, batch=4)
    x = synth_images(4, cfg, seed=7)
    got = sess.infer(x.pin_memory()).clone().numpy()
    exp = SO.SwinOracle(cfg, state).forward(x.numpy())
    assert np.array_equal(got, exp)
    sess2 = compile_frozen(model, batch=4, chunks=2)
    assert np.array_equal(sess2.infer(x.pin_memory()).numpy(), exp)


def _window_attention_case(res, ws, shift, nH, hd, seed, s2_target, use_table, B=2, bias_terms=False):
    """ivit_window_attention_i8 alone against the oracle's statement of WindowAttention.forward (swin_quant.py:131-160)
    on random int8 qkv codes, for a chosen softmax-input scale s2 (decides whether a masked score can be proven to clamp)."""
    from ivit_b200.swin_plan import mask_always_clamps, relative_position_index
    rng = np.random.default_rng(seed)
    C, L, N = nH * hd, res * res, ws * ws
    qkv = np.clip(np.rint(rng.standard_normal((B, L, 3 * C)) * 40), -128, 127).astype(np.int8)
    s1 = np.array([0.05], dtype=f32)
    x = (qkv.astype(f32) * s1).astype(f32).reshape(B, res, res, 3 * C)
    if shift:
        x = np.roll(x, (-shift, -shift), axis=(1, 2))
    xw = SO.window_partition(x, ws).reshape(-1, N, 3, nH, hd).transpose(2, 0, 3, 1, 4)
    q, k, v = xw[0], xw[1], xw[2]
    attn, s, _ = O.quant_matmul(q, s1, np.swapaxes(k, -2, -1), s1)
    sc = f32(hd ** -0.5)
    attn, s = (attn * sc).astype(f32), (s * sc).astype(f32)
    a1 = np.array([np.abs(attn).max() * 0.6], dtype=f32)
    attn, s_a1, _ = O.quant_act(attn, 8, -a1, a1, pre_scale=s)
    tab = (rng.standard_normal(((2 * ws - 1) ** 2, nH)) * 0.5).astype(f32)
    tmax = np.array([np.abs(tab).max()], dtype=f32)
    tabq, s_tab, tab_codes = O.quant_act(tab, 8, -tmax, tmax)
    idx = relative_position_index(ws).reshape(-1)
    bias = tabq[idx].reshape(N, N, nH).transpose(2, 0, 1)
    a2 = np.array([127 * s2_target], dtype=f32)
    attn, s2, _ = O.quant_act(attn, 8, -a2, a2, pre_scale=s_a1, identity=np.broadcast_to(bias[None], attn.shape),
                              identity_scale=s_tab)
    mask = SO.shift_mask(res, ws, shift)
    if mask is not None:
        nW = mask.shape[0]
        attn = (attn.reshape(-1, nW, nH, N, N) + mask[None, :, None]).astype(f32).reshape(-1, nH, N, N)
    pr, s_p, pr_codes = O.ivit_softmax(attn, s2, 8)
    o, s_o, _ = O.quant_matmul(pr, s_p, v, s1)
    o = o.transpose(0, 2, 1, 3).reshape(-1, N, C)
    omax = np.array([np.abs(o).max() * 0.7], dtype=f32)
    _, s3, out_codes = O.quant_act(o, 8, -omax, omax, pre_scale=s_o)
    exp = SO.window_reverse(out_codes.reshape(-1, ws, ws, C), ws, res, res)
    if shift:
        exp = np.roll(exp, (shift, shift), axis=(1, 2))
    p = _cabi.WindowAttentionParams()
    p.mult_attn = float(hm.dyadic_mult(s, s_a1).reshape(-1)[0])
    p.qmin_attn, p.qmax_attn, p.qmin2, p.qmax2, p.qmin_out, p.qmax_out = -128, 127, -128, 127, -128, 127
    p.mult2 = float(hm.dyadic_mult(s_a1, s2).reshape(-1)[0])
    p.mult_bias = float(hm.dyadic_mult(s_tab, s2).reshape(-1)[0])
    p.s2, p.output_bit = float(s2[0]), 8
    p.mult_out = float(hm.dyadic_mult(s_o, s3).reshape(-1)[0])
    bias_codes = torch.from_numpy(np.ascontiguousarray(
        tab_codes[idx].reshape(N, N, nH).transpose(2, 0, 1)).astype(np.int8)).cuda()
    p.bias_codes = bias_codes.data_ptr()
    terms = None
    if bias_terms:              # the identity branch of qact2 evaluated once per layer (what FrozenSwinPlan passes)
        terms = torch.from_numpy(np.rint(bias_codes.cpu().numpy().astype(np.float64) * p.mult_bias).astype(np.int32)).cuda()
        p.bias_terms = terms.data_ptr()
    clamps = mask_always_clamps(s2[0], -128, 127)
    keep = None
    if use_table:
        tp = _cabi.AttentionParams()
        tp.kind, tp.output_bit, tp.s_attn = 0, 8, float(s2[0])
        keep = ops.softmax_table_build(tp, "cuda", collapse=False)
        p.exp_table, p.mask_clamps = tp.exp_table, int(clamps)
        if seed % 2 == 0:
            p.sat_k, p.sat_e = tp.sat_k, tp.sat_e          # the saturated tail of the table need not be gathered
    out, probs = ops.window_attention(torch.from_numpy(qkv).cuda(), B, res, ws, shift, nH, hd, p, want_probs=True)
    assert np.array_equal(probs.cpu().numpy().astype(np.int64), pr_codes.astype(np.int64)), "softmax codes"
    assert np.array_equal(out.cpu().numpy().reshape(exp.shape).astype(np.int64), exp.astype(np.int64)), "output codes"
    # window-order form (params.windowed = 1, the module path's caller): qkv already rolled and partitioned, output in
    # WindowAttention.forward's own [B * nW, N, C] order
    qw = qkv.reshape(B, res, res, 3 * C)
    if shift:
        qw = np.roll(qw, (-shift, -shift), axis=(1, 2))
    qw = np.ascontiguousarray(SO.window_partition(qw, ws).reshape(B, L, 3 * C))
    p.windowed = 1
    out_w, probs_w = ops.window_attention(torch.from_numpy(qw).cuda(), B, res, ws, shift, nH, hd, p, want_probs=True)
    assert np.array_equal(probs_w.cpu().numpy().astype(np.int64), pr_codes.astype(np.int64)), "softmax codes (windowed)"
    assert np.array_equal(out_w.cpu().numpy().reshape(-1, N, C).astype(np.int64), out_codes.reshape(-1, N, C).astype(np.int64)), \
        "output codes (windowed)"
    del keep, terms
    return clamps


@pytest.fixture(params=[0, 1, 2], ids=["rows", "dp4a", "tcgen05"])
def window_variant(request):
    prev = ops.set_window_attention_variant(request.param)
    yield request.param
    ops.set_window_attention_variant(prev)


def test_swin_toy_on_the_dp4a_window_kernel():
    prev = ops.set_window_attention_variant(1)
    try:
        _run_case("swin_toy", 2)
    finally:
        ops.set_window_attention_variant(prev)


@pytest.mark.parametrize("use_table", [False, True])
@pytest.mark.parametrize("s2_target", [0.03, 0.25, 0.6, 1.7])
def test_window_attention_kernel(s2_target, use_table, window_variant):
    """Computed and tabulated exp terms, without and with the shift mask, from scales at which a masked score
    provably clamps (table path) up to scales at which it does not (masked scores can win the row maximum:
    the kernel must then take the computed path even when a table is attached)."""
    _window_attention_case(14, 7, 0, 3, 32, seed=1, s2_target=s2_target, use_table=use_table)
    clamps = _window_attention_case(14, 7, 3, 3, 32, seed=2, s2_target=s2_target, use_table=use_table)
    assert clamps == (s2_target < 0.4)
    _window_attention_case(8, 4, 2, 2, 32, seed=3, s2_target=s2_target, use_table=use_table)      # 4x4 windows
    _window_attention_case(7, 7, 0, 6, 32, seed=4, s2_target=s2_target, use_table=use_table)      # one window, 6 heads
    _window_attention_case(16, 8, 4, 2, 32, seed=5, s2_target=s2_target, use_table=use_table)     # 8x8 windows: 64 tokens
    # odd number of (window, head) units, 16 windows (masks on the last window row AND column), precomputed bias terms
    _window_attention_case(7, 7, 0, 3, 32, seed=7, s2_target=s2_target, use_table=use_table, B=3)
    _window_attention_case(28, 7, 3, 3, 32, seed=8, s2_target=s2_target, use_table=use_table, B=1, bias_terms=True)
    _window_attention_case(14, 7, 3, 6, 32, seed=9, s2_target=s2_target, use_table=use_table, B=5, bias_terms=True)
    if window_variant == 1:
        _window_attention_case(14, 7, 3, 2, 64, seed=6, s2_target=s2_target, use_table=use_table)  # head_dim 64: dp4a only


def test_lifecycle_freeze_runs_a_swin_model_on_the_kernels():
    """lifecycle: load ranges -> freeze_model -> model(x) on CUDA == oracle; unfreeze_model restores the module forward."""
    from ivit_b200 import lifecycle
    from swin_caller import SwinTransformer
    cfg, state, _ = load_swin_case("swin_toy")
    model = SwinTransformer(cfg)
    lifecycle.load_checkpoint(model, {"model": {k: torch.as_tensor(v) for k, v in state.items()}})
    model.cuda().eval()
    lifecycle.freeze_model(model)
    x = synth_images(3, cfg, seed=4)
    with torch.no_grad():
        got = model(x.cuda()).cpu().numpy()
    assert np.array_equal(got, SO.SwinOracle(cfg, state).forward(x.numpy()))
    lifecycle.unfreeze_model(model)
    assert "forward" not in model.__dict__ and model.qact_input.running_stat is True


def test_patch_merge_gather_and_avgpool():
    rng = np.random.default_rng(0)
    B, res, C = 3, 8, 96
    x = rng.integers(-32768, 32768, size=(B, res, res, C)).astype(np.int16)
    exp = np.concatenate([x[:, 0::2, 0::2], x[:, 1::2, 0::2], x[:, 0::2, 1::2], x[:, 1::2, 1::2]], axis=-1)
    got = ops.patch_merge_gather(torch.from_numpy(x).cuda(), B, res, C).cpu().numpy()
    assert np.array_equal(got.reshape(exp.shape), exp)
    for L, Cc in ((49, 768), (196, 192), (7, 96)):
        q = rng.integers(-128, 128, size=(B, L, Cc)).astype(np.int8)
        s = np.array([f32(0.0371)])
        xin = (q.astype(f32) * s).astype(f32)
        pooled = SO.avgpool_tokens(xin)
        x_max = np.array([np.abs(pooled).max() * 0.8], dtype=f32)
        _, s_out, codes = O.quant_act(pooled[:, :, None], 8, -x_max, x_max, pre_scale=s)
        mult = float(hm.dyadic_mult(s, s_out).reshape(-1)[0])
        got = ops.avgpool_requant(torch.from_numpy(q).cuda(), B, L, Cc, float(s[0]), mult, -128, 127).cpu().numpy()
        assert np.array_equal(got.astype(np.int64), codes[:, :, 0].astype(np.int64)), (L, Cc)


@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("C,tokens,bits", [(96, 196, 16), (96, 64, 16), (128, 49, 16), (96, 196, 8), (192, 100, 16),
                                           (96, 64, 8), (96, 160, 16), (96, 3136, 16), (384, 64, 16), (128, 64, 16),
                                           (128, 96, 8)])
def test_layernorm_strided_order(C, tokens, bits, variant):
    ops.set_layernorm_variant(variant)          # 0: two-lanes-per-row kernels where they apply; 1: division kernels
    try:
        _layernorm_strided_order(C, tokens, bits)
    finally:
        ops.set_layernorm_variant(0)


def _layernorm_strided_order(C, tokens, bits):
    """LayerNorm whose reference input is channel-strided: torch's outer-reduction order of the mean, both
    forms (whole 32-token blocks / tail tokens).  16-bit codes make the order matter on ordinary rows."""
    rng = np.random.default_rng(C + tokens)
    B = 2
    lim = 2 ** (bits - 1)
    q = np.clip(np.rint(rng.standard_normal((B, tokens, C)) * lim / 3.5), -lim, lim - 1).astype(np.int64)
    for r in range(0, tokens, 2):            # exact mean ties on half of the rows
        d = int(rng.integers(-3, 4)) * C + C // 2 - int(q[0, r].sum())
        np.add.at(q[0, r], rng.integers(0, C, size=abs(d)), np.sign(d))
        q[0, r] = np.clip(q[0, r], -lim, lim - 1)
    s_in = f32(3.1e-4 if bits == 16 else 0.0371)
    w = (rng.random(C) + 0.5).astype(f32)
    b = (rng.standard_normal(C) * 0.1).astype(f32)
    y, sf, _ = O.ivit_layernorm((q.astype(f32) * s_in).astype(f32), np.array([s_in]), w, b, strided_input=True)
    y_in, _, _ = O.ivit_layernorm((q.astype(f32) * s_in).astype(f32), np.array([s_in]), w, b, strided_input=False)
    x_max = np.array([np.abs(y).max() * 0.7], dtype=f32)
    _, s_out, codes = O.quant_act(y, 8, -x_max, x_max, pre_scale=sf)
    p = hm.layernorm_params(w, b, C)
    mult = hm.dyadic_mult(p["sf"], s_out)
    dt = torch.int16 if bits == 16 else torch.int8
    xt = torch.from_numpy(q.reshape(B * tokens, C)).to(dt).cuda()
    dev = lambda a, d: torch.from_numpy(np.ascontiguousarray(a)).to(d).cuda()
    got = ops.layernorm_requant_strided(0, xt, tokens, float(s_in), dev(p["bias_int"], torch.float32),
                                        dev(p["sf"], torch.float32), dev(mult, torch.float64), -128, 127).cpu().numpy()
    assert np.array_equal(got.astype(np.int64), codes.reshape(B * tokens, C).astype(np.int64))
    assert not np.array_equal(y, y_in) or bits == 8     # the two orders really differ on this input


@pytest.mark.parametrize("case,batch", [("swin_toy", 2), ("swin_toy_w4", 3), ("swin_toy_b", 2), ("swin_tiny", 1)])
def test_frozen_module_path_runs_swin_glue_on_the_carrier(case, batch):
    """Row a16: with only fix() called (the reference's own freeze_model), the caller's forward runs module by module on
    integer carriers -- torch.roll, window_partition / window_reverse, the bias gather, the mask add, PatchMerging's
    slices + cat and the token average pool included (swin_quant.py:18-50, :140-153, :258-289, :339-345, :554).  Every
    QuantAct output against the oracle's tap of the same name, logits against the oracle and the fused plan."""
    from ivit_b200 import lifecycle
    from ivit_b200.quantization_utils import QTensor
    from swin_caller import SwinTransformer
    cfg, state, fix = load_swin_case(case)
    model = SwinTransformer(cfg)
    lifecycle.load_checkpoint(model, {"model": {k: torch.as_tensor(v) for k, v in state.items()}})
    model.cuda().eval()
    lifecycle.freeze_model(model, fused_plan=False)
    assert "forward" not in model.__dict__
    taps = {}

    def hook(name):
        def f(mod, inp, out):
            if isinstance(out, tuple) and isinstance(out[0], QTensor) and out[0].codes is not None:
                taps[name] = out[0].materialize().cpu().numpy().astype(np.int64)
        return f
    for name, mod in model.named_modules():
        if type(mod).__name__ == "QuantAct":
            mod.register_forward_hook(hook(name))
    x = synth_images(batch, cfg, seed=0)
    with torch.no_grad():
        out = model(x.cuda())
        assert isinstance(out, QTensor)
        logits = out.dequantize().cpu().numpy()
        again = model(x.cuda()).float().cpu().numpy()          # cached constants
    o = SO.SwinOracle(cfg, state)
    exp = o.forward(x.numpy(), record=True)
    bad, seen = [], 0
    for name, got in taps.items():
        if name not in o.taps:
            continue
        want = SO.codes_of(o.taps[name]).astype(np.int64)
        if want.size != got.size:
            continue                                            # qact_table: the oracle taps the gathered bias
        seen += 1
        if not np.array_equal(got.reshape(-1), want.reshape(-1)):
            bad.append(f"{name}: {(got.reshape(-1) != want.reshape(-1)).sum()}/{got.size}")
    assert seen >= 30 and not bad, (seen, bad[:12])
    assert np.array_equal(logits, exp) and np.array_equal(again, exp)
    plan = FrozenSwinPlan(cfg, state, "cuda")
    assert np.array_equal(plan.forward(x.cuda()).cpu().numpy(), logits)

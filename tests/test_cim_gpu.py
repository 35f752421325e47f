"""cim_layers path on the B200: quantiser kernels against the pinned oracle (integer codes and
lazily initialised step sizes bit-exact), the LSQ drop-in layers forward + backward, and the ADDA compute-in-memory
layer (bit slices, mapped blocks, ADC) — exact when no weight noise is injected."""
import pytest
import torch

from util import load_golden, relerr

pytestmark = pytest.mark.gpu


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def _bfr(x):
    """Round to bf16-representable fp32: the product's activations are bf16, the oracle then sees the same numbers."""
    return x.to(torch.bfloat16).float()


# --------------------------------------------------------------------------------------------------- kernels
@pytest.mark.parametrize("bits,rows,c,dtype", [(8, 512, 288, torch.bfloat16), (4, 100, 64, torch.bfloat16),
                                               (8, 77, 96, torch.float32), (2, 33, 8, torch.float32)])
def test_lsq_codes_and_lazy_init_are_bit_exact(bits, rows, c, dtype):
    dev = _dev()
    from oracle import cim_torch as OC
    from sdb200 import runtime as rt
    g = torch.Generator().manual_seed(bits * 100 + c)
    x = torch.randn(rows, c, generator=g) * 2.5
    x = _bfr(x) if dtype == torch.bfloat16 else x
    xd = x.to(dev, dtype)
    qp = float(OC.qp_of(bits))
    step = torch.ones((), device=dev)
    rt.lsq_init_step(xd, step, qp)
    want = torch.as_tensor(OC.init_step(x, bits), dtype=torch.float32)
    assert step.item() == want.item()                                   # bit-exact lazy initialisation
    rt.lsq_init_step(xd * 3, step, qp)
    assert step.item() == want.item()                                   # fires only while the step size is 1.0
    for s in (want, want * 0.37):                                       # a step that clamps a good part of the range
        sd = s.to(dev)
        # isint=True output: q * s' / s' with s' the VALUE of the grad-scaled step — the reference's own arithmetic,
        # including the cases where it lands one ulp off the integer
        ref_c, _ = OC.quant_lsq(x, bits, s, True)
        got_c = rt.lsq_quant(xd, sd, qp, codes=True, out_dtype=torch.float32)
        assert torch.equal(got_c.cpu(), ref_c)
        assert torch.equal(torch.round(got_c).cpu(), OC.codes(x, bits, s))          # integer codes bit-exact
        vals = rt.lsq_quant(xd, sd, qp, out_dtype=torch.float32)
        ref, _ = OC.quant_lsq(x, bits, s, False)
        assert torch.equal(vals.cpu(), ref)                                         # isint=False output bit-exact
        # DAC digits computed in the same pass as the codes
        if bits == 8:
            want_d = OC.bit_split(ref_c, 8, 4)
            got_d = rt.bit_split(xd, 2, 4, 8, step=sd, qp=qp)
            assert torch.equal(got_d.float().cpu(), want_d)
    zero = torch.ones((), device=dev)
    rt.lsq_init_step(torch.zeros(8, 8, device=dev, dtype=dtype), zero, qp)
    assert zero.item() == 1.0                                           # all-zero tensor leaves the step at 1


@pytest.mark.parametrize("codes", [False, True])
def test_lsq_quant_backward_vs_oracle_autograd(codes):
    dev = _dev()
    from oracle import cim_torch as OC
    from sdb200 import runtime as rt
    g = torch.Generator().manual_seed(3)
    rows, c, bits = 300, 96, 8
    x = _bfr(torch.randn(rows, c, generator=g) * 2)
    dy = _bfr(torch.randn(rows, c, generator=g))
    s = torch.tensor(0.011)
    xr, sr = x.clone().requires_grad_(True), s.clone().requires_grad_(True)
    y, _ = OC.quant_lsq(xr, bits, sr, codes)
    y.backward(dy)
    dx, ds = rt.lsq_quant_bwd(dy.to(dev, torch.bfloat16), x.to(dev, torch.bfloat16), s.to(dev), float(OC.qp_of(bits)),
                              codes=codes)
    assert relerr(dx.float().cpu(), xr.grad) < 4e-3                     # bf16 rounding of dx (exact mask)
    assert abs(ds.item() - sr.grad.item()) <= 2e-4 * abs(sr.grad.item()) + 1e-7


def test_bit_split_and_adc_combine_match_oracle():
    dev = _dev()
    from oracle import cim_torch as OC
    from sdb200 import runtime as rt
    g = torch.Generator().manual_seed(8)
    q = torch.randint(-127, 128, (64, 40), generator=g).float()
    q[0, :4] = torch.tensor([-127., 127., 0., -16.])
    for x_bit, slice_bit in [(8, 4), (8, 3), (5, 4), (9, 2)]:
        qq = q.clamp(-(2 ** (x_bit - 1) - 1), 2 ** (x_bit - 1) - 1)
        want = OC.bit_split(qq, x_bit, slice_bit)
        slices = want.shape[0] // qq.shape[0]
        got = rt.bit_split(qq.to(dev, torch.bfloat16), slices, slice_bit, x_bit)
        assert torch.equal(got.float().cpu(), want), (x_bit, slice_bit)
        w = 2.0 ** (slice_bit * torch.arange(slices))
        assert torch.equal((want.reshape(slices, *qq.shape) * w[:, None, None]).sum(0), qq)   # digits recombine
    # ADC: scale, noise, clamp, round, recombine, accumulate
    rows, ncol, slices, slice_bit, R = 50, 24, 2, 4, 127.0
    part = torch.randn(slices * rows, ncol, generator=g) * 2000
    alpha = torch.tensor([0.1417])
    gn, on = torch.randn(1000, generator=g) * 0.02, torch.randn(1000, generator=g) * 0.01
    out0 = torch.randn(rows, 64, generator=g).round()
    for noise in (False, True):
        a = alpha * part
        if noise:
            a = a * (1 + gn[8:8 + ncol]) + R * on[8:8 + ncol]
        a = torch.round(torch.clamp(a, -R - 1, R)).reshape(slices, rows, ncol)
        want = out0.clone()
        want[:, 8:8 + ncol] += a[0] + 16 * a[1]
        got = out0.clone().to(dev)
        rt.adc_combine(part.to(dev), got, rows, ncol, 8, slices, slice_bit, alpha.to(dev), R,
                       gn.to(dev) if noise else None, on.to(dev) if noise else None)
        assert torch.equal(got.cpu(), want)


# --------------------------------------------------------------------------------------------------- LSQ layers
def _make_lsq(name, f, dev):
    from cim_layers.layers_qn_lsq import Conv2d_qn_lsq, Linear_qn_lsq
    cfg = f["cfg"]
    w = f["init"]["weight"]
    kw = dict(weight_bit=cfg["weight_bit"], input_bit=cfg["input_bit"], output_bit=cfg["output_bit"],
              noise_scale=cfg["noise_scale"], bias="bias" in f["init"])
    if f["conv"] is None:
        layer = Linear_qn_lsq(w.shape[1], w.shape[0], **kw)
    else:
        layer = Conv2d_qn_lsq(w.shape[1], w.shape[0], w.shape[2], f["conv"]["stride"], f["conv"]["padding"], 1, **kw)
    layer.load_state_dict(f["init"])
    layer = layer.to(dev)
    layer.weight_noise_unit = f["unit"].float().to(dev)
    return layer


@pytest.mark.parametrize("name", ["lsq_linear", "lsq_linear_clean", "lsq_conv3x3", "lsq_conv1x1"])
def test_lsq_layer_forward_backward_vs_oracle(name):
    dev = _dev()
    from oracle import cim_torch as OC
    f = load_golden("cim_layers")[name]
    layer = _make_lsq(name, f, dev)
    unit = f["unit"].float()
    x, x2 = _bfr(f["x"]), _bfr(f["x2"])
    # oracle live on the bf16-representable inputs (it is pinned bit-exact to the reference by tests/test_oracle.py)
    p = {k: v.clone() for k, v in f["init"].items()}
    with torch.no_grad():
        ref1 = OC.lsq_layer(p, f["cfg"], x, unit, f["conv"])
        out1 = layer(x.to(dev))
    torch.cuda.synchronize()
    # lazily initialised step sizes: input and weight exactly, output from the bf16 tensor-core product
    assert layer.step_size_input.item() == p["step_size_input"].item()
    assert layer.step_size_weight.item() == p["step_size_weight"].item()
    assert abs(layer.step_size_output.item() / p["step_size_output"].item() - 1) < 1e-2
    layer.step_size_output.data.fill_(p["step_size_output"].item())     # same output grid for the code comparison
    p2 = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    ref2 = OC.lsq_layer(p2, f["cfg"], x2, unit, f["conv"])
    tgt = f["target"]
    ((ref2 - tgt) ** 2).mean().backward()
    # yardstick for the gradients: the same layer under the reference's own training regime (bf16 autocast)
    p16 = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    with torch.autocast("cpu", dtype=torch.bfloat16):
        ref16 = OC.lsq_layer(p16, f["cfg"], x2, unit, f["conv"])
    ((ref16.float() - tgt) ** 2).mean().backward()
    out2 = layer(x2.to(dev))
    ((out2 - tgt.to(dev)) ** 2).mean().backward()
    torch.cuda.synchronize()
    s_out = p["step_size_output"].item()
    # the layer hands bf16 rows to the next layer (code * step rounded to 8 mantissa bits, < 0.26 step at code 127):
    # compare the integer codes the two outputs stand for
    dcode = (torch.round(out2.detach().cpu() / s_out) - torch.round(ref2.detach() / s_out)).abs()
    frac_same = (dcode < 0.5).float().mean().item()
    print(f"{name}: output codes identical {100 * frac_same:.2f}% max |dcode| {dcode.max().item():.1f}; golden(fp32 inputs) "
          f"rel {relerr(out1.cpu(), f['y1']):.3e}")
    # bf16 operands (the reference's own autocast regime) move a pre-quantisation value across a rounding boundary
    # now and then: at most one output step, on a small fraction of the outputs
    assert dcode.max().item() <= 1 and frac_same > 0.90
    assert relerr(out1.cpu(), f["y1"]) < 3e-2
    # precise parity mode: the product runs in the CODE domain (integer input codes against integer weight codes, or
    # against the noisy fp32 weight as a (hi, lo) pair of bf16 operands) with fp32 accumulation — what is left between
    # the two outputs is fp32 summation order
    from sdb200 import engine as E
    E.set_parity_mode("precise")
    try:
        with torch.no_grad():
            out3 = layer(x2.to(dev))
    finally:
        E.set_parity_mode("fast")
    dcode3 = (torch.round(out3.cpu() / s_out) - torch.round(ref2.detach() / s_out)).abs()
    same3 = (dcode3 < 0.5).float().mean().item()
    print(f"{name}: precise mode: output codes identical {100 * same3:.3f}% max |dcode| {dcode3.max().item():.1f}")
    assert dcode3.max().item() <= 1 and same3 > 0.999
    for k, v in layer.named_parameters():
        ref_g = p2[k].grad
        got = v.grad.cpu()
        err = ((got - ref_g).norm() / ref_g.norm().clamp_min(1e-20)).item()
        yard = ((p16[k].grad - ref_g).norm() / ref_g.norm().clamp_min(1e-20)).item()
        print(f"   grad {k}: rel-L2 {err:.3e} (reference under bf16 autocast: {yard:.3e})")
        # step-size gradients are sums of G * (q - u) with mixed signs: ill-conditioned against any perturbation of the
        # pre-quantisation values, so they are gated on the reference's own bf16-autocast error
        assert err < max(6e-2 if v.dim() == 0 else 3e-2, 1.5 * yard), k


def test_lsq_layer_update_para_and_use_fp():
    dev = _dev()
    from cim_layers.layers_qn_lsq import Linear_qn_lsq
    torch.manual_seed(0)
    layer = Linear_qn_lsq(64, 32, weight_bit=4, input_bit=8, output_bit=8, noise_scale=0.0, bias=True).to(dev)
    x = torch.randn(16, 64, device=dev)
    with torch.no_grad():
        layer(x)
    s = [layer.step_size_input.item(), layer.step_size_weight.item(), layer.step_size_output.item()]
    layer.update_para(weight_bit=6, input_bit=7, output_bit=8, noise_scale=0.0)
    assert layer.step_size_weight.item() == s[1] / 4 and layer.step_size_input.item() == s[0] * 2
    assert layer.step_size_output.item() == s[2]
    layer.update_para(use_FP=True)
    with torch.no_grad():
        y = layer(x)
    ref = torch.nn.functional.linear(x, layer.weight, layer.bias)
    assert relerr(y, ref) < 6e-3
    assert sorted(layer.state_dict()) == ["bias", "step_size_input", "step_size_output", "step_size_weight", "weight"]


# --------------------------------------------------------------------------------------------------- ADDA layer
@pytest.mark.parametrize("name", ["adda_linear_288", "adda_linear_1152", "adda_linear_noise"])
def test_adda_linear_vs_oracle(name):
    dev = _dev()
    from oracle import cim_torch as OC
    from cim_layers.layers_qn_lsq_adda_cim import Linear_lsq_adda_cim, gen_weight_split_dict
    f = load_golden("cim_layers")[name]
    cfg = f["cfg"]
    w = f["init"]["weight"]
    layer = Linear_lsq_adda_cim(w.shape[1], w.shape[0], weight_bit=cfg["weight_bit"], input_bit=cfg["input_bit"],
                                output_bit=cfg["output_bit"], noise_scale=cfg["noise_scale"], dac_bit=cfg["dac_bit"],
                                adc_bit=cfg["adc_bit"], adc_gain_1_scale=cfg["adc_gain_1_scale"],
                                adc_gain_range=cfg["adc_gain_range"], adc_adjust_mode=cfg["adc_adjust_mode"],
                                gain_noise_scale=cfg.get("gain_noise_scale", 0),
                                offset_noise_scale=cfg.get("offset_noise_scale", 0), bias=True)
    layer.load_state_dict(f["init"])
    layer = layer.to(dev).eval()
    layer.weight_noise_unit = f["unit"].float().to(dev)
    layer.weight_mapping_info = f["mapping"]
    block = [max(i["row_num"] for i in f["mapping"].values()), max(i["col_num"] for i in f["mapping"].values())]
    assert gen_weight_split_dict(layer, block) == f["mapping"]           # same block list as the reference mapper
    noise = {}
    if cfg.get("gain_noise_scale"):
        layer.gain_noise, layer.offset_noise = f["gain_noise"].to(dev), f["offset_noise"].to(dev)
        noise = dict(gain_noise=f["gain_noise"], offset_noise=f["offset_noise"])
    x, x2 = _bfr(f["x"]), _bfr(f["x2"])
    p = {k: v.clone() for k, v in f["init"].items()}
    with torch.no_grad():
        ref1 = OC.adda_linear(p, cfg, x, f["mapping"], f["unit"].float(), **noise)
        out1 = layer(x.to(dev))
        assert layer.step_size_input.item() == p["step_size_input"].item()
        assert layer.step_size_weight.item() == p["step_size_weight"].item()
        ref2 = OC.adda_linear(p, cfg, x2, f["mapping"], f["unit"].float(), **noise)
        out2 = layer(x2.to(dev))
    torch.cuda.synchronize()
    exact = cfg["noise_scale"] == 0
    if exact:
        # digits and weight codes are integers: block products, ADC codes and therefore everything downstream agree
        assert abs(layer.adc_gain.item() - p["adc_gain"].item()) <= 1e-7 * p["adc_gain"].item()
        assert layer.step_size_output.item() == pytest.approx(p["step_size_output"].item(), rel=1e-6)
    s_out = p["step_size_output"].item()
    for got, ref in ((out1, ref1), (out2, ref2)):
        dcode = (torch.round(got.cpu() / s_out) - torch.round(ref / s_out)).abs()
        same = (dcode < 0.5).float().mean().item()
        print(f"{name}: output codes identical {100 * same:.2f}% max |dcode| {dcode.max().item():.2f}")
        if exact:
            assert dcode.max().item() == 0
        else:  # non-integer noisy weight codes are rounded to bf16 (as under the reference's autocast): ADC codes flip
            # by one on some boundaries and the upper DAC slice carries weight 16
            assert dcode.max().item() <= 3 and same > 0.7
    assert relerr(out1.cpu(), f["y1"]) < 5e-2                            # vs the reference on the un-rounded inputs
    if not exact:
        # precise parity mode: noisy weight codes as (hi, lo) bf16 operand pairs, so the ADC sees 16 significand bits
        from sdb200 import engine as E
        E.set_parity_mode("precise")
        try:
            with torch.no_grad():
                out3 = layer(x2.to(dev))
        finally:
            E.set_parity_mode("fast")
        dcode = (torch.round(out3.cpu() / s_out) - torch.round(ref2 / s_out)).abs()
        same = (dcode < 0.5).float().mean().item()
        print(f"{name}: precise mode: output codes identical {100 * same:.3f}% max |dcode| {dcode.max().item():.2f}")
        # (a single ADC code of the upper DAC slice still flips where the fp32 partial sum sits within summation-order
        # noise of a rounding boundary; it carries weight 16, hence up to 3 output steps on those few elements)
        assert dcode.max().item() <= 3 and same > 0.999


# --------------------------------------------------------------------------------------------------- inside the DiT
def _convert(model, lin_cls, conv_cls, **kw):
    """What cim_qn_train/progressive_qn_train.py:576-651 does: every leaf of EXACT type nn.Linear / nn.Conv2d is
    replaced by the quantised class sharing its weight / bias Parameters."""
    import torch.nn as nn
    for parent in list(model.modules()):
        for name, m in list(parent.named_children()):
            new = None
            if type(m) is nn.Linear:
                new = lin_cls(m.in_features, m.out_features, bias=m.bias is not None, **kw)
            elif type(m) is nn.Conv2d:
                new = conv_cls(m.in_channels, m.out_channels, m.kernel_size, m.stride, m.padding, m.groups,
                               bias=m.bias is not None, **kw)
            if new is not None:
                new.weight = m.weight
                if m.bias is not None:
                    new.bias = m.bias
                setattr(parent, name, new.to(m.weight.device))
    return model


def test_dit_with_lsq_leaves_vs_oracle():
    """The drop-in DiT after the reference's module surgery (all 25 Linear + the mask Conv2d of the small config
    become LSQ layers, 8-bit activations / 4-bit weights): forward against the oracle DiT with the oracle LSQ layer in
    every leaf, lazily initialised step sizes copied from the oracle's warm-up (SURVEY §8c trap iii); then one captured
    training step must produce finite gradients for every parameter including the step sizes."""
    dev = _dev()
    from oracle import cim_torch as OC
    from oracle import dit_torch as OD
    from cim_layers.layers_qn_lsq import Conv2d_qn_lsq, Linear_qn_lsq
    from models.transformer import DIT
    from util import cond_of, l2err, randomise_zero_init
    fx = load_golden("dit_small_cond")
    cfg, inp = fx["config"], fx["inputs"]
    torch.manual_seed(fx["seed"])
    model = DIT(4, cfg)
    randomise_zero_init(model, fx["rand_seed"])
    qcfg = dict(weight_bit=4, input_bit=8, output_bit=8, noise_scale=0.0)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    states = {}

    def leaf(name, x, w, b, conv):
        p = states.setdefault(name, {"step_size_input": torch.tensor(1.0), "step_size_output": torch.tensor(1.0),
                                     "step_size_weight": torch.tensor(1.0)})
        p["weight"] = w
        if b is not None:
            p["bias"] = b
        return OC.lsq_layer(p, qcfg, x, None, conv)

    cond = cond_of(inp, cfg)
    OD.LEAF = leaf
    try:
        with torch.no_grad():
            OD.dit_forward(sd, cfg, inp["x"], inp["t"], cond)                    # warm-up: initialises every step size
            ref = OD.dit_forward(sd, cfg, inp["x"], inp["t"], cond)
            with torch.autocast("cpu", dtype=torch.bfloat16):
                ref16 = OD.dit_forward(sd, cfg, inp["x"], inp["t"], cond).float()
    finally:
        OD.LEAF = None
    model = _convert(model, Linear_qn_lsq, Conv2d_qn_lsq, **qcfg).to(dev).eval()
    n_q = sum(1 for m in model.modules() if isinstance(m, (Linear_qn_lsq, Conv2d_qn_lsq)))
    # mask conv + patch embed + t_proj (2) + 2 layers x (qkv, out, mlp x2, q/k/v/out, context_proj, adaLN) + adaLN + proj_out
    assert n_q == len(states) == 1 + 1 + 2 + 2 * 10 + 2                         # every leaf went through the hook
    with torch.no_grad():
        for name, m in model.named_modules():
            if name in states:
                for k in ("step_size_input", "step_size_output", "step_size_weight"):
                    getattr(m, k).copy_(states[name][k])
        condd = cond_of(inp, cfg, dev)
        out = model(inp["x"].to(dev), inp["t"].to(dev), condd)
        out_again = model(inp["x"].to(dev), inp["t"].to(dev), condd)             # captured-graph replay
    torch.cuda.synchronize()
    err, yard = l2err(out.cpu(), ref), l2err(ref16, ref)
    print(f"DiT with LSQ leaves: ours-vs-oracle rel-L2 {err:.3e}; oracle bf16-autocast-vs-fp32 {yard:.3e}")
    assert torch.equal(out, out_again)
    assert err < max(5e-2, 1.5 * yard)
    # one training step through the quantised network
    model.train()
    x, t = inp["x"].to(dev), inp["t"].to(dev)
    tgt = torch.randn(x.shape, generator=torch.Generator().manual_seed(1)).to(dev)
    for _ in range(2):                                                            # second pass replays the graphs
        model.zero_grad()
        loss = torch.nn.functional.mse_loss(model(x, t, condd), tgt)
        loss.backward()
    torch.cuda.synchronize()
    missing = [k for k, p in model.named_parameters() if p.grad is None or not torch.isfinite(p.grad).all()]
    assert not missing, missing[:6]
    steps = [p.grad.abs().item() for k, p in model.named_parameters() if "step_size" in k]
    assert sum(1 for s in steps if s > 0) > 0.9 * len(steps)


@pytest.mark.parametrize("name", ["adda_linear_288", "adda_linear_1152", "adda_linear_noise"])
def test_adda_linear_backward_vs_oracle(name):
    """Gradients of the ADDA layer (input, weight, bias, the three LSQ step sizes and the ADC gain) against autograd
    through the oracle (itself pinned to the reference's gradients exactly, tests/test_oracle.py)."""
    dev = _dev()
    from oracle import cim_torch as OC
    from cim_layers.layers_qn_lsq_adda_cim import Linear_lsq_adda_cim
    f = load_golden("cim_layers")[name]
    cfg = f["cfg"]
    w = f["init"]["weight"]
    layer = Linear_lsq_adda_cim(w.shape[1], w.shape[0], weight_bit=cfg["weight_bit"], input_bit=cfg["input_bit"],
                                output_bit=cfg["output_bit"], noise_scale=cfg["noise_scale"], dac_bit=cfg["dac_bit"],
                                adc_bit=cfg["adc_bit"], adc_gain_1_scale=cfg["adc_gain_1_scale"],
                                adc_gain_range=cfg["adc_gain_range"], adc_adjust_mode=cfg["adc_adjust_mode"],
                                gain_noise_scale=cfg.get("gain_noise_scale", 0),
                                offset_noise_scale=cfg.get("offset_noise_scale", 0), bias=True)
    # start from the state AFTER the reference's first forward (step sizes and ADC gain initialised)
    state = dict(f["init"])
    state.update(f["after"])
    layer.load_state_dict(state)
    layer = layer.to(dev).train()
    layer.__dict__["_gain_init_pending"] = False
    layer.weight_noise_unit = f["unit"].float().to(dev)
    layer.weight_mapping_info = f["mapping"]
    noise = {}
    if cfg.get("gain_noise_scale"):
        layer.gain_noise, layer.offset_noise = f["gain_noise"].to(dev), f["offset_noise"].to(dev)
        noise = dict(gain_noise=f["gain_noise"], offset_noise=f["offset_noise"])
    x2 = _bfr(f["x2"])
    tgt = f["target"]
    p = {k: v.clone().requires_grad_(True) for k, v in state.items()}
    xo = x2.clone().requires_grad_(True)
    ref = OC.adda_linear(p, cfg, xo, f["mapping"], f["unit"].float(), **noise)
    ((ref - tgt) ** 2).mean().backward()
    p16 = {k: v.clone().requires_grad_(True) for k, v in state.items()}
    with torch.autocast("cpu", dtype=torch.bfloat16):
        ref16 = OC.adda_linear(p16, cfg, x2, f["mapping"], f["unit"].float(), **noise)
    ((ref16.float() - tgt) ** 2).mean().backward()
    xd = x2.to(dev).requires_grad_(True)
    out = layer(xd)
    ((out - tgt.to(dev)) ** 2).mean().backward()
    torch.cuda.synchronize()
    got = {k: v.grad.cpu() for k, v in layer.named_parameters()}
    got["x"] = xd.grad.cpu()
    want = {k: v.grad for k, v in p.items()}
    want["x"] = xo.grad
    for k in want:
        err = ((got[k] - want[k]).norm() / want[k].norm().clamp_min(1e-20)).item()
        yard = ((p16[k].grad - want[k]).norm() / want[k].norm().clamp_min(1e-20)).item() if k in p16 else 0.0
        print(f"   {name} grad {k}: rel-L2 {err:.3e} (oracle under bf16 autocast: {yard:.3e})")
        # scalar gradients (step sizes, ADC gain) are ill-conditioned sums; everything is gated on the reference's own
        # bf16-autocast error like the LSQ layers
        assert err < max(8e-2 if want[k].dim() == 0 else 4e-2, 1.5 * yard), k


def test_dit_with_adda_leaves_vs_oracle_and_trains():
    """configs[4] structure: every Linear and the mask conv of the DiT as ADDA compute-in-memory layers (8-bit inputs in
    two 4-bit DAC slices, 4-bit weights, 8-bit ADC, array height 576) — forward against the oracle DiT with the oracle
    ADDA layer in every leaf (lazily initialised parameters copied from the oracle's warm-up), then a captured training
    step with finite gradients for every parameter including step sizes and ADC gains."""
    dev = _dev()
    from oracle import cim_torch as OC
    from oracle import dit_torch as OD
    from cim_layers.layers_qn_lsq_adda_cim import Conv2d_lsq_adda_cim, Linear_lsq_adda_cim, gen_weight_split_dict
    from models.transformer import DIT
    from util import cond_of, l2err, randomise_zero_init
    fx = load_golden("dit_small_cond")
    cfg, inp = fx["config"], fx["inputs"]
    torch.manual_seed(fx["seed"])
    model = DIT(4, cfg)
    randomise_zero_init(model, fx["rand_seed"])
    acfg = dict(weight_bit=4, input_bit=8, output_bit=8, noise_scale=0.0, dac_bit=5, adc_bit=8,
                adc_gain_1_scale=9.071428571, adc_gain_range=[1 / 64, 1 / 64], adc_adjust_mode='current')
    ocfg = dict(acfg, adc_gain_min=1 / 64, adc_gain_max=1 / 64)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    states, maps = {}, {}

    class _W:   # the mapper only looks at module.weight
        def __init__(self, w):
            self.weight = w

    def leaf(name, x, w, b, conv):
        p = states.setdefault(name, {"step_size_input": torch.tensor(1.0), "step_size_output": torch.tensor(1.0),
                                     "step_size_weight": torch.tensor(1.0), "adc_gain": torch.tensor(1 / 64)})
        p["weight"] = w.reshape(w.shape[0], -1)
        if b is not None:
            p["bias"] = b
        maps.setdefault(name, gen_weight_split_dict(_W(p["weight"]), [576, 2048]))
        if conv is not None:   # 1x1 convolution = the linear layer applied per pixel (Unfold order (c, 1, 1))
            y = OC.adda_linear(p, ocfg, x.permute(0, 2, 3, 1).contiguous(), maps[name], None)
            return y.permute(0, 3, 1, 2).contiguous()
        return OC.adda_linear(p, ocfg, x, maps[name], None)

    cond = cond_of(inp, cfg)
    OD.LEAF = leaf
    try:
        with torch.no_grad():
            OD.dit_forward(sd, cfg, inp["x"], inp["t"], cond)
            ref = OD.dit_forward(sd, cfg, inp["x"], inp["t"], cond)
            with torch.autocast("cpu", dtype=torch.bfloat16):
                ref16 = OD.dit_forward(sd, cfg, inp["x"], inp["t"], cond).float()
    finally:
        OD.LEAF = None
    model = _convert(model, Linear_lsq_adda_cim, Conv2d_lsq_adda_cim, **acfg).to(dev).eval()
    with torch.no_grad():
        for name, m in model.named_modules():
            if name in states:
                for k in ("step_size_input", "step_size_output", "step_size_weight", "adc_gain"):
                    getattr(m, k).copy_(states[name][k])
                m.__dict__["_gain_init_pending"] = False
                m.weight_mapping_info = gen_weight_split_dict(m, [576, 2048])
        condd = cond_of(inp, cfg, dev)
        out = model(inp["x"].to(dev), inp["t"].to(dev), condd)
        out_again = model(inp["x"].to(dev), inp["t"].to(dev), condd)
    torch.cuda.synchronize()
    err, yard = l2err(out.cpu(), ref), l2err(ref16, ref)
    print(f"DiT with ADDA leaves: ours-vs-oracle rel-L2 {err:.3e}; oracle bf16-autocast-vs-fp32 {yard:.3e}")
    assert torch.equal(out, out_again)
    assert err < max(5e-2, 1.5 * yard)
    model.train()
    x, t = inp["x"].to(dev), inp["t"].to(dev)
    tgt = torch.randn(x.shape, generator=torch.Generator().manual_seed(1)).to(dev)
    for _ in range(2):
        model.zero_grad()
        loss = torch.nn.functional.mse_loss(model(x, t, condd), tgt)
        loss.backward()
    torch.cuda.synchronize()
    missing = [(k, "none" if p.grad is None else "nonfinite") for k, p in model.named_parameters()
               if p.grad is None or not torch.isfinite(p.grad).all()]
    assert not missing, (len(missing), missing[:40])
    gains = [p.grad.abs().item() for k, p in model.named_parameters() if k.endswith("adc_gain")]
    assert len(gains) == 26


# ------------------------------------------------------------------------------------- k x k ADDA convolutions
def _adda_conv(f, dev, train=False, after=False):
    from cim_layers.layers_qn_lsq_adda_cim import Conv2d_lsq_adda_cim, gen_weight_split_dict
    cfg, cv = f["cfg"], f["conv"]
    cin, cout = f["shape"]
    layer = Conv2d_lsq_adda_cim(cin, cout, cv["kernel_size"], cv["stride"], cv["padding"], 1, weight_bit=cfg["weight_bit"],
                                input_bit=cfg["input_bit"], output_bit=cfg["output_bit"], noise_scale=cfg["noise_scale"],
                                dac_bit=cfg["dac_bit"], adc_bit=cfg["adc_bit"], adc_gain_1_scale=cfg["adc_gain_1_scale"],
                                adc_gain_range=cfg["adc_gain_range"], adc_adjust_mode=cfg["adc_adjust_mode"], bias=True)
    state = dict(f["init"])
    if after:
        state.update(f["after"])
    layer.load_state_dict(state)
    layer = layer.to(dev)
    layer = layer.train() if train else layer.eval()
    if after:
        layer.__dict__["_gain_init_pending"] = False
    layer.weight_noise_unit = f["unit"].float().to(dev)
    layer.weight_mapping_info = f["mapping"]
    block = [max(i["row_num"] for i in f["mapping"].values()), max(i["col_num"] for i in f["mapping"].values())]
    assert gen_weight_split_dict(layer, block) == f["mapping"]
    return layer, state


@pytest.mark.parametrize("name", ["adda_conv3x3", "adda_conv3x3_clean", "adda_conv4x4s2"])
def test_adda_conv_vs_oracle(name):
    """Conv2d_lsq_adda_cim with k x k kernels (3x3 same, 4x4 stride 2; K split into row blocks in nn.Unfold's (c, ky, kx)
    order): output codes against the oracle (== the reference, tests/test_oracle.py), bit-exact without weight noise."""
    dev = _dev()
    from oracle import cim_torch as OC
    f = load_golden("cim_adda_conv")[name]
    cfg = f["cfg"]
    layer, _ = _adda_conv(f, dev)
    x, x2 = _bfr(f["x"]), _bfr(f["x2"])
    p = {k: v.clone() for k, v in f["init"].items()}
    with torch.no_grad():
        ref1 = OC.adda_linear(p, cfg, x, f["mapping"], f["unit"].float(), conv=f["conv"])
        out1 = layer(x.to(dev))
        assert layer.step_size_input.item() == p["step_size_input"].item()
        assert layer.step_size_weight.item() == p["step_size_weight"].item()
        ref2 = OC.adda_linear(p, cfg, x2, f["mapping"], f["unit"].float(), conv=f["conv"])
        out2 = layer(x2.to(dev))
    torch.cuda.synchronize()
    assert out1.shape == ref1.shape
    exact = cfg["noise_scale"] == 0
    if exact:
        assert abs(layer.adc_gain.item() - p["adc_gain"].item()) <= 1e-7 * p["adc_gain"].item()
        assert layer.step_size_output.item() == pytest.approx(p["step_size_output"].item(), rel=1e-6)
    s_out = p["step_size_output"].item()
    for got, ref in ((out1, ref1), (out2, ref2)):
        dcode = (torch.round(got.cpu() / s_out) - torch.round(ref / s_out)).abs()
        same = (dcode < 0.5).float().mean().item()
        print(f"{name}: output codes identical {100 * same:.2f}% max |dcode| {dcode.max().item():.2f}")
        if exact:
            assert dcode.max().item() == 0
        else:
            assert dcode.max().item() <= 3 and same > 0.7
    assert relerr(out1.cpu(), f["y1"]) < 5e-2


@pytest.mark.parametrize("name", ["adda_conv3x3", "adda_conv4x4s2"])
def test_adda_conv_backward_vs_oracle(name):
    dev = _dev()
    from oracle import cim_torch as OC
    f = load_golden("cim_adda_conv")[name]
    cfg = f["cfg"]
    layer, state = _adda_conv(f, dev, train=True, after=True)
    x2, tgt = _bfr(f["x2"]), f["target"]
    p = {k: v.clone().requires_grad_(True) for k, v in state.items()}
    xo = x2.clone().requires_grad_(True)
    ref = OC.adda_linear(p, cfg, xo, f["mapping"], f["unit"].float(), conv=f["conv"])
    ((ref - tgt) ** 2).mean().backward()
    p16 = {k: v.clone().requires_grad_(True) for k, v in state.items()}
    with torch.autocast("cpu", dtype=torch.bfloat16):
        ref16 = OC.adda_linear(p16, cfg, x2, f["mapping"], f["unit"].float(), conv=f["conv"])
    ((ref16.float() - tgt) ** 2).mean().backward()
    xd = x2.to(dev).requires_grad_(True)
    out = layer(xd)
    ((out - tgt.to(dev)) ** 2).mean().backward()
    torch.cuda.synchronize()
    got = {k: v.grad.cpu() for k, v in layer.named_parameters()}
    got["x"] = xd.grad.cpu()
    want = {k: v.grad for k, v in p.items()}
    want["x"] = xo.grad
    for k in want:
        err = ((got[k] - want[k]).norm() / want[k].norm().clamp_min(1e-20)).item()
        yard = ((p16[k].grad - want[k]).norm() / want[k].norm().clamp_min(1e-20)).item() if k in p16 else 0.0
        print(f"   {name} grad {k}: rel-L2 {err:.3e} (oracle under bf16 autocast: {yard:.3e})")
        assert err < max(8e-2 if want[k].dim() == 0 else 4e-2, 1.5 * yard), k


# --------------------------------------------------------------------------------------- inside the ReLU UNet
@pytest.mark.parametrize("kind", ["lsq", "adda"])
def test_relu_unet_with_cim_leaves_runs_and_trains(kind):
    """The ReLU / CustomMultiheadAttention UNet variant after the reference's module surgery (Model_Unet_train.py: every
    nn.Conv2d — 3x3, 1x1 shortcut, 4x4 stride-2 down-sample, the 18 -> 3 mask conv — and every nn.Linear — time
    embedding, q/k/v/out projections, context projections — becomes a cim_layers leaf; ConvTranspose2d stays float).
    With 8-bit weights, no weight noise the quantised network tracks the float one; a training step gives every
    parameter, step size and ADC gain a finite gradient."""
    dev = _dev()
    import torch.nn as nn
    from cim_layers.layers_qn_lsq import Conv2d_qn_lsq, Linear_qn_lsq
    from cim_layers.layers_qn_lsq_adda_cim import Conv2d_lsq_adda_cim, Linear_lsq_adda_cim, gen_weight_split_dict
    from models.unet_cond_base_relu import Unet
    from util import cond_of, l2err
    fx = load_golden("unet_small_cond_relu")
    cfg, inp = fx["config"], fx["inputs"]
    torch.manual_seed(fx["seed"])
    model = Unet(4, cfg).eval().to(dev)
    x, t, cond = inp["x"].to(dev), inp["t"].to(dev), cond_of(inp, cfg, dev)
    with torch.no_grad():
        plain = model(x, t, cond)
    n_leaves = sum(1 for m in model.modules() if type(m) in (nn.Conv2d, nn.Linear))
    kw = dict(weight_bit=8, input_bit=8, output_bit=8, noise_scale=0.0)
    if kind == "lsq":
        _convert(model, Linear_qn_lsq, Conv2d_qn_lsq, **kw)
        leaf_types = (Linear_qn_lsq, Conv2d_qn_lsq)
    else:
        kw.update(dac_bit=5, adc_bit=12, adc_gain_1_scale=9.071428571, adc_gain_range=[1 / 64, 1 / 64],
                  adc_adjust_mode="current")
        _convert(model, Linear_lsq_adda_cim, Conv2d_lsq_adda_cim, **kw)
        leaf_types = (Linear_lsq_adda_cim, Conv2d_lsq_adda_cim)
        for m in model.modules():
            if type(m) in leaf_types:
                m.weight_mapping_info = gen_weight_split_dict(m, [576, 2048])
    leaves = [m for m in model.modules() if type(m) in leaf_types]
    assert len(leaves) == n_leaves and n_leaves > 40
    assert any(type(m) is Conv2d_qn_lsq or type(m) is Conv2d_lsq_adda_cim for m in leaves if getattr(m, "kernel_size", None) == (3, 3))
    model.invalidate_plan()
    with torch.no_grad():
        q1 = model(x, t, cond)      # first forward initialises the step sizes (and ADC gains)
        q2 = model(x, t, cond)
    err = l2err(q2.cpu(), plain.cpu())
    print(f"ReLU UNet with {kind} leaves ({n_leaves} of them): rel-L2 vs the float network {err:.3e}")
    assert torch.isfinite(q1).all() and torch.isfinite(q2).all() and q2.shape == plain.shape
    assert err < (0.25 if kind == "lsq" else 0.5)
    model.train()
    out = model(x, t, cond)
    out.square().mean().backward()
    torch.cuda.synchronize()
    missing = [k for k, p in model.named_parameters() if p.requires_grad and (p.grad is None or not torch.isfinite(p.grad).all())]
    assert not missing, missing[:6]


@pytest.mark.parametrize("kind", ["lsq", "adda"])
def test_captured_step_packs_the_quantised_weights_in_one_launch(kind):
    """Inside a captured training step the quantised weights of all cim_layers leaves live in the weight bank's buffer and
    their bf16 GEMM forms (forward, transposed, convolution, flipped-tap dgrad, mapped array blocks) come from ONE
    table-driven launch after the weight side (bank.refresh_derived) instead of one or two launches per leaf and
    direction. Same conversions, so output and every gradient must be bit-identical to the per-leaf route."""
    dev = _dev()
    from cim_layers.layers_qn_lsq import Conv2d_qn_lsq, Linear_qn_lsq
    from cim_layers.layers_qn_lsq_adda_cim import Conv2d_lsq_adda_cim, Linear_lsq_adda_cim, gen_weight_split_dict
    from models.unet_cond_base_relu import Unet
    from sdb200 import quant as Q
    from util import cond_of
    fx = load_golden("unet_small_cond_relu")
    cfg, inp = fx["config"], fx["inputs"]
    torch.manual_seed(fx["seed"])
    model = Unet(4, cfg).eval().to(dev)
    x, t, cond = inp["x"].to(dev), inp["t"].to(dev), cond_of(inp, cfg, dev)
    kw = dict(weight_bit=4, input_bit=8, output_bit=8, noise_scale=0.05)
    if kind == "lsq":
        _convert(model, Linear_qn_lsq, Conv2d_qn_lsq, **kw)
    else:
        kw.update(dac_bit=5, adc_bit=12, adc_gain_1_scale=9.071428571, adc_gain_range=[1 / 64, 1 / 64],
                  adc_adjust_mode="current")
        _convert(model, Linear_lsq_adda_cim, Conv2d_lsq_adda_cim, **kw)
        for m in model.modules():
            if type(m) in (Linear_lsq_adda_cim, Conv2d_lsq_adda_cim):
                m.weight_mapping_info = gen_weight_split_dict(m, [144, 64])   # several row and column blocks per leaf
    g = torch.Generator().manual_seed(21)
    for m in model.modules():
        if hasattr(m, "_wq_spec"):
            m.weight_noise_unit = torch.randn(m.weight.shape, generator=g).to(dev)   # the same noise in both runs
    model.invalidate_plan()
    with torch.no_grad():
        model(x, t, cond)           # lazy initialisation of the step sizes / ADC gains
    model.train()

    def step(banked):
        Q.BANK_DERIVED = banked
        try:
            model.invalidate_plan()
            model.zero_grad(set_to_none=True)
            out = model(x, t, cond)
            out.square().mean().backward()
            torch.cuda.synchronize()
            tg = model.__dict__.get("_train_graph")
            assert tg is not None, "the training step was not captured"
            return out.detach().clone(), {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}, \
                len(tg.bank.djobs), tg.launches
        finally:
            Q.BANK_DERIVED = True

    out_a, ga, jobs_a, launches_a = step(True)
    out_b, gb, jobs_b, launches_b = step(False)
    print(f"   {kind}: {jobs_a} packed forms of quantised weights from one launch; captured launches {launches_b} -> {launches_a}")
    # (tg.launches counts the library's own launches: the per-leaf route packs with torch copy kernels, which it does not
    #  see, so the banked route shows up as exactly one launch more — the table-driven pack)
    assert jobs_a > 40 and jobs_b == 0 and launches_a == launches_b + 1
    assert torch.equal(out_a, out_b)
    assert ga.keys() == gb.keys() and len(ga) > 100
    bad = [k for k in ga if not torch.equal(ga[k], gb[k])]
    assert not bad, bad[:6]


# --------------------------------------------------------------------------------- fused weight-side kernels
@pytest.mark.parametrize("isint", [False, True])
@pytest.mark.parametrize("noise", [0.0, 0.08])
@pytest.mark.parametrize("shape,bits", [((192, 288), 4), ((64, 32, 3, 3), 8), ((3, 18, 1, 1), 4)])
def test_fused_weight_side_matches_the_torch_formulation(isint, noise, shape, bits):
    """csrc/wquant.cu (lazy step init + weight_quant_lsq + add_noise in two launches, backward in two) against the
    op-by-op torch formulation of the same reference formulas (sdb200/quant.py, which the oracle tests pin to the
    reference): quantised weights bit-identical, lazily initialised step identical, gradients of the weight and of the
    step size (through the clamp mask, the straight-through round and the noise range's max / min with tied codes)."""
    dev = _dev()
    from sdb200 import quant as Q
    g = torch.Generator().manual_seed(sum(shape) + bits)
    w0 = (torch.randn(shape, generator=g) * 0.05).to(dev)
    unit = torch.randn(shape, generator=g).to(dev) if noise else None
    gout = torch.randn(shape, generator=g).to(dev)

    def torch_path():
        w = w0.clone().requires_grad_(True)
        step = torch.ones((), device=dev, requires_grad=True)
        Q.init_weight_step_(step, w, bits)
        wq, _ = Q.quantize_weight(w, step, bits, isint)
        out = Q.add_weight_noise(wq, noise, unit)
        (out * gout).sum().backward()
        return out.detach(), step.detach().clone(), w.grad, step.grad

    def fused_path():
        w = w0.clone().requires_grad_(True)
        step = torch.ones((), device=dev, requires_grad=True)
        owner = torch.nn.Module()
        out, _ = Q.weight_qn(owner, w, step, bits, isint, noise, unit)
        (out * gout).sum().backward()
        out2, _ = Q.weight_qn(owner, w, step, bits, isint, noise, unit)     # second call: step no longer 1, scratch reused
        return out.detach(), step.detach().clone(), w.grad, step.grad, out2.detach()

    ro, rs, rdw, rds = torch_path()
    fo, fs, fdw, fds, fo2 = fused_path()
    torch.cuda.synchronize()
    assert fs.item() == rs.item() and rs.item() != 1.0
    assert torch.equal(fo, ro) and torch.equal(fo2, ro)
    assert relerr(fdw, rdw) < 1e-5
    assert abs(fds.item() - rds.item()) <= 2e-4 * max(abs(rds.item()), 1e-6), (fds.item(), rds.item())


@pytest.mark.parametrize("rows,c,bias,quant", [(1024, 768, True, True), (77, 24, False, True), (300, 64, True, False)])
def test_adda_output_side_matches_the_torch_formulation(rows, c, bias, quant):
    """b200sd_adda_out / _bwd (one launch each way) against the chain of torch ops they replace — the reference's
    `x / w_scale / in_scale / adc_scale + bias` and data_quant_lsq (layers_qn_lsq_adda_cim.py:482-500): forward
    bit-exact (same IEEE divisions in the same order), gradients to fp32 rounding."""
    dev = _dev()
    from sdb200 import quant as Q
    g = torch.Generator().manual_seed(rows + c)
    acc = torch.round(torch.randn(rows, c, generator=g) * 900).to(dev)
    step_w = torch.tensor(0.0123, device=dev)
    step_in = torch.tensor(0.0456, device=dev)
    b = (torch.randn(c, generator=g) * 0.1).to(dev) if bias else None
    gout = (torch.randn(rows, c, generator=g) * 0.01).to(torch.bfloat16).to(dev)
    bits = 8
    qp = float(Q.qp_of(bits))

    def leaves():
        a = acc.clone().requires_grad_(True)
        al = torch.tensor(0.1417, device=dev, requires_grad=True)
        so = torch.tensor(0.0211, device=dev, requires_grad=True) if quant else None
        bb = b.clone().requires_grad_(True) if bias else None
        return a, al, so, bb

    a, al, so, bb = leaves()
    y = a / (1.0 / step_w).detach() / (1.0 / step_in).detach() / al
    if bias:
        y = y + bb.view(1, -1)
    if quant:
        s = Q._grad_scale(so, 1.0 / (qp * rows * c) ** 0.5)
        ref = Q._round_ste(torch.clamp(y / s, -qp, qp)) * s
    else:
        ref = y
    ref.backward(gout.float())

    class Owner:
        pass
    a2, al2, so2, bb2 = leaves()
    out = Q.adda_out(Owner(), a2, al2, bb2, so2, step_w, step_in, bits, rows * c)
    assert out.dtype == torch.bfloat16
    assert torch.equal(out, ref.detach().to(torch.bfloat16))
    out.backward(gout)
    assert relerr(a2.grad, a.grad) < 1e-5
    assert relerr(al2.grad, al.grad) < 1e-4
    if bias:
        assert relerr(bb2.grad, bb.grad) < 1e-4
    if quant:
        assert relerr(so2.grad, so.grad) < 1e-4
    # a second call in a row reuses the layer's scratch (the last CTA re-arms the arrival counter)
    a3, al3, so3, bb3 = leaves()
    owner = Owner()
    for _ in range(2):
        for t in (a3, al3, so3, bb3):
            if t is not None:
                t.grad = None
        Q.adda_out(owner, a3, al3, bb3, so3, step_w, step_in, bits, rows * c).backward(gout)
    assert torch.equal(a3.grad, a2.grad) and torch.equal(al3.grad, al2.grad)


@pytest.mark.parametrize("mode", ["gain", "current"])
def test_adc_scale_kernel_matches_the_torch_formulation(mode):
    """get_adc_scale (layers_utils_adda.py:40-49) as one launch: value bit-exact, gradient through the straight-through
    clamp / round as autograd gives it."""
    dev = _dev()
    from sdb200 import quant as Q
    cases = {"gain": ((1 / 64, 1 / 64, 1 / 64), (3.3, 1.0, 8.0), (0.013, 1 / 128, 1 / 16), (20.0, 1.0, 8.0)),
             # 'current' rounds the range 1 / gain: gains above 2 would round it to zero (the reference divides by it)
             "current": ((1 / 64, 1 / 64, 1 / 64), (0.013, 1 / 128, 1 / 16), (0.3, 1 / 128, 0.4), (0.001, 1 / 128, 0.4))}
    for g0, lo, hi in cases[mode]:
        scale1 = 9.071428571
        gain = torch.tensor(g0, device=dev, requires_grad=True)
        gc = torch.clamp(gain, lo, hi)
        g = (gc - gain).detach() + gain
        if mode == "gain":
            g = (torch.round(g) - g).detach() + g
        else:
            r = 1 / g
            g = 1 / ((torch.round(r) - r).detach() + r)
        ref = g * scale1
        ref.backward()
        gain2 = torch.tensor(g0, device=dev, requires_grad=True)
        got = Q.AdcScaleFn.apply(gain2, lo, hi, mode == "gain", scale1)
        assert got.item() == ref.item()
        got.backward()
        assert abs(gain2.grad.item() - gain.grad.item()) <= 1e-5 * abs(gain.grad.item())


def test_model_wide_weight_side_equals_the_per_layer_launches():
    """sdb200.quant.prepare_weight_side (b200sd_weight_qn_multi: one launch per phase for every quantised leaf of a
    model) against the per-layer launches: quantised weights, lazily initialised step sizes and all gradients
    bit-identical (same kernel bodies, same per-layer CTA split); a bypassed leaf (use_FP) and a leaf the pass never runs
    are handled; more than 64 leaves exercise the table chunking."""
    dev = _dev()
    import torch.nn as nn
    from cim_layers.layers_qn_lsq import Linear_qn_lsq
    from cim_layers.layers_qn_lsq_adda_cim import Linear_lsq_adda_cim
    from sdb200 import quant as Q
    kw = dict(weight_bit=4, input_bit=8, output_bit=8, noise_scale=0.08)
    akw = dict(kw, dac_bit=5, adc_bit=8, adc_gain_1_scale=9.071428571, adc_gain_range=[1 / 64, 1 / 64],
               adc_adjust_mode="current")

    def build():
        torch.manual_seed(5)
        leaves = [Linear_qn_lsq(24 + 8 * (i % 5), 16 + 8 * (i % 3), **kw) for i in range(66)]
        leaves += [Linear_lsq_adda_cim(32, 24, **akw), Linear_qn_lsq(16, 16, **dict(kw, noise_scale=0.0))]
        m = nn.Sequential(*leaves).to(dev)
        g = torch.Generator().manual_seed(9)
        for lf in m:
            lf.weight_noise_unit = torch.randn(lf.weight.shape, generator=g).to(dev)
        m[3].use_FP = True
        return m

    def run(model, banked):
        Q.FUSED_WEIGHT_BANK = banked
        try:
            served = Q.prepare_weight_side(model)
            outs = []
            for i, lf in enumerate(model):
                if i == 7:
                    continue                                   # a leaf this pass does not run
                if lf.use_FP:
                    continue
                outs.append(lf._w_qn(isint=lf._WQ_ISINT)[0])
            g = torch.Generator().manual_seed(11)
            loss = sum((o * torch.randn(o.shape, generator=g).to(dev)).sum() for o in outs)
            loss.backward()
        finally:
            Q.FUSED_WEIGHT_BANK = True
        return served, [o.detach() for o in outs]

    a, b = build(), build()
    served_a, outs_a = run(a, True)
    served_b, outs_b = run(b, False)
    torch.cuda.synchronize()
    assert served_a == 67 and served_b == 0
    assert all(torch.equal(x, y) for x, y in zip(outs_a, outs_b))
    for i, (la, lb) in enumerate(zip(a, b)):
        if i != 7:      # (the model-wide pass initialises the step size of a leaf the forward then never runs)
            assert la.step_size_weight.item() == lb.step_size_weight.item()
        for pa, pb in ((la.weight, lb.weight), (la.step_size_weight, lb.step_size_weight)):
            assert (pa.grad is None) == (pb.grad is None)
            if pa.grad is not None:
                assert torch.equal(pa.grad, pb.grad)
    assert a[7].weight.grad is None and a[3].weight.grad is None
    assert "_wq_pending" in a[7].__dict__                       # parked, dropped by the next pass
    Q.prepare_weight_side(a)
    with torch.no_grad():                                       # the no-grad route (sampling) serves the same numbers
        w_ng = a[0]._w_qn()[0]
    Q.FUSED_WEIGHT_BANK = False
    try:
        with torch.no_grad():
            w_pl = a[0]._w_qn()[0]
    finally:
        Q.FUSED_WEIGHT_BANK = True
    assert torch.equal(w_ng, w_pl)


@pytest.mark.parametrize("noise,gain_noise", [(0.0, 0.0), (0.08, 0.0), (0.08, 0.02)])
def test_single_block_adda_layer_equals_the_block_loop(noise, gain_noise):
    """A compute-in-memory Linear whose weight is one mapped block runs digits -> GEMM -> ONE pass (b200sd_adda_tail: ADC,
    recombination, division chain, bias, output quantiser; the fp32 block product kept for the backward pass) instead of
    the block loop's adc_combine + adda_out (and adda_out_bwd + adc_bwd with a recomputed product): outputs bit-identical,
    every gradient equal to fp32 rounding."""
    dev = _dev()
    import copy
    import cim_layers.layers_qn_lsq_adda_cim as A
    torch.manual_seed(21)
    base = A.Linear_lsq_adda_cim(288, 320, weight_bit=4, input_bit=8, output_bit=8, noise_scale=noise, dac_bit=5, adc_bit=8,
                                 adc_gain_1_scale=9.071428571, adc_gain_range=[1 / 64, 1 / 64], adc_adjust_mode="current",
                                 gain_noise_scale=gain_noise, offset_noise_scale=gain_noise, bias=True).to(dev)
    base.weight_noise_unit = torch.randn(base.weight.shape, generator=torch.Generator().manual_seed(4)).to(dev)
    g = torch.Generator().manual_seed(8)
    x = (torch.randn(512, 288, generator=g) * 1.5).to(dev)
    gout = torch.randn(512, 320, generator=g).to(dev)
    with torch.no_grad():
        base(x)                                        # lazy step sizes and the ADC gain settle on the block-loop path
    runs = {}
    for fused in (False, True):
        A.FUSED_SINGLE_BLOCK = fused
        try:
            layer = copy.deepcopy(base)
            layer.weight_noise_unit = base.weight_noise_unit
            xi = x.clone().requires_grad_(True)
            out = layer(xi)
            (out * gout).sum().backward()
            with torch.no_grad():
                out_ng = layer(x)
            runs[fused] = (out.detach(), out_ng, xi.grad, {k: v.grad for k, v in layer.named_parameters()})
        finally:
            A.FUSED_SINGLE_BLOCK = True
    torch.cuda.synchronize()
    o0, n0, dx0, g0 = runs[False]
    o1, n1, dx1, g1 = runs[True]
    assert torch.equal(o0, o1) and torch.equal(n0, n1) and torch.equal(o1, n1)
    assert relerr(dx1, dx0) < 1e-5
    for k in g0:
        assert (g0[k] is None) == (g1[k] is None), k
        if g0[k] is not None:
            assert relerr(g1[k], g0[k]) < 2e-4, (k, relerr(g1[k], g0[k]))

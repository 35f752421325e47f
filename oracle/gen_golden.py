"""Generate tests/golden/*.pt by running the REAL reference (/root/reference, importable only in the build
container) on seeded inputs, and check the oracle restatements against it while doing so.

    python oracle/gen_golden.py            # regenerates every fixture; exits non-zero if the oracle disagrees

A fixture holds: the model config, the construction seed (weights are NOT stored: `torch.manual_seed(seed)` followed
by constructing the module reproduces them bit-for-bit — tests/test_structure.py proves our modules consume the RNG
exactly like the reference's), the inputs, the reference fp32 output, and for information the error of the
reference's own bf16-autocast run against its fp32 run (the yardstick for the bf16 tolerance).
"""
import importlib.util
import os
import sys

import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
GOLD = os.path.join(ROOT, "tests", "golden")


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def relerr(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max()).item()


def l2err(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm()).item()


SMALL_COND = {
    'down_channels': [64, 128, 192, 256], 'mid_channels': [256, 192], 'down_sample': [True, True, True],
    'attn_down': [True, True, True], 'time_emb_dim': 128, 'norm_channels': 16, 'num_heads': 8,
    'conv_out_channels': 64, 'num_down_layers': 2, 'num_mid_layers': 2, 'num_up_layers': 2,
    'condition_config': {
        'condition_types': ['text', 'image'],
        'text_condition_config': {'text_embed_model': 'clip', 'train_text_embed_model': False, 'text_embed_dim': 128,
                                  'cond_drop_prob': 0.1},
        'image_condition_config': {'image_condition_input_channels': 18, 'image_condition_output_channels': 3,
                                   'image_condition_h': 64, 'image_condition_w': 64, 'cond_drop_prob': 0.1}}}
SMALL_UNCOND = {k: v for k, v in SMALL_COND.items() if k != 'condition_config'}
SMALL_UNCOND['attn_down'] = [True, False, True]
SMALL_UNCOND['down_sample'] = [True, True, False]


def make_inputs(seed, batch, cfg, hw=32, ctx_len=77):
    """Synthetic inputs of SURVEY §8d: latents N(0,1), t ~ U[0,1000), text N(0,1), mask one-hot of a random class
    map with the background class dropped (stored compactly as the uint8 class map)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(batch, 4, hw, hw, generator=g)
    t = torch.randint(0, 1000, (batch,), generator=g)
    out = {"x": x, "t": t}
    cc = cfg.get('condition_config')
    if cc:
        if 'text' in cc['condition_types']:
            out["text"] = torch.randn(batch, ctx_len, cc['text_condition_config']['text_embed_dim'], generator=g)
        if 'image' in cc['condition_types']:
            ic = cc['image_condition_config']
            out["mask_classes"] = torch.randint(0, ic['image_condition_input_channels'] + 1,
                                                (batch, ic['image_condition_h'], ic['image_condition_w']),
                                                generator=g).to(torch.uint8)
    return out


def expand_mask(classes, channels):
    """dataset/celeb_dataset.py:172-179: one-hot over channels+1 classes, background (class 0) dropped."""
    oh = torch.nn.functional.one_hot(classes.long(), channels + 1)[..., 1:]
    return oh.permute(0, 3, 1, 2).float().contiguous()


def cond_of(inp, cfg):
    cc = cfg.get('condition_config')
    if not cc:
        return None
    c = {}
    if "text" in inp:
        c["text"] = inp["text"]
    if "mask_classes" in inp:
        c["image"] = expand_mask(inp["mask_classes"], cc['image_condition_config']['image_condition_input_channels'])
    return c


def _cond(types, text_dim=512, mask_hw=512):
    cc = {'condition_types': types}
    if 'text' in types:
        cc['text_condition_config'] = {'text_embed_model': 'clip', 'train_text_embed_model': False,
                                       'text_embed_dim': text_dim, 'cond_drop_prob': 0.9}
    if 'image' in types:
        cc['image_condition_config'] = {'image_condition_input_channels': 18, 'image_condition_output_channels': 3,
                                        'image_condition_h': mask_hw, 'image_condition_w': mask_hw,
                                        'cond_drop_prob': 0.9}
    return cc


# Model_DiT_9L_config.py:139-148 / Model_DiT_12L_config.py:125-133 (timestep_emb_dim = ldm_scaling(2.58) = 192)
DIT_9L = {'hidden_size': 288, 'patch_size': 2, 'timestep_emb_dim': 192, 'num_layers': 9, 'num_heads': 9,
          'head_dim': 32, 'condition_config': _cond(['text', 'image'])}
DIT_12L = dict(DIT_9L, num_layers=12, condition_config=_cond(['image']))
DIT_SMALL = {'hidden_size': 96, 'patch_size': 2, 'timestep_emb_dim': 64, 'num_layers': 2, 'num_heads': 3,
             'head_dim': 32, 'condition_config': _cond(['text', 'image'], text_dim=128, mask_hw=64)}
DIT_SMALL_PLAIN = {k: v for k, v in DIT_SMALL.items() if k != 'condition_config'}


def randomise_zero_init(model, seed):
    """SURVEY §8c trap (i): adaLN projections and proj_out are zero-initialised (transformer.py:147-151,
    transformer_layer.py:73-74) — a default-init DiT returns exactly 0 and every gate is closed. Fixtures draw those
    parameters from N(0, 0.02) (biases too) so the whole network is exercised."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if "adaptive_norm_layer" in name or name.startswith("proj_out"):
                p.copy_(torch.randn(p.shape, generator=g) * 0.02)


def gen_dit():
    sys.path.insert(0, REF)
    from models.transformer import DIT as RefDIT
    sys.path.pop(0)
    O = _load(os.path.join(ROOT, "oracle", "dit_torch.py"), "oracle_dit_torch")
    ok = True
    for name, cfg, batch, hw in [("dit_small_cond", DIT_SMALL, 3, 16), ("dit_small_plain", DIT_SMALL_PLAIN, 2, 16),
                                 ("dit_9l_text_image", DIT_9L, 2, 32), ("dit_12l_image", DIT_12L, 2, 32)]:
        seed = 1111
        torch.manual_seed(seed)
        model = RefDIT(4, cfg).eval()
        randomise_zero_init(model, seed + 7)
        inp = make_inputs(seed + 1, batch, cfg, hw=hw)
        cond = cond_of(inp, cfg)
        with torch.no_grad():
            args = (inp["x"], inp["t"]) + ((cond,) if cond is not None else ())
            ref = model(*args)
            ref_scalar_t = model(inp["x"], torch.as_tensor(int(inp["t"][0])), *args[2:])
            with torch.autocast("cpu", dtype=torch.bfloat16):
                ref_bf16 = model(*args).float()
            orc = O.dit_forward(model.state_dict(), cfg, inp["x"], inp["t"], cond)
            orc_scalar = O.dit_forward(model.state_dict(), cfg, inp["x"], torch.as_tensor(int(inp["t"][0])), cond)
        e_or, e_sc, e_bf = relerr(orc, ref), relerr(orc_scalar, ref_scalar_t), relerr(ref_bf16, ref)
        print(f"{name}: oracle-vs-ref {e_or:.2e} (scalar t {e_sc:.2e})  ref bf16-autocast-vs-fp32 {e_bf:.2e} "
              f"rel-L2 {l2err(ref_bf16, ref):.2e}  |out|max {ref.abs().max():.3f}")
        ok &= e_or < 1e-5 and e_sc < 1e-5
        torch.save({"config": cfg, "seed": seed, "rand_seed": seed + 7, "inputs": inp, "output": ref,
                    "output_scalar_t": ref_scalar_t,
                    "ref_bf16_autocast_relerr": e_bf, "ref_bf16_autocast_l2err": l2err(ref_bf16, ref)},
                   os.path.join(GOLD, name + ".pt"))
    return ok


LSQ_CFG = dict(weight_bit=4, input_bit=8, output_bit=8, noise_scale=0.08)                       # Model_DiT_9L_train.py:541-548
ADDA_CFG = dict(LSQ_CFG, dac_bit=5, adc_bit=8, adc_gain_1_scale=9.071428571, adc_gain_range=[1 / 64, 1 / 64],
                adc_adjust_mode='current')                                                       # :578-590


def gen_cim():
    """cim_layers fixtures: the real reference layers on seeded inputs, weight noise injected (trap iv)."""
    from unittest.mock import MagicMock
    for mod in ("onnx", "onnx.shape_inference", "onnx.numpy_helper", "matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(mod, MagicMock())   # absent here; only plotting / export helpers of the mapper use them
    sys.path.insert(0, REF)
    from cim_layers.layers_qn_lsq import Conv2d_qn_lsq as RConv, Linear_qn_lsq as RLin
    from cim_layers.layers_qn_lsq_adda_cim import Linear_lsq_adda_cim as RAdda
    from cim_weight_mapper.weight_process import gen_weight_split_dict
    sys.path.pop(0)
    O = _load(os.path.join(ROOT, "oracle", "cim_torch.py"), "oracle_cim_torch")
    ok = True
    out = {}
    real_randn_like = torch.randn_like

    def run(layer, x, unit):
        torch.randn_like = lambda t, **k: unit.to(t)       # the ONLY randn_like on the path is the weight noise
        try:
            return layer(x)
        finally:
            torch.randn_like = real_randn_like

    def params_of(layer):
        return {k: v.detach().clone() for k, v in layer.state_dict().items()}

    g = torch.Generator().manual_seed(1111)
    # ---- Linear_qn_lsq: DiT MLP shape, two forwards (first one initialises the step sizes)
    for name, mk, x, conv in [
        ("lsq_linear", lambda: RLin(288, 192, bias=True, **LSQ_CFG), torch.randn(2, 32, 288, generator=g), None),
        ("lsq_linear_clean", lambda: RLin(96, 64, bias=False, **dict(LSQ_CFG, noise_scale=0)),
         torch.randn(48, 96, generator=g) * 3, None),
        ("lsq_conv3x3", lambda: RConv(32, 64, 3, 1, 1, 1, bias=True, **LSQ_CFG), torch.randn(2, 32, 16, 16, generator=g),
         dict(stride=1, padding=1)),
        ("lsq_conv1x1", lambda: RConv(64, 32, 1, 1, 0, 1, bias=True, **dict(LSQ_CFG, weight_bit=8)),
         torch.randn(2, 64, 8, 8, generator=g), dict(stride=1, padding=0)),
    ]:
        torch.manual_seed(5)
        layer = mk()
        init = params_of(layer)
        unit = torch.randn(layer.weight.shape, generator=g).to(torch.float16).float()   # stored as fp16: keep it exact
        x2 = torch.randn(x.shape, generator=g) * 1.3
        with torch.no_grad():
            y1 = run(layer, x, unit)
        after = params_of(layer)
        layer.zero_grad()
        y2 = run(layer, x2, unit)
        tgt = torch.randn(y2.shape, generator=g)
        ((y2 - tgt) ** 2).mean().backward()
        grads = {k: v.grad.detach().clone() for k, v in layer.named_parameters() if v.grad is not None}
        cfg = dict(LSQ_CFG, weight_bit=layer.weight_bit, noise_scale=layer.noise_scale)
        p = {k: v.clone() for k, v in init.items()}
        with torch.no_grad():
            o1 = O.lsq_layer(p, cfg, x, unit, conv)
        e1 = (o1 - y1).abs().max().item()
        ep = max((p[k] - after[k]).abs().max().item() for k in after)
        p2 = {k: v.clone().requires_grad_(True) for k, v in after.items()}
        assert all(torch.equal(after[k], init[k]) for k in after if after[k].dim() > 0)
        o2 = O.lsq_layer(p2, cfg, x2, unit, conv)
        ((o2 - tgt) ** 2).mean().backward()
        eg = max(((p2[k].grad - grads[k]).abs().max() / grads[k].abs().max().clamp_min(1e-20)).item() for k in grads)
        print(f"cim {name}: oracle-vs-ref out {e1:.1e} lazy-init params {ep:.1e} second forward {(o2 - y2).abs().max().item():.1e} grads {eg:.1e}")
        ok &= e1 == 0 and ep == 0 and (o2 - y2).abs().max().item() == 0 and eg < 1e-5
        out[name] = {"cfg": cfg, "conv": conv, "init": init, "after": {k: v for k, v in after.items() if v.dim() == 0},
                     "x": x, "x2": x2, "unit": unit.to(torch.float16), "y1": y1, "y2": y2.detach(), "target": tgt,
                     "grads": grads}

    # ---- Linear_lsq_adda_cim: one and two row blocks (array height 576), with and without ADC noise
    for name, (K, N), extra, block in [("adda_linear_288", (288, 160), {}, [576, 2048]),
                                       ("adda_linear_1152", (1152, 96), {}, [576, 2048]),
                                       ("adda_linear_noise", (192, 96), dict(gain_noise_scale=0.02, offset_noise_scale=0.01,
                                                                             noise_scale=0.0), [64, 40])]:
        cfgk = dict(ADDA_CFG, **extra)
        torch.manual_seed(9)
        layer = RAdda(K, N, bias=True, **cfgk)
        layer.weight_mapping_info = gen_weight_split_dict(layer, block)
        init = params_of(layer)
        unit = torch.randn(layer.weight.shape, generator=g).to(torch.float16).float()
        x = torch.randn(2, 32, K, generator=g)
        x2 = torch.randn(2, 32, K, generator=g) * 0.7
        with torch.no_grad():
            y1 = run(layer, x, unit)
            after = params_of(layer)
        layer.zero_grad()
        x2r = x2.clone().requires_grad_(True)
        y2 = run(layer, x2r, unit)
        tgt = torch.randn(y2.shape, generator=g)
        ((y2 - tgt) ** 2).mean().backward()
        grads = {k: v.grad.detach().clone() for k, v in layer.named_parameters() if v.grad is not None}
        grads["x"] = x2r.grad.detach().clone()
        y2 = y2.detach()
        c = dict(cfgk, adc_gain_min=layer.adc_gain_min, adc_gain_max=layer.adc_gain_max)
        p = {k: v.clone() for k, v in init.items()}
        noise = dict(gain_noise=layer.gain_noise, offset_noise=layer.offset_noise) if extra else {}
        with torch.no_grad():
            o1 = O.adda_linear(p, c, x, layer.weight_mapping_info, unit, **noise)
            ep = max((p[k] - after[k]).abs().max().item() for k in after)
        p2 = {k: v.clone().requires_grad_(True) for k, v in p.items()}
        x2o = x2.clone().requires_grad_(True)
        o2 = O.adda_linear(p2, c, x2o, layer.weight_mapping_info, unit, **noise)
        ((o2 - tgt) ** 2).mean().backward()
        og = {k: v.grad for k, v in p2.items()}
        og["x"] = x2o.grad
        eg = max(((og[k] - grads[k]).abs().max() / grads[k].abs().max().clamp_min(1e-20)).item() for k in grads)
        e1, e2 = (o1 - y1).abs().max().item(), (o2.detach() - y2).abs().max().item()
        print(f"cim {name}: oracle-vs-ref out {e1:.1e} lazy-init params {ep:.1e} second forward {e2:.1e} grads {eg:.1e} "
              f"blocks {len(layer.weight_mapping_info)} adc_gain {after['adc_gain'].item():.6f}")
        ok &= e1 == 0 and ep == 0 and e2 == 0 and eg < 1e-5
        out[name] = {"cfg": {k: v for k, v in c.items()}, "mapping": layer.weight_mapping_info, "init": init,
                     "after": {k: v for k, v in after.items() if v.dim() == 0}, "x": x, "x2": x2,
                     "unit": unit.to(torch.float16), "y1": y1, "y2": y2, "target": tgt, "grads": grads,
                     "gain_noise": layer.gain_noise.clone(), "offset_noise": layer.offset_noise.clone()}
    torch.save(out, os.path.join(GOLD, "cim_layers.pt"))
    return ok


def gen_cim_conv():
    """k x k Conv2d_lsq_adda_cim fixtures (the leaves of the ReLU-UNet CIM pipeline, Model_Unet_train.py): the real
    reference layer on seeded inputs with injected weight noise; oracle/cim_torch.py adda_linear(conv=...) must reproduce
    outputs, lazily initialised parameters and every gradient. -> tests/golden/cim_adda_conv.pt"""
    from unittest.mock import MagicMock
    for mod in ("onnx", "onnx.shape_inference", "onnx.numpy_helper", "matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(mod, MagicMock())
    sys.path.insert(0, REF)
    from cim_layers.layers_qn_lsq_adda_cim import Conv2d_lsq_adda_cim as RConv
    from cim_weight_mapper.weight_process import gen_weight_split_dict
    sys.path.pop(0)
    O = _load(os.path.join(ROOT, "oracle", "cim_torch.py"), "oracle_cim_torch")
    real_randn_like = torch.randn_like

    def run(layer, x, unit):
        torch.randn_like = lambda t, **k: unit.to(t)
        try:
            return layer(x)
        finally:
            torch.randn_like = real_randn_like

    g = torch.Generator().manual_seed(2222)
    ok, out = True, {}
    for name, (cin, cout, k, st, pd), hw, extra, block in [
            ("adda_conv3x3", (32, 48, 3, 1, 1), 12, {}, [128, 2048]),                       # K = 288 -> 3 row blocks
            ("adda_conv3x3_clean", (24, 40, 3, 1, 1), 8, dict(noise_scale=0.0), [576, 2048]),
            ("adda_conv4x4s2", (16, 32, 4, 2, 1), 12, dict(noise_scale=0.0), [576, 2048])]:
        cfgk = dict(ADDA_CFG, **extra)
        torch.manual_seed(9)
        layer = RConv(cin, cout, k, st, pd, 1, bias=True, **cfgk)
        layer.weight_mapping_info = gen_weight_split_dict(layer, block)
        init = {kk: v.detach().clone() for kk, v in layer.state_dict().items()}
        unit = torch.randn(layer.weight.shape, generator=g).to(torch.float16).float()
        x = torch.randn(2, cin, hw, hw, generator=g)
        x2 = torch.randn(2, cin, hw, hw, generator=g) * 0.7
        with torch.no_grad():
            y1 = run(layer, x, unit)
            after = {kk: v.detach().clone() for kk, v in layer.state_dict().items()}
        layer.zero_grad()
        x2r = x2.clone().requires_grad_(True)
        y2 = run(layer, x2r, unit)
        tgt = torch.randn(y2.shape, generator=g)
        ((y2 - tgt) ** 2).mean().backward()
        grads = {kk: v.grad.detach().clone() for kk, v in layer.named_parameters() if v.grad is not None}
        grads["x"] = x2r.grad.detach().clone()
        y2 = y2.detach()
        c = dict(cfgk, adc_gain_min=layer.adc_gain_min, adc_gain_max=layer.adc_gain_max)
        conv = dict(kernel_size=k, stride=st, padding=pd)
        p = {kk: v.clone() for kk, v in init.items()}
        with torch.no_grad():
            o1 = O.adda_linear(p, c, x, layer.weight_mapping_info, unit, conv=conv)
            ep = max((p[kk] - after[kk]).abs().max().item() for kk in after)
        p2 = {kk: v.clone().requires_grad_(True) for kk, v in p.items()}
        x2o = x2.clone().requires_grad_(True)
        o2 = O.adda_linear(p2, c, x2o, layer.weight_mapping_info, unit, conv=conv)
        ((o2 - tgt) ** 2).mean().backward()
        og = {kk: v.grad for kk, v in p2.items()}
        og["x"] = x2o.grad
        eg = max(((og[kk] - grads[kk]).abs().max() / grads[kk].abs().max().clamp_min(1e-20)).item() for kk in grads)
        e1, e2 = (o1 - y1).abs().max().item(), (o2.detach() - y2).abs().max().item()
        print(f"cim {name}: oracle-vs-ref out {e1:.1e} lazy-init params {ep:.1e} second forward {e2:.1e} grads {eg:.1e} "
              f"blocks {len(layer.weight_mapping_info)} adc_gain {after['adc_gain'].item():.6f}")
        ok &= e1 == 0 and ep == 0 and e2 == 0 and eg < 1e-5
        out[name] = {"cfg": dict(c), "conv": conv, "shape": (cin, cout), "mapping": layer.weight_mapping_info, "init": init,
                     "after": {kk: v for kk, v in after.items() if v.dim() == 0}, "x": x, "x2": x2,
                     "unit": unit.to(torch.float16), "y1": y1, "y2": y2, "target": tgt, "grads": grads}
    torch.save(out, os.path.join(GOLD, "cim_adda_conv.pt"))
    return ok


def sampler_model(x, t, cond=None):
    """Deterministic stand-in denoiser whose arithmetic (*, +, / on fp32) is bit-reproducible on any device."""
    s = (t.float() * 0.001).view(-1, 1, 1, 1)   # (not "/ 1000": CUDA divides by a scalar through its reciprocal)
    out = x * 0.5 + s * x.flip(-1)
    if cond is not None:
        out = out + cond["image"]
    return out


def gen_samplers():
    """DDPMSampler / DDIMSampler of the real reference on a stand-in model with every randn_like draw recorded."""
    sys.path.insert(0, REF)
    import scheduler.linear_noise_scheduler as ref_sched
    sys.path.pop(0)
    import numpy as np
    S = _load(os.path.join(ROOT, "oracle", "scheduler_np.py"), "oracle_scheduler_np")
    os.makedirs(GOLD, exist_ok=True)
    g = torch.Generator().manual_seed(77)
    draws = []
    real_randn_like = torch.randn_like

    def recording_randn_like(x, *a, **k):
        z = torch.randn(x.shape, generator=g)
        draws.append(z)
        return z

    out, ok = {}, True
    x_T = torch.randn(3, 4, 8, 8, generator=g)
    cond = {"image": torch.randn(3, 4, 8, 8, generator=g) * 0.1}
    torch.randn_like = recording_randn_like
    try:
        for name, kw in (("ddim_linear_eta0", dict(steps=10, method="linear", eta=0.0)),
                         ("ddim_quadratic_eta05", dict(steps=7, method="quadratic", eta=0.5)),
                         ("ddim_linear_eta1_trace", dict(steps=20, method="linear", eta=1.0, only_return_x_0=False,
                                                         interval=4))):
            draws.clear()
            smp = ref_sched.DDIMSampler(sampler_model, (0.0015, 0.0195), 1000)
            res = smp.forward(x_T.clone(), cond, None, **kw)
            out[name] = {"kw": kw, "beta": (0.0015, 0.0195), "T": 1000, "out": res.clone(), "z": [z.clone() for z in draws]}
            # oracle restatement against the reference
            bar = S.sampler_alpha_bar((0.0015, 0.0195), 1000)[2]
            ts, tp = S.ddim_timesteps(1000, kw["steps"], kw["method"])
            x = x_T.numpy()
            for k, i in enumerate(reversed(range(kw["steps"]))):
                t = torch.full((3,), int(ts[i]), dtype=torch.long)
                x = S.ddim_step(bar, x, sampler_model(torch.from_numpy(x), t, cond).numpy(), draws[k].numpy(), int(ts[i]),
                                int(tp[i]), kw["eta"])
            final = res if res.dim() == 4 else res[:, -1]
            want = final.numpy() if res.dim() == 4 else np.clip(x, -1, 1)
            same = np.array_equal(x if res.dim() == 4 else want, final.numpy())
            print(f"{name}: oracle == reference: {same}")
            ok &= bool(same)
        draws.clear()
        smp = ref_sched.DDPMSampler(sampler_model, (1e-4, 0.02), 40)
        res = smp.forward(x_T.clone())
        out["ddpm_T40"] = {"beta": (1e-4, 0.02), "T": 40, "out": res.clone(), "z": [z.clone() for z in draws],
                           "state_keys": list(smp.state_dict().keys())}
        x = x_T.numpy()
        for k, ts_ in enumerate(reversed(range(40))):
            t = torch.full((3,), ts_, dtype=torch.long)
            z = draws[k].numpy() if ts_ > 0 else None
            x = S.ddpm_step((1e-4, 0.02), 40, x, sampler_model(torch.from_numpy(x), t).numpy(), z, ts_)
        same = np.array_equal(x, res.numpy())
        print(f"ddpm_T40: oracle == reference: {same}")
        ok &= bool(same)
    finally:
        torch.randn_like = real_randn_like
    out["x_T"], out["cond_image"] = x_T, cond["image"]
    torch.save(out, os.path.join(GOLD, "samplers.pt"))
    return ok


def gen_vqvae_encode():
    """VQVAE encoder + quantiser (SURVEY f-4): the call that produces the latents the LDM trainers consume."""
    sys.path.insert(0, REF)
    from models.vqvae import VQVAE as RefVQVAE
    sys.path.pop(0)
    O = _load(os.path.join(ROOT, "oracle", "unet_torch.py"), "oracle_unet_torch")
    acfg = yaml.safe_load(open(os.path.join(REF, "config/celebhq.yaml")))['autoencoder_params']
    torch.manual_seed(1111)
    vae = RefVQVAE(3, acfg).eval()
    x_im = torch.rand(2, 3, 64, 64, generator=torch.Generator().manual_seed(6)) * 2 - 1
    feats = {}
    hook = vae.pre_quant_conv.register_forward_hook(lambda m, i, o: feats.__setitem__("pre", o.detach().float().clone()))
    with torch.no_grad():
        with torch.autocast("cpu", dtype=torch.bfloat16):
            vae.encode(x_im)
        pre_bf16 = feats["pre"]
        zq, losses = vae.encode(x_im)
        pre = feats["pre"]
        hook.remove()
        _, _, ref_idx = vae.quantize(pre)
        o_pre = O.vqvae_encode_latent(vae.state_dict(), acfg, x_im)
        o_q, o_mse, o_idx, d2 = O.vq_quantize(pre, vae.embedding.weight.detach())
    e_or, e_bf = relerr(o_pre, pre), relerr(pre_bf16, pre)
    same = (o_idx == ref_idx).float().mean().item()
    rows = torch.arange(d2.shape[0])   # where the reference's cdist picked another code it must be a numerical tie
    best = d2[rows, o_idx.reshape(-1)]
    gap = ((d2[rows, ref_idx.reshape(-1)] - best) / best.clamp_min(1e-30)).max().item()
    print(f"vqvae_encode: oracle-vs-ref latent {e_or:.2e}  ref bf16-autocast-vs-fp32 {e_bf:.2e}  indices equal "
          f"{same:.4f}  worst relative distance gap of a differing index {gap:.2e}  "
          f"mse {o_mse.item():.6f} vs {losses['codebook_loss'].item():.6f}  quant_out equal {torch.equal(o_q, zq) if same == 1 else 'n/a'}")
    ok = e_or < 1e-5 and gap < 1e-5 and abs(o_mse.item() - losses['codebook_loss'].item()) < 1e-6
    torch.save({"config": acfg, "seed": 1111, "inputs": {"x": x_im}, "pre_quant": pre, "quant_out": zq,
                "indices": ref_idx, "codebook_loss": losses["codebook_loss"], "commitment_loss": losses["commitment_loss"],
                "ref_bf16_autocast_relerr": e_bf, "ref_bf16_autocast_l2err": l2err(pre_bf16, pre)},
               os.path.join(GOLD, "vqvae_encode.pt"))
    return ok


def gen_percent_opt():
    """PercentOptimizerFP (cim_layers/IBA_optimizer.py:80-143): the real class, 4 steps on 4 tensors with recorded
    gradients (one tensor receives no gradient at step 2: its step counter falls behind, :93-94,105)."""
    OT = _load(os.path.join(ROOT, "oracle", "optim_torch.py"), "oracle_optim_torch")
    spec = importlib.util.spec_from_file_location("ref_IBA_optimizer", os.path.join(REF, "cim_layers", "IBA_optimizer.py"))
    ref_mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref_mod)
    g = torch.Generator().manual_seed(321)
    shapes = [(96, 40), (33000,), (7,), (3, 5, 2, 2)]
    init = [torch.randn(s, generator=g) * (0.5 if i else 0.05) for i, s in enumerate(shapes)]
    grads = [[torch.randn(s, generator=g) * 10 ** float(torch.randint(-4, 2, (1,), generator=g)) for s in shapes]
             for _ in range(4)]
    grads[1][2] = None
    lr, betas, eps = 2e-3, (0.9, 0.999), 1e-8
    params = [torch.nn.Parameter(t.clone()) for t in init]
    opt = ref_mod.PercentOptimizerFP(params, lr=lr, betas=betas, eps=eps)
    mine = [t.clone() for t in init]
    m, v, steps = [torch.zeros_like(t) for t in init], [torch.zeros_like(t) for t in init], [0] * len(init)
    after = []
    for gs in grads:
        for p, gr in zip(params, gs):
            p.grad = None if gr is None else gr.clone()
        opt.step()
        for k, gr in enumerate(gs):       # the oracle takes one step count per call: group tensors by their own count
            if gr is not None:
                steps[k] += 1
        for st in sorted(set(steps)):
            sel = [k for k in range(len(init)) if steps[k] == st and gs[k] is not None]
            OT.percent_step([mine[k] for k in sel], [gs[k] for k in sel], [m[k] for k in sel], [v[k] for k in sel], st,
                            lr, betas, eps)
        for a, b in zip(params, mine):
            assert torch.equal(a.data, b), "oracle PercentOptimizerFP differs from the reference"
        after.append([p.data.clone() for p in params])
    torch.save({"init": init, "grads": grads, "lr": lr, "betas": betas, "eps": eps, "after": {2: after[1], 4: after[3]},
                "exp_avg": [opt.state[p]["exp_avg"].clone() for p in params],
                "exp_avg_sq": [opt.state[p]["exp_avg_sq"].clone() for p in params],
                "steps": [opt.state[p]["step"] for p in params]}, os.path.join(GOLD, "percent_opt.pt"))
    print("percent_opt: oracle == reference over 4 steps; fixture written")


def gen_enhance():
    """Enhance wrappers, Linear family (cim_qn_train/layers_enhance.py): the real modules on CPU with perturbed
    parameters; outputs and gradients of every parameter and of the input; plus a table of the group / rank search."""
    import itertools
    import json
    from unittest.mock import MagicMock
    for m in ("matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(m, MagicMock())
    spec = importlib.util.spec_from_file_location("ref_layers_enhance", os.path.join(REF, "cim_qn_train", "layers_enhance.py"))
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    OE = _load(os.path.join(ROOT, "oracle", "enhance_torch.py"), "oracle_enhance_torch")
    g = torch.Generator().manual_seed(77)
    torch.manual_seed(77)   # the branch convolutions keep nn.Conv2d's default initialisation (global generator)
    out = {"cases": []}

    def bf16_grad_err(mod_, x_, r_, grads_):
        """The reference module's own bf16-autocast gradient error (relative L2 per parameter): the yardstick the GPU
        tests hold the kernels to where a ReLU makes the gradient ill-conditioned."""
        for p_ in mod_.parameters():
            p_.grad = None
        xa = x_.detach().clone().requires_grad_(True)
        with torch.autocast("cpu", dtype=torch.bfloat16):
            ya = mod_(xa)
        (ya.float() * r_).sum().backward()
        return {k_: float((p_.grad - grads_[k_]).norm() / (grads_[k_].norm() + 1e-12)) for k_, p_ in mod_.named_parameters()}

    for kind, fin, fout, ops_f, relu, sig in (("branch", 288, 288, 0.05, False, True), ("branch", 96, 64, 0.1, True, True),
                                              ("branch", 64, 192, 0.05, False, False), ("layer", 96, 192, 0.05, None, None),
                                              ("layer", 96, 64, 0.25, None, None)):
        lin = torch.nn.Linear(fin, fout)
        with torch.no_grad():
            lin.weight.copy_(torch.randn(fout, fin, generator=g) / fin ** 0.5)
            lin.bias.copy_(torch.randn(fout, generator=g) * 0.1)
        mod = ref.EnhanceBranchLinear_LoR(lin, ops_f, relu=relu, sigmoid=sig) if kind == "branch" else \
            ref.EnhanceLayerLinear(lin, ops_f)
        with torch.no_grad():
            for n_, p_ in mod.named_parameters():
                if "original_linear" in n_:
                    continue
                if n_ == "blend_factor":
                    p_.copy_(torch.tensor(-0.4 if sig else 0.3))
                else:
                    p_.add_(torch.randn(p_.shape, generator=g) * 0.1)
        x = torch.randn(2, 24, fin, generator=g).requires_grad_(True)
        r = torch.randn(2, 24, fout, generator=g)
        y = mod(x)
        (y * r).sum().backward()
        sd = {k: v.detach().clone() for k, v in mod.state_dict().items()}
        grads = {k: p_.grad.clone() for k, p_ in mod.named_parameters()}
        # oracle == reference
        ps = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
        x2 = x.detach().clone().requires_grad_(True)
        if kind == "branch":
            y2, _, _ = OE.enhance_branch_linear(x2, ps["original_linear.weight"], ps["original_linear.bias"],
                                                ps["enhance_branch_1.weight"], ps["enhance_branch_1.bias"],
                                                ps["enhance_branch_2.weight"], ps["enhance_branch_2.bias"],
                                                ps["blend_factor"], relu, sig)
        else:
            y2 = OE.enhance_layer_linear(x2, ps["original_linear.weight"], ps["original_linear.bias"],
                                         ps["fc1.weight"], ps["fc1.bias"])
        (y2 * r).sum().backward()
        assert torch.equal(y, y2), "oracle Enhance forward differs"
        # gradients: same formulas, but autograd may order the accumulation of a tensor used twice differently
        for name_, a_, b_ in [("x", x.grad, x2.grad)] + [(k, grads[k], ps[k].grad) for k in grads]:
            d_ = (a_ - b_).abs().max().item() / (a_.abs().max().item() + 1e-30)
            assert d_ < 1e-5, f"oracle Enhance gradient of {name_} differs: {d_:.2e}"
        out["cases"].append({"kind": kind, "in": fin, "out": fout, "ops_factor": ops_f, "relu": relu, "sigmoid": sig,
                             "state": sd, "x": x.detach().clone(), "r": r, "y": y.detach().clone(),
                             "dx": x.grad.clone(), "grads": grads, "ref_bf16_grad_err": bf16_grad_err(mod, x, r, grads),
                             "std_mean": None if kind != "branch" else [float(mod.std_ori.detach()), float(mod.mean_ori.detach()),
                                                                         float(mod.std_enhance.detach()), float(mod.mean_enhance.detach())]})
    out["conv_cases"] = []
    for kind, ci, co, k, pad, ops_f, relu, sig in (("branch", 16, 8, 1, 0, 0.1, False, True), ("branch", 32, 16, 1, 0, 0.25, True, False),
                                                    ("layer", 16, 32, 3, 1, 0.05, None, None), ("layer", 16, 8, 1, 0, 0.25, None, None)):
        conv = torch.nn.Conv2d(ci, co, k, padding=pad)
        with torch.no_grad():
            conv.weight.copy_(torch.randn(conv.weight.shape, generator=g) / (ci * k * k) ** 0.5)
            conv.bias.copy_(torch.randn(co, generator=g) * 0.1)
        mod = ref.EnhanceBranchConv2d_LoR(conv, ops_f, relu=relu, sigmoid=sig) if kind == "branch" else \
            ref.EnhanceLayerConv2d(conv, ops_f)
        with torch.no_grad():
            for n_, p_ in mod.named_parameters():
                if "original_conv" in n_:
                    continue
                if n_ == "blend_factor":
                    p_.copy_(torch.tensor(-0.4 if sig else 0.3))
                else:
                    p_.add_(torch.randn(p_.shape, generator=g) * 0.1)
        x = torch.randn(3, ci, 8, 8, generator=g).requires_grad_(True)
        r = torch.randn(3, co, 8, 8, generator=g)
        y = mod(x)
        (y * r).sum().backward()
        sd = {k_: v.detach().clone() for k_, v in mod.state_dict().items()}
        grads = {k_: p_.grad.clone() for k_, p_ in mod.named_parameters()}
        ps = {k_: v.clone().requires_grad_(True) for k_, v in sd.items()}
        x2 = x.detach().clone().requires_grad_(True)
        if kind == "branch":
            y2 = OE.enhance_branch_conv1x1(x2, ps["original_conv.weight"], ps["original_conv.bias"],
                                           ps["enhance_branch_1.weight"], ps["enhance_branch_1.bias"], mod.enhance_branch_1.groups,
                                           ps["enhance_branch_2.weight"], ps["enhance_branch_2.bias"], mod.enhance_branch_2.groups,
                                           ps["blend_factor"], relu, sig)
        else:
            y2 = OE.enhance_layer_conv(x2, ps["original_conv.weight"], ps["original_conv.bias"], 1, pad,
                                       ps["conv_enhance.weight"], ps["conv_enhance.bias"], mod.groups)
        (y2 * r).sum().backward()
        assert torch.equal(y, y2), "oracle Enhance conv forward differs"
        for name_, a_, b_ in [("x", x.grad, x2.grad)] + [(k_, grads[k_], ps[k_].grad) for k_ in grads]:
            d_ = (a_ - b_).abs().max().item() / (a_.abs().max().item() + 1e-30)
            assert d_ < 1e-5, f"oracle Enhance conv gradient of {name_} differs: {d_:.2e}"
        out["conv_cases"].append({"kind": kind, "cin": ci, "cout": co, "k": k, "pad": pad, "ops_factor": ops_f, "relu": relu,
                                  "sigmoid": sig, "state": sd, "x": x.detach().clone(), "r": r, "y": y.detach().clone(),
                                  "dx": x.grad.clone(), "grads": grads,
                                  "ref_bf16_grad_err": bf16_grad_err(mod, x, r, grads)})
    table = []
    for i, o, f in itertools.product([28, 96, 192, 288, 512, 864, 1152], [64, 288, 512, 864, 1152, 1728],
                                     [0.02, 0.05, 0.1, 0.25]):
        table.append(["lora_linear", i, o, f, ref.analyze_lora_linear_ops(i, o, f)])
        table.append(["enhance_linear", i, o, f, ref.analyze_enhance_linear_ops(i, o, f)])
    for i, o, k, f in itertools.product([18, 64, 128, 384], [3, 64, 256, 384], [1, 3], [0.05, 0.1]):
        table.append(["lora_conv", i, o, k, f, ref.analyze_lora_conv_ops(i, o, k, k, f)])
        table.append(["enhance_conv", i, o, k, f, ref.analyze_enhance_conv_ops(i, o, k, k, f)])
    torch.save(out, os.path.join(GOLD, "enhance_linear.pt"))
    with open(os.path.join(GOLD, "enhance_search_table.json"), "w") as fjs:
        json.dump(table, fjs)
    print(f"enhance: oracle == reference on {len(out['cases'])} + {len(out['conv_cases'])} modules; {len(table)} search results recorded")


def gen_enhance_conv_kxk():
    """EnhanceBranchConv2d_LoR around k x k / strided original convolutions (the ReLU-UNet's 3x3 and 4x4-stride-2 leaves,
    layers_enhance.py:439-535): the real reference module, oracle/enhance_torch.py enhance_branch_conv must reproduce
    output and gradients. -> tests/golden/enhance_conv_kxk.pt"""
    from unittest.mock import MagicMock
    for m in ("matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(m, MagicMock())
    spec = importlib.util.spec_from_file_location("ref_layers_enhance", os.path.join(REF, "cim_qn_train", "layers_enhance.py"))
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    OE = _load(os.path.join(ROOT, "oracle", "enhance_torch.py"), "oracle_enhance_torch")
    g = torch.Generator().manual_seed(78)
    torch.manual_seed(78)
    cases = []
    for ci, co, k, st, pad, hw, ops_f, relu, sig in ((16, 16, 3, 1, 1, 8, 0.1, False, True), (16, 24, 3, 1, 1, 8, 0.1, True, False),
                                                      (16, 16, 4, 2, 1, 8, 0.1, False, True)):
        conv = torch.nn.Conv2d(ci, co, k, stride=st, padding=pad)
        with torch.no_grad():
            conv.weight.copy_(torch.randn(conv.weight.shape, generator=g) / (ci * k * k) ** 0.5)
            conv.bias.copy_(torch.randn(co, generator=g) * 0.1)
        mod = ref.EnhanceBranchConv2d_LoR(conv, ops_f, relu=relu, sigmoid=sig)
        with torch.no_grad():
            for n_, p_ in mod.named_parameters():
                if "original_conv" in n_:
                    continue
                if n_ == "blend_factor":
                    p_.copy_(torch.tensor(-0.4 if sig else 0.3))
                else:
                    p_.add_(torch.randn(p_.shape, generator=g) * 0.1)
        x = torch.randn(3, ci, hw, hw, generator=g).requires_grad_(True)
        y = mod(x)
        r = torch.randn(y.shape, generator=g)
        (y * r).sum().backward()
        sd = {k_: v.detach().clone() for k_, v in mod.state_dict().items()}
        grads = {k_: p_.grad.clone() for k_, p_ in mod.named_parameters()}
        ps = {k_: v.clone().requires_grad_(True) for k_, v in sd.items()}
        x2 = x.detach().clone().requires_grad_(True)
        y2 = OE.enhance_branch_conv(x2, ps["original_conv.weight"], ps["original_conv.bias"], st, pad,
                                    ps["enhance_branch_1.weight"], ps["enhance_branch_1.bias"], mod.enhance_branch_1.groups,
                                    ps["enhance_branch_2.weight"], ps["enhance_branch_2.bias"], mod.enhance_branch_2.groups,
                                    ps["blend_factor"], relu, sig)
        (y2 * r).sum().backward()
        assert torch.equal(y, y2), "oracle Enhance k x k conv forward differs"
        for name_, a_, b_ in [("x", x.grad, x2.grad)] + [(k_, grads[k_], ps[k_].grad) for k_ in grads]:
            d_ = (a_ - b_).abs().max().item() / (a_.abs().max().item() + 1e-30)
            assert d_ < 1e-5, f"oracle Enhance k x k conv gradient of {name_} differs: {d_:.2e}"
        for p_ in mod.parameters():
            p_.grad = None
        xa = x.detach().clone().requires_grad_(True)
        with torch.autocast("cpu", dtype=torch.bfloat16):
            ya = mod(xa)
        (ya.float() * r).sum().backward()
        bf_err = {k_: float((p_.grad - grads[k_]).norm() / (grads[k_].norm() + 1e-12)) for k_, p_ in mod.named_parameters()}
        cases.append({"kind": "branch", "cin": ci, "cout": co, "k": k, "stride": st, "pad": pad, "ops_factor": ops_f,
                      "relu": relu, "sigmoid": sig, "state": sd, "x": x.detach().clone(), "r": r, "y": y.detach().clone(),
                      "dx": x.grad.clone(), "grads": grads, "ref_bf16_grad_err": bf_err,
                      "groups": (mod.enhance_branch_1.groups, mod.enhance_branch_2.groups)})
    torch.save({"conv_cases": cases}, os.path.join(GOLD, "enhance_conv_kxk.pt"))
    print(f"enhance k x k: oracle == reference on {len(cases)} modules")
    return True


def main():
    if "--only-enhance-kxk" in sys.argv:
        gen_enhance_conv_kxk()
        return
    if "--only-enhance" in sys.argv:
        gen_enhance()
        return
    if "--only-percent-opt" in sys.argv:
        gen_percent_opt()
        return
    if "--only-vqvae-encode" in sys.argv:
        if not gen_vqvae_encode():
            print("ORACLE DISAGREES WITH THE REFERENCE")
            sys.exit(1)
        return
    if "--only-samplers" in sys.argv:
        if not gen_samplers():
            print("ORACLE DISAGREES WITH THE REFERENCE")
            sys.exit(1)
        return
    if "--only-cim-conv" in sys.argv:
        if not gen_cim_conv():
            print("ORACLE DISAGREES WITH THE REFERENCE")
            sys.exit(1)
        return
    if "--only-cim" in sys.argv:
        if not gen_cim():
            print("ORACLE DISAGREES WITH THE REFERENCE")
            sys.exit(1)
        return
    if "--only-dit" in sys.argv:
        if not gen_dit():
            print("ORACLE DISAGREES WITH THE REFERENCE")
            sys.exit(1)
        return
    sys.path.insert(0, REF)
    from models.unet_cond_base import Unet as RefCond
    from models.unet_base import Unet as RefUncond
    from models.unet_cond_base_relu import Unet as RefCondRelu
    from models.vqvae import VQVAE as RefVQVAE
    sys.path.pop(0)
    O = _load(os.path.join(ROOT, "oracle", "unet_torch.py"), "oracle_unet_torch")
    os.makedirs(GOLD, exist_ok=True)
    full_cond = yaml.safe_load(open(os.path.join(REF, "config/celebhq_text_image_cond.yaml")))
    full_uncond = yaml.safe_load(open(os.path.join(REF, "config/celebhq.yaml")))
    cases = [
        ("unet_small_cond", RefCond, SMALL_COND, 3, False),
        ("unet_small_uncond", RefUncond, SMALL_UNCOND, 2, False),
        ("unet_small_cond_relu", RefCondRelu, SMALL_COND, 2, True),
        ("unet_full_cond", RefCond, full_cond['ldm_params'], 2, False),
        ("unet_full_uncond", RefUncond, full_uncond['ldm_params'], 2, False),
    ]
    ok = True
    for name, cls, cfg, batch, relu in cases:
        seed = 1111
        torch.manual_seed(seed)
        model = cls(4, cfg).eval()
        inp = make_inputs(seed + 1, batch, cfg)
        cond = cond_of(inp, cfg)
        with torch.no_grad():
            args = (inp["x"], inp["t"]) + ((cond,) if cond is not None else ())
            ref = model(*args)
            model.train()  # train-mode MHA takes the non-fused path (SURVEY §8c trap ii): must agree
            ref_train = model(*args)
            model.eval()
            with torch.autocast("cpu", dtype=torch.bfloat16):
                ref_bf16 = model(*args).float()
            orc = O.unet_forward(model.state_dict(), cfg, inp["x"], inp["t"], cond, relu=relu)
        e_or, e_tr, e_bf = relerr(orc, ref), relerr(ref_train, ref), relerr(ref_bf16, ref)
        print(f"{name}: oracle-vs-ref {e_or:.2e}  ref train-vs-eval {e_tr:.2e}  ref bf16-autocast-vs-fp32 {e_bf:.2e}")
        ok &= e_or < 1e-5
        print(f"    rel-L2: ref bf16-autocast-vs-fp32 {l2err(ref_bf16, ref):.2e}")
        torch.save({"config": cfg, "seed": seed, "relu": relu, "inputs": inp, "output": ref,
                    "ref_bf16_autocast_relerr": e_bf, "ref_bf16_autocast_l2err": l2err(ref_bf16, ref)},
                   os.path.join(GOLD, name + ".pt"))

    # VQVAE decoder
    acfg = full_uncond['autoencoder_params']
    torch.manual_seed(1111)
    vae = RefVQVAE(3, acfg).eval()
    z = torch.randn(1, 4, 32, 32, generator=torch.Generator().manual_seed(5))
    with torch.no_grad():
        ref = vae.decode(z)
        orc = O.vqvae_decode(vae.state_dict(), acfg, z)
        with torch.autocast("cpu", dtype=torch.bfloat16):
            ref_bf16 = vae.decode(z).float()
    e_or, e_bf = relerr(orc, ref), relerr(ref_bf16, ref)
    print(f"vqvae_decode: oracle-vs-ref {e_or:.2e}  ref bf16-autocast-vs-fp32 {e_bf:.2e}")
    ok &= e_or < 1e-5
    torch.save({"config": acfg, "seed": 1111, "inputs": {"z": z}, "output": ref.to(torch.float16),
                "ref_bf16_autocast_relerr": e_bf, "ref_bf16_autocast_l2err": l2err(ref_bf16, ref)},
               os.path.join(GOLD, "vqvae_decode.pt"))

    # scheduler: outputs of the reference class itself
    sys.path.insert(0, REF)
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler as RefSched
    sys.path.pop(0)
    import numpy as np
    fx = {}
    for T, b0, b1 in [(1000, 0.00085, 0.012), (1000, 0.0015, 0.0195)]:
        s = RefSched(T, b0, b1)
        g = torch.Generator().manual_seed(T + int(b0 * 1e6))
        x, n = torch.randn(4, 4, 32, 32, generator=g), torch.randn(4, 4, 32, 32, generator=g)
        t = torch.tensor([0, 1, 500, T - 1])
        key = f"T{T}_b{b0}"
        fx[key + "_x"], fx[key + "_n"], fx[key + "_t"] = x.numpy(), n.numpy(), t.numpy()
        fx[key + "_add_noise"] = s.add_noise(x, n, t).numpy()
        for step in [T - 1, 500, 1, 0]:
            torch.manual_seed(step)
            xp, x0 = s.sample_prev_timestep(x, n, torch.as_tensor(step))
            torch.manual_seed(step)
            z = torch.randn(x.shape)
            fx[f"{key}_prev{step}_z"], fx[f"{key}_prev{step}_xp"], fx[f"{key}_prev{step}_x0"] = \
                z.numpy(), xp.numpy(), x0.numpy()
        fx[key + "_betas"] = s.betas.numpy()
        fx[key + "_acp"] = s.alpha_cum_prod.numpy()
    np.savez_compressed(os.path.join(GOLD, "scheduler.npz"), **fx)
    print("scheduler fixtures written")
    ok &= gen_dit()
    ok &= gen_cim()
    ok &= gen_samplers()
    ok &= gen_vqvae_encode()
    if not ok:
        print("ORACLE DISAGREES WITH THE REFERENCE")
        sys.exit(1)


if __name__ == "__main__":
    main()

"""The oracle restatements against the committed outputs of the REAL reference (tests/golden, made by
oracle/gen_golden.py in the build container). CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import scheduler_np as onp
from oracle import unet_torch as O
from util import GOLD, cond_of, load_golden, relerr


def test_scheduler_oracle_bit_exact_vs_reference_outputs():
    fx = np.load(os.path.join(GOLD, "scheduler.npz"))
    for T, b0, b1 in [(1000, 0.00085, 0.012), (1000, 0.0015, 0.0195)]:
        key = f"T{T}_b{b0}"
        tb = onp.make_tables(T, b0, b1)
        assert np.array_equal(tb["betas"], fx[key + "_betas"])
        assert np.array_equal(tb["alpha_cum_prod"], fx[key + "_acp"])
        x, n, t = fx[key + "_x"], fx[key + "_n"], fx[key + "_t"]
        assert np.array_equal(onp.add_noise(tb, x, n, t), fx[key + "_add_noise"])
        for step in [T - 1, 500, 1, 0]:
            xp, x0 = onp.sample_prev_timestep(tb, x, n, step, fx[f"{key}_prev{step}_z"])
            assert np.array_equal(xp, fx[f"{key}_prev{step}_xp"]), step
            assert np.array_equal(x0, fx[f"{key}_prev{step}_x0"]), step


def _product_cls(name, relu):
    if "uncond" in name:
        from models.unet_base import Unet
    elif relu:
        from models.unet_cond_base_relu import Unet
    else:
        from models.unet_cond_base import Unet
    return Unet


@pytest.mark.parametrize("name", ["unet_small_cond", "unet_small_uncond", "unet_small_cond_relu", "unet_full_cond"])
def test_unet_oracle_vs_reference_golden(name):
    """Weights come from constructing the PRODUCT module under the fixture seed — so this also proves the product's
    parameter creation order equals the reference's (any difference would change every weight)."""
    fx = load_golden(name)
    torch.manual_seed(fx["seed"])
    model = _product_cls(name, fx["relu"])(4, fx["config"])
    inp = fx["inputs"]
    with torch.no_grad():
        out = O.unet_forward(model.state_dict(), fx["config"], inp["x"], inp["t"], cond_of(inp, fx["config"]),
                             relu=fx["relu"])
    assert relerr(out, fx["output"]) < 1e-5  # fp32 vs fp32: summation-order noise only


def test_time_embedding_oracles_agree():
    t = torch.tensor([0, 3, 999])
    a = O.time_embedding(t, 512).numpy()
    b = onp.time_embedding(t.numpy(), 512)
    assert np.abs(a - b).max() < 1e-4


@pytest.mark.parametrize("name", ["dit_small_cond", "dit_small_plain", "dit_9l_text_image", "dit_12l_image"])
def test_dit_oracle_vs_reference_golden(name):
    """DiT restatement against the committed outputs of the real reference DIT; weights from constructing the
    PRODUCT module under the fixture seed (proves identical parameter creation order and initialisation)."""
    from models.transformer import DIT
    from oracle import dit_torch as OD
    from util import randomise_zero_init
    fx = load_golden(name)
    torch.manual_seed(fx["seed"])
    model = DIT(4, fx["config"])
    randomise_zero_init(model, fx["rand_seed"])
    inp = fx["inputs"]
    cond = cond_of(inp, fx["config"])
    with torch.no_grad():
        out = OD.dit_forward(model.state_dict(), fx["config"], inp["x"], inp["t"], cond)
        out_s = OD.dit_forward(model.state_dict(), fx["config"], inp["x"], torch.as_tensor(int(inp["t"][0])), cond)
    assert relerr(out, fx["output"]) < 1e-5
    assert relerr(out_s, fx["output_scalar_t"]) < 1e-5
    assert fx["output"].abs().max() > 0.1  # not the vacuous all-zero default-init output


def test_cim_oracle_vs_reference_golden():
    """cim_layers restatement against the committed outputs of the real reference layers: outputs of two forwards,
    the lazily initialised step sizes / ADC gain, and (LSQ layers) the parameter gradients — all exactly."""
    from oracle import cim_torch as OC
    fx = load_golden("cim_layers")
    for name, f in fx.items():
        unit = f["unit"].float()
        p = {k: v.clone() for k, v in f["init"].items()}
        if name.startswith("lsq"):
            with torch.no_grad():
                y1 = OC.lsq_layer(p, f["cfg"], f["x"], unit, f["conv"])
            assert torch.equal(y1, f["y1"]), name
            for k, v in f["after"].items():
                assert torch.equal(p[k], v), (name, k)
            p2 = {k: v.clone().requires_grad_(True) for k, v in p.items()}
            y2 = OC.lsq_layer(p2, f["cfg"], f["x2"], unit, f["conv"])
            assert torch.equal(y2.detach(), f["y2"]), name
            ((y2 - f["target"]) ** 2).mean().backward()
            for k, g in f["grads"].items():
                assert relerr(p2[k].grad, g) < 1e-5, (name, k)
        else:
            noise = {}
            if f["cfg"].get("gain_noise_scale"):
                noise = dict(gain_noise=f["gain_noise"], offset_noise=f["offset_noise"])
            with torch.no_grad():
                y1 = OC.adda_linear(p, f["cfg"], f["x"], f["mapping"], unit, **noise)
                for k, v in f["after"].items():
                    assert torch.equal(p[k], v), (name, k)
            assert torch.equal(y1, f["y1"]), name
            p2 = {k: v.clone().requires_grad_(True) for k, v in p.items()}
            x2 = f["x2"].clone().requires_grad_(True)
            y2 = OC.adda_linear(p2, f["cfg"], x2, f["mapping"], unit, **noise)
            assert torch.equal(y2.detach(), f["y2"]), name
            ((y2 - f["target"]) ** 2).mean().backward()
            got = {k: v.grad for k, v in p2.items()}
            got["x"] = x2.grad
            for k, g in f["grads"].items():
                assert relerr(got[k], g) < 1e-5, (name, k)


def test_cim_conv_oracle_vs_reference_golden():
    """k x k Conv2d_lsq_adda_cim (nn.Unfold row order, row-block split of K): the oracle against the committed outputs,
    lazily initialised parameters and gradients of the real reference layer — exactly."""
    from oracle import cim_torch as OC
    for name, f in load_golden("cim_adda_conv").items():
        unit = f["unit"].float()
        p = {k: v.clone() for k, v in f["init"].items()}
        with torch.no_grad():
            y1 = OC.adda_linear(p, f["cfg"], f["x"], f["mapping"], unit, conv=f["conv"])
        for k, v in f["after"].items():
            assert torch.equal(p[k], v), (name, k)
        assert torch.equal(y1, f["y1"]), name
        p2 = {k: v.clone().requires_grad_(True) for k, v in p.items()}
        x2 = f["x2"].clone().requires_grad_(True)
        y2 = OC.adda_linear(p2, f["cfg"], x2, f["mapping"], unit, conv=f["conv"])
        assert torch.equal(y2.detach(), f["y2"]), name
        ((y2 - f["target"]) ** 2).mean().backward()
        got = {k: v.grad for k, v in p2.items()}
        got["x"] = x2.grad
        for k, g in f["grads"].items():
            assert relerr(got[k], g) < 1e-5, (name, k)


def _sampler_model(x, t, cond=None):
    """The stand-in denoiser of oracle/gen_golden.py:gen_samplers (fp32 * and + only: bit-reproducible anywhere)."""
    s = (t.float() * 0.001).view(-1, 1, 1, 1)   # (not "/ 1000": CUDA divides by a scalar through its reciprocal)
    out = x * 0.5 + s * x.flip(-1)
    if cond is not None:
        out = out + cond["image"]
    return out


def test_sampler_oracle_bit_exact_vs_reference_outputs():
    fx = torch.load(os.path.join(GOLD, "samplers.pt"), weights_only=False)
    x_T, cond = fx["x_T"], {"image": fx["cond_image"]}
    for name in ("ddim_linear_eta0", "ddim_quadratic_eta05", "ddim_linear_eta1_trace"):
        c = fx[name]
        kw = c["kw"]
        bar = onp.sampler_alpha_bar(c["beta"], c["T"])[2]
        ts, tp = onp.ddim_timesteps(c["T"], kw["steps"], kw["method"])
        x = x_T.numpy()
        for k, i in enumerate(reversed(range(kw["steps"]))):
            t = torch.full((x.shape[0],), int(ts[i]), dtype=torch.long)
            x = onp.ddim_step(bar, x, _sampler_model(torch.from_numpy(x), t, cond).numpy(), c["z"][k].numpy(),
                              int(ts[i]), int(tp[i]), kw["eta"])
        want = c["out"] if c["out"].dim() == 4 else c["out"][:, -1]
        got = x if c["out"].dim() == 4 else np.clip(x, -1, 1)
        assert np.array_equal(got, want.numpy()), name
    c = fx["ddpm_T40"]
    x = x_T.numpy()
    for k, ts_ in enumerate(reversed(range(c["T"]))):
        t = torch.full((x.shape[0],), ts_, dtype=torch.long)
        x = onp.ddpm_step(c["beta"], c["T"], x, _sampler_model(torch.from_numpy(x), t).numpy(),
                          c["z"][k].numpy() if ts_ > 0 else None, ts_)
    assert np.array_equal(x, c["out"].numpy())


def test_vqvae_encode_oracle_vs_reference_golden():
    from models.vqvae import VQVAE  # only for the seeded weights (same constructor RNG consumption as the reference)
    fx = load_golden("vqvae_encode")
    torch.manual_seed(fx["seed"])
    vae = VQVAE(3, fx["config"])
    with torch.no_grad():
        pre = O.vqvae_encode_latent(vae.state_dict(), fx["config"], fx["inputs"]["x"])
        q, mse, idx, _ = O.vq_quantize(fx["pre_quant"], vae.embedding.weight.detach())
    assert relerr(pre, fx["pre_quant"]) < 1e-5
    assert torch.equal(idx, fx["indices"]) and torch.equal(q, fx["quant_out"])
    torch.testing.assert_close(mse, fx["codebook_loss"], rtol=1e-6, atol=0)


def test_percent_optimizer_oracle_matches_reference_fixture():
    """oracle/optim_torch.percent_step reproduces the reference's PercentOptimizerFP (cim_layers/IBA_optimizer.py:80-143)
    bit for bit on the recorded gradients, including a tensor that skips a step."""
    from oracle import optim_torch as OT
    fx = load_golden("percent_opt")
    p = [t.clone() for t in fx["init"]]
    m, v, steps = [torch.zeros_like(t) for t in p], [torch.zeros_like(t) for t in p], [0] * len(p)
    for n, gs in enumerate(fx["grads"], 1):
        for k, g in enumerate(gs):
            steps[k] += g is not None
        for st in sorted(set(steps)):
            sel = [k for k in range(len(p)) if steps[k] == st and gs[k] is not None]
            OT.percent_step([p[k] for k in sel], [gs[k] for k in sel], [m[k] for k in sel], [v[k] for k in sel], st,
                            fx["lr"], fx["betas"], fx["eps"])
        if n in fx["after"]:
            assert all(torch.equal(a, b) for a, b in zip(p, fx["after"][n]))
    assert all(torch.equal(a, b) for a, b in zip(m, fx["exp_avg"]))
    assert all(torch.equal(a, b) for a, b in zip(v, fx["exp_avg_sq"]))
    assert steps == [int(s) for s in fx["steps"]]

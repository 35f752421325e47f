"""Sampling-path parity on the GPU: VQVAE.decode vs the reference's golden output (a14), hoisted conditioning and
CUDA-graph replay vs the plain per-call path (must be bit-identical: same kernels, same order), and the
device-resident sampling loop vs the drop-in loop with the scheduler oracle in between."""
import numpy as np
import pytest
import torch

from util import cond_of, l2err, load_golden, relerr

pytestmark = pytest.mark.gpu


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def test_vqvae_decode_vs_reference_golden():
    dev = _dev()
    from models.vqvae import VQVAE
    fx = load_golden("vqvae_decode")
    torch.manual_seed(fx["seed"])
    vae = VQVAE(3, fx["config"]).eval().to(dev)
    with torch.no_grad():
        out = vae.decode(fx["inputs"]["z"].to(dev))
    ref = fx["output"].float()
    err, l2 = relerr(out.cpu(), ref), l2err(out.cpu(), ref)
    print(f"vqvae decode: max-norm {err:.3e} L2 {l2:.3e} (reference bf16 autocast {fx['ref_bf16_autocast_relerr']:.3e} / "
          f"{fx['ref_bf16_autocast_l2err']:.3e})")
    assert out.shape == ref.shape and out.dtype == torch.float32
    # gate: never less accurate than the reference's own bf16-autocast decode of the same latent (4.0e-2 / 3.5e-2 here;
    # the decoder's 128^2 / 256^2 stages are long chains of bf16-rounded activations), capped at 5e-2
    assert err <= min(5e-2, fx["ref_bf16_autocast_relerr"])
    assert l2 <= min(5e-2, fx["ref_bf16_autocast_l2err"])


def test_vqvae_quantize_vs_reference_golden():
    """Codebook search on the reference's own pre-quantisation latent: same indices (the reference's cdist may pick
    another code only between numerically tied ones), bit-identical straight-through output, same loss value."""
    dev = _dev()
    from models.vqvae import VQVAE
    from oracle import unet_torch as O
    fx = load_golden("vqvae_encode")
    torch.manual_seed(fx["seed"])
    vae = VQVAE(3, fx["config"]).eval().to(dev)
    pre = fx["pre_quant"]
    with torch.no_grad():
        zq, losses, idx = vae.quantize(pre.to(dev))
    assert idx.dtype == torch.int64 and idx.shape == fx["indices"].shape and zq.shape == fx["quant_out"].shape
    same = idx.cpu() == fx["indices"]
    if not same.all():
        _, _, _, d2 = O.vq_quantize(pre, vae.embedding.weight.detach().cpu())
        rows = torch.arange(d2.shape[0])
        ours, ref = d2[rows, idx.cpu().reshape(-1)], d2[rows, fx["indices"].reshape(-1)]
        assert ((ours - ref).abs() / ref.clamp_min(1e-30)).max() < 1e-5
    assert same.float().mean() > 0.999
    m = same[:, None].expand_as(zq)
    assert torch.equal(zq.cpu()[m], fx["quant_out"][m])
    for k in ("codebook_loss", "commitment_loss"):
        torch.testing.assert_close(losses[k].cpu(), fx[k], rtol=1e-5, atol=0)


def test_vqvae_encode_vs_reference_golden():
    dev = _dev()
    from models.vqvae import VQVAE
    fx = load_golden("vqvae_encode")
    torch.manual_seed(fx["seed"])
    vae = VQVAE(3, fx["config"]).eval().to(dev)
    x = fx["inputs"]["x"].to(dev)
    with torch.no_grad():
        lat = vae.encode_latent(x)
        zq, losses = vae.encode(x)
        zq2, losses2, idx = vae.quantize(lat)
        out, z, ql = vae(x)
    ref = fx["pre_quant"]
    err, l2 = relerr(lat.cpu(), ref), l2err(lat.cpu(), ref)
    print(f"vqvae encoder latent: max-norm {err:.3e} L2 {l2:.3e} (reference bf16 autocast "
          f"{fx['ref_bf16_autocast_relerr']:.3e} / {fx['ref_bf16_autocast_l2err']:.3e})")
    # gate: never less accurate than the reference's own bf16-autocast encoder on the same image, capped at 5e-2
    assert err <= min(5e-2, fx["ref_bf16_autocast_relerr"]) and l2 <= min(5e-2, fx["ref_bf16_autocast_l2err"])
    assert torch.equal(zq, zq2) and torch.equal(losses["codebook_loss"], losses2["codebook_loss"])
    assert zq.shape == fx["quant_out"].shape and idx.shape == fx["indices"].shape
    assert out.shape == x.shape and torch.equal(z, zq) and set(ql) == {"codebook_loss", "commitment_loss"}
    # every output vector is (numerically) a codebook row
    cb = vae.embedding.weight.detach()
    pick = cb[idx.reshape(-1)].reshape(*idx.shape, -1).permute(0, 3, 1, 2)
    torch.testing.assert_close(zq, pick, rtol=1e-6, atol=1e-6)


def test_vqvae_has_no_cpu_path_and_no_backward():
    _dev()
    from models.vqvae import VQVAE
    from sdb200.runtime import B200Error
    fx = load_golden("vqvae_decode")
    vae = VQVAE(3, fx["config"])
    with pytest.raises(B200Error):
        vae.encode(torch.zeros(1, 3, 64, 64))  # CPU tensor
    with pytest.raises(B200Error):
        vae.decode(torch.zeros(1, 4, 8, 8))  # CPU tensor
    vae = vae.cuda()
    with pytest.raises(B200Error):
        vae.encode(torch.zeros(1, 3, 64, 64, device="cuda"))  # autograd recording: stage-1 training is out of scope


def _small_cond(dev):
    from models.unet_cond_base import Unet
    fx = load_golden("unet_small_cond")
    torch.manual_seed(fx["seed"])
    return fx, Unet(4, fx["config"]).eval().to(dev)


def test_graph_and_hoisting_are_bit_identical_to_plain_calls():
    dev = _dev()
    from sdb200 import engine as E
    from sdb200 import graphs as G
    fx, model = _small_cond(dev)
    inp, cfg = fx["inputs"], fx["config"]
    cond = cond_of(inp, cfg, dev)
    x, t = inp["x"].to(dev), inp["t"].to(dev)
    with torch.no_grad():
        G.ENABLE_FORWARD_GRAPHS, E.HOIST_CONDITIONING = False, False
        try:
            plain = model(x, t, cond)
            E.HOIST_CONDITIONING = True
            hoist1 = model(x, t, cond)
            hoist2 = model(x, t, cond)  # second call runs on the cached K/V and mask channels
            G.ENABLE_FORWARD_GRAPHS = True
            g1 = model(x, t, cond)      # capture
            g2 = model(x, t, cond)      # replay
            g3 = model(x * 0.5, t, cond)
            G.ENABLE_FORWARD_GRAPHS = False
            p3 = model(x * 0.5, t, cond)
        finally:
            G.ENABLE_FORWARD_GRAPHS, E.HOIST_CONDITIONING = True, True
    for other in (hoist1, hoist2, g1, g2):
        assert torch.equal(plain, other)
    assert torch.equal(g3, p3) and not torch.equal(g3, plain)


def test_condition_cache_sees_in_place_edits_and_weight_updates():
    dev = _dev()
    fx, model = _small_cond(dev)
    inp, cfg = fx["inputs"], fx["config"]
    cond = cond_of(inp, cfg, dev)
    x, t = inp["x"].to(dev), inp["t"].to(dev)
    with torch.no_grad():
        a = model(x, t, cond)
        cond["text"].mul_(0.5)                    # same storage, new version
        b = model(x, t, cond)
        model.downs[0].context_proj[0].weight.mul_(1.5)
        c = model(x, t, cond)
    assert not torch.equal(a, b) and not torch.equal(b, c)


def test_condition_cache_is_not_fooled_by_a_recycled_address():
    """The next batch's condition tensors usually get the ADDRESS of the previous batch's (caching allocator, same shape,
    _version 0 again — batch_condition_image_generator.py builds them with torch.cat(...).to(device) per batch). The
    hoisted state and the captured forward must follow the tensors' identity, not their address."""
    dev = _dev()
    from sdb200.graphs import SamplingPipeline
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    fx, model = _small_cond(dev)
    inp, cfg = fx["inputs"], fx["config"]
    x, t = inp["x"].to(dev), inp["t"].to(dev)
    host = cond_of(inp, cfg)
    pipe = SamplingPipeline(model, LinearNoiseScheduler(1000, 0.00085, 0.012), None)
    z = {i: torch.zeros_like(x) for i in range(1000)}
    with torch.no_grad():
        cond = {k: v.to(dev) for k, v in host.items()}
        ptrs = {k: v.data_ptr() for k, v in cond.items()}
        a = model(x, t, cond)
        pa, _, _ = pipe.sample(x, cond, t_start=999, num_steps=2, z_fn=lambda i: z[i], decode=False)
        del cond
        cond2 = {k: (v.flip(0) * 0.5).to(dev) for k, v in host.items()}     # same shapes, other contents
        recycled = all(cond2[k].data_ptr() == ptrs[k] for k in cond2)
        b = model(x, t, cond2)
        pb, _, _ = pipe.sample(x, cond2, t_start=999, num_steps=2, z_fn=lambda i: z[i], decode=False)
        model.invalidate_plan()
        want = model(x, t, cond2)
    print("condition tensors re-used the freed addresses:", recycled)
    assert not torch.equal(a, b) and torch.equal(b, want)
    assert not torch.equal(pa, pb)


def test_a_new_condition_reuses_the_captured_forward():
    """A sampling job with fresh condition tensors (the previous job's are gone) must not re-capture the forward: the
    old CondState lends its storage (engine.rebind_cond: mask channels, text rows, per-site K/V recomputed), the graph
    captured on it is replayed — results bit-identical to a freshly planned model, for the conditional and the
    unconditional (CFG) state of two successive jobs."""
    dev = _dev()
    fx, model = _small_cond(dev)
    inp, cfg = fx["inputs"], fx["config"]
    x, t = inp["x"].to(dev), inp["t"].to(dev)
    host = cond_of(inp, cfg)
    with torch.no_grad():
        cond = {k: v.to(dev) for k, v in host.items()}
        unc = {k: torch.zeros_like(v) for k, v in cond.items()}
        model(x, t, cond), model(x, t, unc)
        graphs1 = dict(model.__dict__["_graphed"])
        assert len(graphs1) == 2
        del cond, unc
        cond2 = {k: (v.flip(0) * 0.5).to(dev) for k, v in host.items()}
        unc2 = {k: torch.full_like(v, 0.25) for k, v in cond2.items()}
        b, bu = model(x, t, cond2), model(x, t, unc2)
        graphs2 = model.__dict__["_graphed"]
        assert len(graphs2) == 2 and all(graphs2[k] is graphs1[k] for k in graphs1)      # replayed, not re-captured
        b_again = model(x, t, cond2)                                                    # plain cache hit afterwards
        model.invalidate_plan()
        want, want_u = model(x, t, cond2), model(x, t, unc2)
    assert torch.equal(b, want) and torch.equal(bu, want_u) and torch.equal(b_again, want)
    assert not torch.equal(want, want_u)


@pytest.mark.parametrize("cfg_scale", [1.0, 3.0])
def test_sampling_pipeline_matches_drop_in_loop(cfg_scale):
    """Device-resident loop (graph, t on device, in-place x update, fused CFG) vs the reference-shaped Python loop
    built from the drop-in modules, same injected z: bit-identical latents after 6 timesteps incl. t = 0."""
    dev = _dev()
    from oracle import scheduler_np as onp
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    from sdb200.graphs import SamplingPipeline
    fx, model = _small_cond(dev)
    inp, cfg = fx["inputs"], fx["config"]
    cond = cond_of(inp, cfg, dev)
    uncond = {k: torch.zeros_like(v) for k, v in cond.items()}
    sched = LinearNoiseScheduler(1000, 0.00085, 0.012)
    x = inp["x"].to(dev)
    n = x.shape[0]
    g = torch.Generator().manual_seed(9)
    zs = {i: torch.randn(x.shape, generator=g).to(dev) for i in range(6)}
    pipe = SamplingPipeline(model, sched, None)
    xf, x0f, _ = pipe.sample(x, cond, uncond if cfg_scale > 1 else None, cfg_scale, t_start=5, z_fn=lambda i: zs[i])
    tb = onp.make_tables(1000, 0.00085, 0.012)
    xt = x.clone()
    with torch.no_grad():
        for i in reversed(range(6)):
            t = torch.full((n,), i, device=dev, dtype=torch.long)
            eps = model(xt, t, cond)
            if cfg_scale > 1:
                eu = model(xt, t, uncond)
                eps_np = (eu + cfg_scale * (eps - eu)).cpu().numpy()  # tools/sample_ddpm_text_image_cond.py:86
                xt2, x0 = sched.sample_prev_timestep(xt, eps, i, z=zs[i], noise_pred_uncond=eu,
                                                     cf_guidance_scale=cfg_scale)
            else:
                eps_np = eps.cpu().numpy()
                xt2, x0 = sched.sample_prev_timestep(xt, eps, torch.as_tensor(i), z=zs[i])
            # the scheduler oracle (numpy restatement of the reference) on the same eps
            rp, r0 = onp.sample_prev_timestep(tb, xt.cpu().numpy(), eps_np, i, zs[i].cpu().numpy())
            assert np.array_equal(xt2.cpu().numpy(), rp) and np.array_equal(x0.cpu().numpy(), r0)
            xt = xt2
    assert torch.equal(xf, xt) and torch.equal(x0f, x0)


def test_in_kernel_noise_is_fresh_per_step_under_graph_replay():
    dev = _dev()
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    from sdb200.graphs import SamplingPipeline
    fx, model = _small_cond(dev)
    cond = cond_of(fx["inputs"], fx["config"], dev)
    sched = LinearNoiseScheduler(1000, 0.00085, 0.012)
    pipe = SamplingPipeline(model, sched, None)
    x = fx["inputs"]["x"].to(dev)
    a, _, _ = pipe.sample(x, cond, t_start=999, num_steps=3, decode=False, noise_call=7)
    b, _, _ = pipe.sample(x, cond, t_start=999, num_steps=3, decode=False, noise_call=7)
    assert torch.isfinite(a).all() and torch.equal(a, b)  # same seed, same call index, same t sequence -> reproducible
    c, _, _ = pipe.sample(x, cond, t_start=999, num_steps=3, decode=False)
    d, _, _ = pipe.sample(x, cond, t_start=999, num_steps=3, decode=False)
    assert not torch.equal(c, d) and not torch.equal(a, c)   # successive calls draw fresh noise trajectories
    sched.seed = 12345                                        # a changed seed takes effect without a new plan
    plan_before = pipe._plan
    e, _, _ = pipe.sample(x, cond, t_start=999, num_steps=3, decode=False, noise_call=7)
    assert pipe._plan is plan_before and not torch.equal(a, e)
    sched.seed = 0
    # statistics of the drawn noise: x_{t-1} - mean is sigma*z with z ~ N(0,1); compare two consecutive steps
    P = pipe.plan(x, cond)
    P.xt.copy_(x); P.t.fill_(999); pipe.step(P); x1 = P.xt.clone(); pipe.step(P); x2 = P.xt.clone()
    assert not torch.equal(x1 - x, x2 - x1)


def test_unconditional_drop_in_loop_replays_one_captured_forward():
    """tools/sample_ddpm_vqvae.py:29-34 calls model(xt, t) once per timestep: after the first call every step must be a
    replay of the same captured forward (a regression re-captured the graph on every call of an unconditional model)."""
    dev = _dev()
    from models.unet_base import Unet
    fx = load_golden("unet_small_uncond")
    torch.manual_seed(fx["seed"])
    model = Unet(4, fx["config"]).eval().to(dev)
    x = fx["inputs"]["x"].to(dev)
    with torch.no_grad():
        model(x, torch.as_tensor(9).unsqueeze(0).to(dev))
        graphs = dict(model.__dict__["_graphed"])
        assert len(graphs) == 1
        for i in reversed(range(9)):
            model(x, torch.as_tensor(i).unsqueeze(0).to(dev))
    after = model.__dict__["_graphed"]
    assert len(after) == 1 and all(after[k] is graphs[k] for k in graphs)

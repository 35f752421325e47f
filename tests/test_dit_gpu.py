"""DiT path on the B200: the token-row kernels against fp32 torch, the drop-in `DIT` against the
committed outputs of the real reference (tests/golden/dit_*.pt, weights rebuilt from the fixture seed with the
zero-initialised adaLN / proj_out parameters re-randomised exactly as the generator did), gradients against torch
autograd through the fp32 oracle, and graph replay / conditioning hoist consistency."""
import pytest
import torch

from util import cond_of, l2err, load_golden, randomise_zero_init, relerr

pytestmark = pytest.mark.gpu
NET_CAP = 2.5e-2


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def _bf(t):
    return t.to(torch.bfloat16)


# --------------------------------------------------------------------------------------------------- row kernels
@pytest.mark.parametrize("n,L,c,mod", [(3, 256, 288, True), (2, 64, 96, True), (2, 50, 1024, True), (4, 77, 288, False),
                                       (1, 1, 8, True)])
def test_layernorm_mod_forward_backward(n, L, c, mod):
    dev = _dev()
    from sdb200 import runtime as rt
    g = torch.Generator().manual_seed(n * 1000 + L + c)
    x = _bf(torch.randn(n * L, c, generator=g) * 2 + 0.5).to(dev)
    big = _bf(torch.randn(n, 3 * c + 8, generator=g) * 0.3).to(dev)      # shift | scale as slices of a wider parent
    shift, scale = (big[:, 8:8 + c], big[:, 8 + c:8 + 2 * c]) if mod else (None, None)
    dy = _bf(torch.randn(n * L, c, generator=g)).to(dev)
    y, stats = rt.layernorm_mod(x, n, L, c, shift=shift, scale=scale, eps=1e-6, want_stats=True)
    y2 = rt.layernorm_mod(x, n, L, c, shift=shift, scale=scale, eps=1e-6)
    assert torch.equal(y, y2)

    xf = x.float().reshape(n, L, c).requires_grad_(True)
    shf = shift.float().clone().requires_grad_(True) if mod else None
    scf = scale.float().clone().requires_grad_(True) if mod else None
    ref = torch.nn.functional.layer_norm(xf, (c,), None, None, 1e-6)
    if mod:
        ref = ref * (1 + scf[:, None]) + shf[:, None]
    ref.backward(dy.float().reshape(n, L, c))
    assert relerr(y.float(), ref.detach().reshape(n * L, c)) < 6e-3          # bf16 output rounding
    mean = x.float().mean(-1)
    assert (stats[:, 0] - mean).abs().max() < 1e-4 * (1 + mean.abs().max())

    dx, dsh, dsc = rt.layernorm_mod_bwd(dy, x, stats, n, L, c, scale=scale)
    assert relerr(dx.float(), xf.grad.reshape(n * L, c)) < 6e-3
    if mod:
        assert relerr(dsh, shf.grad) < 1e-4 and relerr(dsc, scf.grad) < 1e-3    # fp32 fixed-order sums
        dx_b, dsh_b, _ = rt.layernorm_mod_bwd(dy, x, stats, n, L, c, scale=scale)
        assert torch.equal(dx, dx_b) and torch.equal(dsh, dsh_b)              # deterministic
    else:
        assert dsh is None and dsc is None
    # fused second-consumer gradient
    dres = _bf(torch.randn(n * L, c, generator=g)).to(dev)
    dx2, _, _ = rt.layernorm_mod_bwd(dy, x, stats, n, L, c, scale=scale, dres=dres)
    assert relerr(dx2.float(), xf.grad.reshape(n * L, c) + dres.float()) < 6e-3


@pytest.mark.parametrize("n,L,c", [(3, 256, 288), (2, 33, 96)])
def test_gate_residual_and_relu_backward(n, L, c):
    dev = _dev()
    from sdb200 import runtime as rt
    g = torch.Generator().manual_seed(L)
    x, h, dy = (_bf(torch.randn(n * L, c, generator=g)).to(dev) for _ in range(3))
    gate = _bf(torch.randn(n, 2 * c, generator=g)).to(dev)[:, c:]
    y = rt.gate_residual(x, h, gate, n, L, c)
    ref = x.float().reshape(n, L, c) + gate.float()[:, None] * h.float().reshape(n, L, c)
    assert relerr(y.float(), ref.reshape(n * L, c)) < 6e-3
    dh, dgate = rt.gate_residual_bwd(dy, h, gate, n, L, c)
    assert relerr(dh.float(), (gate.float()[:, None] * dy.float().reshape(n, L, c)).reshape(n * L, c)) < 6e-3
    assert relerr(dgate, (dy.float() * h.float()).reshape(n, L, c).sum(1)) < 1e-4
    dr = rt.relu_bwd(dy, h)
    assert torch.equal(dr, torch.where(h > 0, dy, torch.zeros_like(dy)))


def test_patchify_unpatchify_match_einops():
    dev = _dev()
    from einops import rearrange
    from sdb200 import runtime as rt
    g = torch.Generator().manual_seed(2)
    n, ch, gh, gw, p = 3, 4, 16, 8, 2
    tok = torch.randn(n * gh * gw, p * p * ch, generator=g).to(dev)
    want = rearrange(tok.reshape(n, gh * gw, -1), 'b (nh nw) (ph pw c) -> b c (nh ph) (nw pw)', ph=p, pw=p, nw=gw, nh=gh)
    assert torch.equal(rt.unpatchify(tok, n, ch, gh, gw, p, p), want)
    assert torch.equal(rt.unpatchify(_bf(tok), n, ch, gh, gw, p, p), _bf(want).float())
    img = torch.randn(n, ch, gh * p, gw * p, generator=g).to(dev)
    back = rt.patchify(img, p, p, ld=24)
    want_t = rearrange(img, 'b c (nh ph) (nw pw) -> b (nh nw) (ph pw c)', ph=p, pw=p).reshape(n * gh * gw, -1)
    assert torch.equal(back[:, :16], _bf(want_t)) and back[:, 16:].abs().max() == 0


def test_patch_embedding_module_vs_torch():
    dev = _dev()
    from einops import rearrange
    from models.patch_embed import PatchEmbedding, get_patch_position_embedding
    torch.manual_seed(4)
    pe = PatchEmbedding(32, 32, 7, 2, 2, 288).to(dev)
    x = torch.randn(2, 7, 32, 32, device=dev)
    out = pe(x)
    tok = rearrange(x, 'b c (nh ph) (nw pw) -> b (nh nw) (ph pw c)', ph=2, pw=2)
    ref = torch.nn.functional.linear(tok, pe.patch_embed[0].weight, pe.patch_embed[0].bias) \
        + get_patch_position_embedding(288, (16, 16), dev)
    assert out.shape == ref.shape
    assert relerr(out, ref.detach()) < 6e-3


# --------------------------------------------------------------------------------------------------- whole model
def _build(name, dev=None, train=False):
    from models.transformer import DIT
    fx = load_golden(name)
    torch.manual_seed(fx["seed"])
    model = DIT(4, fx["config"])
    randomise_zero_init(model, fx["rand_seed"])
    model = model.train() if train else model.eval()
    return fx, (model.to(dev) if dev is not None else model)


@pytest.mark.parametrize("name", ["dit_small_cond", "dit_small_plain", "dit_9l_text_image", "dit_12l_image"])
def test_dit_forward_vs_reference_golden(name):
    dev = _dev()
    fx, model = _build(name, dev)
    inp, cfg = fx["inputs"], fx["config"]
    cond = cond_of(inp, cfg, dev)
    extra = (cond,) if cond is not None else ()
    with torch.no_grad():
        out = model(inp["x"].to(dev), inp["t"].to(dev), *extra)
        out_again = model(inp["x"].to(dev), inp["t"].to(dev), *extra)          # graph replay
        out_s = model(inp["x"].to(dev), torch.as_tensor(int(inp["t"][0])), *extra)   # 0-dim t shared by the batch
    torch.cuda.synchronize()
    err, l2 = relerr(out.cpu(), fx["output"]), l2err(out.cpu(), fx["output"])
    print(f"{name}: ours-vs-reference-fp32 max-norm {err:.3e} L2 {l2:.3e} (reference bf16 autocast: "
          f"{fx['ref_bf16_autocast_relerr']:.3e} / {fx['ref_bf16_autocast_l2err']:.3e})")
    assert out.shape == fx["output"].shape and out.dtype == torch.float32
    assert torch.equal(out, out_again)
    # same gate as the UNet (tests/test_parity_gpu.py): never less accurate than the reference's own bf16 autocast
    assert err <= min(NET_CAP, 1.25 * fx["ref_bf16_autocast_relerr"])
    assert l2 <= min(NET_CAP, fx["ref_bf16_autocast_l2err"])
    assert relerr(out_s.cpu(), fx["output_scalar_t"]) <= min(NET_CAP, 1.25 * fx["ref_bf16_autocast_relerr"])


def test_dit_graph_and_hoisting_are_bit_identical_to_eager():
    dev = _dev()
    from sdb200 import engine as E
    from sdb200 import graphs as G
    fx, model = _build("dit_small_cond", dev)
    inp, cfg = fx["inputs"], fx["config"]
    cond = cond_of(inp, cfg, dev)
    x, t = inp["x"].to(dev), inp["t"].to(dev)
    with torch.no_grad():
        a = model(x, t, cond)
        G.ENABLE_FORWARD_GRAPHS = False
        E.HOIST_CONDITIONING = False
        E.BATCH_TEMB = False
        try:
            model.invalidate_plan()
            b = model(x, t, cond)
        finally:
            G.ENABLE_FORWARD_GRAPHS = True
            E.HOIST_CONDITIONING = True
            E.BATCH_TEMB = True
            model.invalidate_plan()
    assert torch.equal(a, b)


def test_dit_sampling_pipeline_matches_dropin_loop():
    dev = _dev()
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    from sdb200.graphs import SamplingPipeline
    fx, model = _build("dit_small_cond", dev)
    inp, cfg = fx["inputs"], fx["config"]
    cond = cond_of(inp, cfg, dev)
    sched = LinearNoiseScheduler(1000, 0.00085, 0.012)
    xt = inp["x"].to(dev)
    zs = {i: torch.randn(xt.shape, generator=torch.Generator().manual_seed(i)).to(dev) for i in range(995, 1000)}
    with torch.no_grad():
        x = xt.clone()
        for i in reversed(range(995, 1000)):
            eps = model(x, torch.as_tensor(i).unsqueeze(0).to(dev), cond)
            x, _ = sched.sample_prev_timestep(x, eps, torch.as_tensor(i).to(dev), z=zs[i])
        pipe = SamplingPipeline(model, sched)
        got, _, _ = pipe.sample(xt, cond, t_start=999, num_steps=5, z_fn=lambda i: zs[i], decode=False)
    assert torch.equal(got, x)


def _oracle_grads(model, cfg, x, t, cond, target, bf16=False):
    from oracle import dit_torch as OD
    sd = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in model.state_dict().items()}
    with torch.autocast("cpu", dtype=torch.bfloat16, enabled=bf16):
        out = OD.dit_forward(sd, cfg, x, t, cond)
    loss = torch.nn.functional.mse_loss(out.float(), target)
    loss.backward()
    return loss.item(), {k: v.grad for k, v in sd.items() if v.grad is not None}


def _l2(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-20)).item()


@pytest.mark.parametrize("name,graphs", [("dit_small_cond", True), ("dit_small_cond", False), ("dit_small_plain", True)])
def test_dit_gradients_vs_oracle(name, graphs):
    dev = _dev()
    from sdb200 import graphs as G
    fx, model = _build(name, train=True)
    inp, cfg = fx["inputs"], fx["config"]
    target = torch.randn(inp["x"].shape, generator=torch.Generator().manual_seed(11))
    cond_cpu = cond_of(inp, cfg)
    ref_loss, ref = _oracle_grads(model, cfg, inp["x"], inp["t"], cond_cpu, target)
    _, ref16 = _oracle_grads(model, cfg, inp["x"], inp["t"], cond_cpu, target, bf16=True)
    model = model.to(dev)
    cond = cond_of(inp, cfg, dev)
    old = G.ENABLE_TRAIN_GRAPHS
    G.ENABLE_TRAIN_GRAPHS = graphs
    try:
        out = model(inp["x"].to(dev), inp["t"].to(dev), *((cond,) if cond is not None else ()))
        loss = torch.nn.functional.mse_loss(out, target.to(dev))
        loss.backward()
    finally:
        G.ENABLE_TRAIN_GRAPHS = old
    torch.cuda.synchronize()
    assert abs(loss.item() - ref_loss) / ref_loss < 2e-2
    norms = sorted(v.norm().item() for v in ref.values())
    typical = norms[len(norms) // 2]
    worst, yard, missing = [], [], []
    for k, p in model.named_parameters():
        if k not in ref:
            continue
        if p.grad is None:
            missing.append(k)
            continue
        if ref[k].norm().item() < 1e-4 * typical:   # key bias of a softmax: analytically zero
            assert p.grad.norm().item() < 2e-2 * typical, k
            continue
        worst.append((_l2(p.grad.cpu(), ref[k]), k))
        yard.append(_l2(ref16[k], ref[k]))
    assert not missing, f"parameters without gradient: {missing[:5]}"
    worst.sort(reverse=True)
    yard.sort(reverse=True)
    med, ymed = worst[len(worst) // 2][0], yard[len(yard) // 2]
    print(f"{name} graphs={graphs}: loss {loss.item():.5f} (oracle {ref_loss:.5f}); per-parameter rel-L2 vs fp32 oracle: "
          f"worst {worst[0][0]:.3e} ({worst[0][1]}), median {med:.3e}; reference bf16-autocast: worst {yard[0]:.3e}, "
          f"median {ymed:.3e}")
    # same gate as the UNet gradients: <= 1.5x the reference's own bf16-autocast gradient error, floors 8e-2 / 4e-2
    assert worst[0][0] < max(8e-2, 1.5 * yard[0]), worst[:5]
    assert med < max(4e-2, 1.5 * ymed)


def test_dit_training_is_deterministic_and_learns():
    dev = _dev()
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    fx = load_golden("dit_small_cond")
    inp, cfg = fx["inputs"], fx["config"]
    cond = cond_of(inp, cfg, dev)
    sched = LinearNoiseScheduler(1000, 0.00085, 0.012)

    def run():
        from models.transformer import DIT
        torch.manual_seed(fx["seed"])
        model = DIT(4, cfg)
        randomise_zero_init(model, fx["rand_seed"])
        model = model.train().to(dev)
        opt = torch.optim.Adam(model.parameters(), lr=2e-4)
        x0 = inp["x"].to(dev)
        noise = torch.randn(x0.shape, generator=torch.Generator().manual_seed(3)).to(dev)
        t = inp["t"].to(dev)
        losses = []
        for _ in range(8):
            noisy = sched.add_noise(x0, noise, t)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = torch.nn.functional.mse_loss(model(noisy, t, cond), noise)
            opt.zero_grad()
            loss.backward()
            torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
            opt.step()
            losses.append(loss.item())
        return losses

    a, b = run(), run()
    assert a == b, "training is not bit-reproducible"
    assert a[-1] < a[0]

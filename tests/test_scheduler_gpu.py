"""Scheduler parity: CUDA kernels through the product class vs the numpy oracle, BIT-EXACT
(fp32, every op individually rounded), plus timestep-index exactness at the table ends and large-batch properties."""
import numpy as np
import pytest
import torch

from oracle import scheduler_np as onp

pytestmark = pytest.mark.gpu

CFGS = [(1000, 0.00085, 0.012), (1000, 0.0015, 0.0195), (50, 1e-4, 0.02)]


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


@pytest.mark.parametrize("T,b0,b1", CFGS)
@pytest.mark.parametrize("shape", [(4, 4, 32, 32), (64, 4, 32, 32), (3, 5, 7), (1, 1)])
def test_add_noise_bit_exact(T, b0, b1, shape):
    dev = _dev()
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    s = LinearNoiseScheduler(T, b0, b1)
    tb = onp.make_tables(T, b0, b1)
    g = torch.Generator().manual_seed(1111)
    x, n = torch.randn(shape, generator=g), torch.randn(shape, generator=g)
    t = torch.randint(0, T, (shape[0],), generator=g)
    t[0], t[-1] = T - 1, 0
    out = s.add_noise(x.to(dev), n.to(dev), t.to(dev)).cpu().numpy()
    assert np.array_equal(out, onp.add_noise(tb, x.numpy(), n.numpy(), t.numpy()))


@pytest.mark.parametrize("T,b0,b1", CFGS)
@pytest.mark.parametrize("shape", [(4, 4, 32, 32), (2, 3, 5)])
def test_prev_step_bit_exact(T, b0, b1, shape):
    dev = _dev()
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    s = LinearNoiseScheduler(T, b0, b1)
    tb = onp.make_tables(T, b0, b1)
    g = torch.Generator().manual_seed(7)
    x, e, z = (torch.randn(shape, generator=g) for _ in range(3))
    for t in [T - 1, T // 2, 1, 0]:
        for tt in (t, torch.as_tensor(t), torch.as_tensor(t).to(dev)):
            xp, x0 = s.sample_prev_timestep(x.to(dev), e.to(dev), tt, z=z.to(dev))
            rp, r0 = onp.sample_prev_timestep(tb, x.numpy(), e.numpy(), t, z.numpy())
            assert np.array_equal(xp.cpu().numpy(), rp), t
            assert np.array_equal(x0.cpu().numpy(), r0), t


def test_prev_step_cfg_and_rng():
    dev = _dev()
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    s = LinearNoiseScheduler(1000, 0.00085, 0.012)
    tb = onp.make_tables(1000, 0.00085, 0.012)
    g = torch.Generator().manual_seed(9)
    shape = (8, 4, 32, 32)
    x, ec, eu, z = (torch.randn(shape, generator=g) for _ in range(4))
    xp, x0 = s.sample_prev_timestep(x.to(dev), ec.to(dev), 400, z=z.to(dev), noise_pred_uncond=eu.to(dev),
                                    cf_guidance_scale=3.5)
    e = (eu.numpy() + (np.float32(3.5) * (ec.numpy() - eu.numpy()).astype(np.float32)).astype(np.float32))
    rp, r0 = onp.sample_prev_timestep(tb, x.numpy(), e.astype(np.float32), 400, z.numpy())
    assert np.array_equal(xp.cpu().numpy(), rp) and np.array_equal(x0.cpu().numpy(), r0)
    # in-kernel Philox: implied z = (x_prev - mean)/sigma must be standard normal, fresh on every call, and t=0 draws nothing
    big = torch.zeros(256, 4, 32, 32, device=dev)
    xa, _ = s.sample_prev_timestep(big, big, 500)
    xb, _ = s.sample_prev_timestep(big, big, 500)
    sig = float(np.sqrt((1 - tb["alpha_cum_prod"][499]) / (1 - tb["alpha_cum_prod"][500]) * tb["betas"][500]))
    za = (xa / sig).flatten()
    assert abs(za.mean().item()) < 5e-3 and abs(za.std().item() - 1) < 5e-3
    assert abs((za ** 4).mean().item() - 3) < 0.05
    assert not torch.equal(xa, xb)
    x0_, _ = s.sample_prev_timestep(big, big, 0)
    assert x0_.abs().max().item() == 0


def test_scheduler_rejects_cpu_and_bad_t():
    dev = _dev()
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    from sdb200.runtime import B200Error
    s = LinearNoiseScheduler(10, 1e-4, 0.02)
    x = torch.zeros(2, 4)
    with pytest.raises(B200Error):
        s.add_noise(x, x, torch.zeros(2, dtype=torch.long))
    with pytest.raises(B200Error):
        s.sample_prev_timestep(x.to(dev), x.to(dev), 10)
    with pytest.raises(RuntimeError):
        s.add_noise(x.to(dev), x.to(dev), torch.zeros(3, dtype=torch.long))
    empty = torch.zeros(0, 4, device=dev)
    assert s.add_noise(empty, empty, torch.zeros(0, dtype=torch.long)).shape == (0, 4)

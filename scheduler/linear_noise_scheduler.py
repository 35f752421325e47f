"""Drop-in for the reference's scheduler/linear_noise_scheduler.py: same class name, constructor, public table
attributes and method signatures; the arithmetic runs as one fused sm_100a kernel per call
(csrc/sched.cu through the C-ABI in include/b200sd.h).

Differences a caller can observe: none in values when `z` is injected (bit-identical fp32 results); when `z` is
not injected it is drawn on the device with Philox instead of the reference's CPU torch.randn + H2D copy
(scheduler/linear_noise_scheduler.py:72), so the random stream differs. CPU tensors are rejected: this package
has no CPU implementation (the reference itself is the CPU path).
"""
import torch

import sdb200  # noqa: F401  (registers the `sdb200` alias for the hyphenated package directory)
from sdb200 import runtime as _rt


class LinearNoiseScheduler:
    r"""Linear (in sqrt-beta) DDPM noise schedule — reference scheduler/linear_noise_scheduler.py:8-78."""

    def __init__(self, num_timesteps, beta_start, beta_end):
        self.num_timesteps = num_timesteps
        self.beta_start = beta_start
        self.beta_end = beta_end
        # Host tables, built once with the same torch calls as the reference (:18-24) so they are bit-identical;
        # callers and GUIs read these attributes directly.
        self.betas = torch.linspace(beta_start ** 0.5, beta_end ** 0.5, num_timesteps) ** 2
        self.alphas = 1. - self.betas
        self.alpha_cum_prod = torch.cumprod(self.alphas, dim=0)
        self.sqrt_alpha_cum_prod = torch.sqrt(self.alpha_cum_prod)
        self.sqrt_one_minus_alpha_cum_prod = torch.sqrt(1 - self.alpha_cum_prod)
        self._dev = {}          # device -> resident copies (the reference re-uploads them on every call)
        self._philox_offset = 0
        self.seed = 0

    def _tables(self, device):
        key = (device.type, device.index)
        tb = self._dev.get(key)
        if tb is None:
            tb = tuple(t.to(device=device, dtype=torch.float32).contiguous() for t in (
                self.betas, self.alphas, self.alpha_cum_prod, self.sqrt_one_minus_alpha_cum_prod,
                self.sqrt_alpha_cum_prod))
            self._dev[key] = tb
        return tb

    def add_noise(self, original, noise, t):
        r"""q(x_t | x_0): sqrt(acp[t]) * original + sqrt(1 - acp[t]) * noise, t of shape (B,)."""
        if not original.is_cuda:
            raise _rt.B200Error("add_noise: CUDA tensors required (no CPU implementation in this package)")
        betas, alphas, acp, s1m, sa = self._tables(original.device)
        t = torch.as_tensor(t).to(device=original.device, dtype=torch.int64).reshape(-1)
        batch_size = original.shape[0]
        if t.numel() not in (1, batch_size):
            raise RuntimeError(f"shape '[{batch_size}]' is invalid for input of size {t.numel()}")
        return _rt.sched_add_noise(original.float(), noise.to(original.device).float(), t, sa, s1m)

    def sample_prev_timestep(self, xt, noise_pred, t, z=None, noise_pred_uncond=None, cf_guidance_scale=1.0):
        r"""One ancestral step x_t -> (x_{t-1}, x_0 estimate); `t` is one timestep shared by the batch (Python int,
        0-dim CPU tensor as the reference's loops pass, or a CUDA tensor for graph capture).
        Extensions (keyword-only in spirit, defaults reproduce the reference): `z` injects the N(0,1) draw;
        `noise_pred_uncond` + `cf_guidance_scale` fuse the CFG combine of tools/sample_ddpm_text_image_cond.py:86."""
        if not xt.is_cuda:
            raise _rt.B200Error("sample_prev_timestep: CUDA tensors required (no CPU implementation in this package)")
        betas, alphas, acp, s1m, _ = self._tables(xt.device)
        if isinstance(t, torch.Tensor) and not t.is_cuda:
            t = int(t.reshape(-1)[0].item())
        if z is not None:
            z = z.to(device=xt.device, dtype=torch.float32).contiguous()
        if noise_pred_uncond is not None:
            noise_pred_uncond = noise_pred_uncond.float().contiguous()
        off = self._philox_offset
        self._philox_offset += (xt.numel() + 3) // 4 + 4
        return _rt.sched_prev_step(xt.float(), noise_pred.float(), t, (betas, alphas, acp, s1m), z=z,
                                   eps_uncond=noise_pred_uncond, cfg_scale=cf_guidance_scale, seed=self.seed,
                                   offset=off)


# --------------------------------------------------------------------------------------------------------------
# DDPMSampler / DDIMSampler — reference scheduler/linear_noise_scheduler.py:80-256 (the samplers the GUIs and
# batch_condition_image_generator.py:323 drive, usually through a classifier-free-guidance wrapper as `model`).
# Same constructors, buffers, forward() signatures and return values. Per step the reference uploads the alpha table,
# gathers two (B,1,1,1) coefficient tensors and runs ~15 elementwise kernels; here the coefficients of the step are
# computed once on the host with the same fp32 torch operations (so they are the same numbers) and the update is one
# kernel, bit-identical to the reference's result for the same model output and the same z.
import numpy as np
import torch.nn as nn
import torch.nn.functional as F
from tqdm import tqdm


def extract(v, i, shape):
    """v[i] reshaped to (batch, 1, 1, ...) — reference :80-91."""
    out = torch.gather(v, index=i, dim=0)
    out = out.to(device=i.device, dtype=torch.float32)
    return out.view([i.shape[0]] + [1] * (len(shape) - 1))


def _require_cuda(x_t, who):
    if not x_t.is_cuda:
        raise _rt.B200Error(f"{who}: CUDA tensors required (no CPU implementation in this package)")


class DDPMSampler(nn.Module):
    """Ancestral sampler over all T steps with an unconditional `model(x_t, t)` — reference :94-170."""

    NAN_CHECK_EVERY = 50   # the reference syncs the host on every step for its NaN check (:136); NaNs do not heal,
                           # so checking every 50th step and the last raises the same error without the 1000 syncs

    def __init__(self, model, beta, T):
        super().__init__()
        self.model = model
        self.T = T
        self.register_buffer("beta_t", torch.linspace(*beta, T, dtype=torch.float32))
        alpha_t = 1.0 - self.beta_t
        alpha_t_bar = torch.cumprod(alpha_t, dim=0)
        alpha_t_bar_prev = F.pad(alpha_t_bar[:-1], (1, 0), value=1.0)
        self.register_buffer("coeff_1", torch.sqrt(1.0 / alpha_t))
        self.register_buffer("coeff_2", self.coeff_1 * (1.0 - alpha_t) / torch.sqrt(1.0 - alpha_t_bar))
        self.register_buffer("posterior_variance", self.beta_t * (1.0 - alpha_t_bar_prev) / (1.0 - alpha_t_bar))
        self._host = None

    def _coeffs(self, time_step):
        if self._host is None or self._host[0] is not self.coeff_1:
            self._host = (self.coeff_1, self.coeff_1.detach().float().cpu(), self.coeff_2.detach().float().cpu(),
                          torch.sqrt(self.posterior_variance.detach().float().cpu()))
        _, c1, c2, sd = self._host
        return c1[time_step].item(), c2[time_step].item(), sd[time_step].item()

    @torch.no_grad()
    def cal_mean_variance(self, x_t, t):
        epsilon_theta = self.model(x_t, t)
        ts = int(t.reshape(-1)[0].item())
        c1, c2, _ = self._coeffs(ts)
        mean = _rt.sampler_step(x_t.float().contiguous(), epsilon_theta.float().contiguous(), None, c1, -c2, 0.0)
        return mean, extract(self.posterior_variance.to(x_t.device), t, x_t.shape)

    @torch.no_grad()
    def sample_one_step(self, x_t, time_step, z=None):
        _require_cuda(x_t, "DDPMSampler")
        t = torch.full((x_t.shape[0],), time_step, device=x_t.device, dtype=torch.long)
        epsilon_theta = self.model(x_t, t)
        c1, c2, sd = self._coeffs(int(time_step))
        if time_step > 0 and z is None:
            z = torch.randn_like(x_t)
        x_prev = _rt.sampler_step(x_t.float().contiguous(), epsilon_theta.float().contiguous(),
                                  z if time_step > 0 else None, c1, -c2, sd)
        if time_step % self.NAN_CHECK_EVERY == 0 and torch.isnan(x_prev).any().item():
            raise ValueError("nan in tensor!")
        return x_prev

    @torch.no_grad()
    def forward(self, x_t, only_return_x_0=True, interval=1, **kwargs):
        x = [x_t]
        with tqdm(reversed(range(self.T)), colour="#6565b5", total=self.T) as sampling_steps:
            for time_step in sampling_steps:
                x_t = self.sample_one_step(x_t, time_step)
                if not only_return_x_0 and ((self.T - time_step) % interval == 0 or time_step == 0):
                    x.append(torch.clip(x_t, -1.0, 1.0))
                sampling_steps.set_postfix(ordered_dict={"step": time_step + 1, "sample": len(x)})
        if only_return_x_0:
            return x_t
        return torch.stack(x, dim=1)


class DDIMSampler(nn.Module):
    """DDIM sampler (eta = 0) .. DDPM-like (eta = 1) over a linear or quadratic sub-sequence of the T timesteps, with a
    conditional `model(x_t, t, cond_input)` — reference :174-256. `uncond_input` is stored, as in the reference; the
    classifier-free-guidance combine is the job of whatever module is passed as `model`
    (batch_condition_image_generator.py:241-254)."""

    draw_noise_when_eta_is_zero = True   # the reference draws randn_like(x_t) every step even when sigma = 0 (:194);
                                         # keeping the draw keeps torch's CUDA generator in the reference's state

    def __init__(self, model, beta, T):
        super().__init__()
        self.model = model
        self.T = T
        beta_t = torch.linspace(*beta, T, dtype=torch.float32)
        alpha_t = 1.0 - beta_t
        self.alpha_t_bar = torch.cumprod(alpha_t, dim=0)

    def step_coefficients(self, time_step, prev_time_step, eta):
        """(c1, c2, sigma) of x_{t-1} = c1 x_t + c2 eps + sigma z, with the reference's fp32 operations (:186-203)."""
        a = self.alpha_t_bar.detach().float().cpu()
        alpha_t, alpha_t_prev = a[int(time_step)], a[int(prev_time_step)]
        sigma_t = eta * torch.sqrt((1 - alpha_t_prev) / (1 - alpha_t) * (1 - alpha_t / alpha_t_prev))
        c1 = torch.sqrt(alpha_t_prev / alpha_t)
        c2 = torch.sqrt(1 - alpha_t_prev - sigma_t ** 2) - torch.sqrt((alpha_t_prev * (1 - alpha_t)) / alpha_t)
        return c1.item(), c2.item(), sigma_t.item()

    @torch.no_grad()
    def sample_one_step(self, x_t, time_step, prev_time_step, eta, z=None):
        _require_cuda(x_t, "DDIMSampler")
        t = torch.full((x_t.shape[0],), int(time_step), device=x_t.device, dtype=torch.long)
        c1, c2, sigma = self.step_coefficients(time_step, prev_time_step, eta)
        epsilon_theta_t = self.model(x_t, t, self.cond_input)
        if z is None and (sigma != 0.0 or self.draw_noise_when_eta_is_zero):
            z = torch.randn_like(x_t)
        return _rt.sampler_step(x_t.float().contiguous(), epsilon_theta_t.float().contiguous(), z, c1, c2, sigma)

    @torch.no_grad()
    def forward(self, x_t, cond_input, uncond_input, steps=1, method="linear", eta=0.0, only_return_x_0=True,
                interval=1):
        self.cond_input = cond_input
        self.uncond_input = uncond_input
        if method == "linear":
            a = self.T // steps
            time_steps = np.asarray(list(range(0, self.T, a)))
        elif method == "quadratic":
            time_steps = (np.linspace(0, np.sqrt(self.T * 0.8), steps) ** 2).astype(np.int32)
        else:
            raise NotImplementedError(f"sampling method {method} is not implemented!")
        time_steps = time_steps + 1
        time_steps_prev = np.concatenate([[0], time_steps[:-1]])
        x = [x_t]
        with tqdm(reversed(range(0, steps)), colour="#6565b5", total=steps) as sampling_steps:
            for i in sampling_steps:
                x_t = self.sample_one_step(x_t, time_steps[i], time_steps_prev[i], eta)
                if not only_return_x_0 and ((steps - i) % interval == 0 or i == 0):
                    x.append(torch.clip(x_t, -1.0, 1.0))
                sampling_steps.set_postfix(ordered_dict={"step": i + 1, "sample": len(x)})
        if only_return_x_0:
            return x_t
        return torch.stack(x, dim=1)

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

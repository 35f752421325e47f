"""Shared machinery of the cim_layers drop-ins (cim_layers/layers_qn_lsq.py, layers_qn_lsq_adda_cim.py in this repo):

  activations  — quantised by the CUDA kernels of csrc/quant.cu (forward and backward), step sizes read on the device;
  weights      — a few elementwise torch ops on parameter-sized tensors (differentiable, so the straight-through
                 round, the clamp mask, the LSQ step-size gradient and the max-min range of the injected noise follow
                 the reference's autograd graph exactly: cim_layers/quant_noise_utils.py:51-69,92-98,160-179);
  GEMM / conv  — the igemm / wgrad kernels through the same autograd Functions as the float layers.

Reference semantics kept on purpose: lazy step-size initialisation from the first batch while the parameter equals
1.0 (layers_utils_lsq.py:40-42,57-59,74-76) — done on the device, without the reference's host synchronisation.
"""
import math

import torch
from torch.autograd import Function

from . import runtime as rt


def qp_of(bits):
    return float(2 ** (bits - 1) - 1)


class LsqQuantFn(Function):
    """data_quant_lsq on [rows, c] activations; `numel` is the element count the reference's gradient scale sees
    (padding columns excluded)."""

    @staticmethod
    def forward(ctx, x, step, qp, codes, numel):
        ctx.save_for_backward(x, step)
        ctx.cfg = (qp, codes, numel)
        return rt.lsq_quant(x, step, qp, codes=codes, numel=numel)

    @staticmethod
    def backward(ctx, dy):
        x, step = ctx.saved_tensors
        qp, codes, numel = ctx.cfg
        if dy.dtype != torch.bfloat16 or dy.stride(1) != 1 or dy.stride(0) % 8:
            dy = dy.to(torch.bfloat16).contiguous()
        dx, ds = rt.lsq_quant_bwd(dy, x, step, qp, codes=codes, numel=numel)
        return (dx if ctx.needs_input_grad[0] else None), ds.reshape(step.shape), None, None, None


def quantize_rows(x, step, bits, codes=False, numel=None, lazy_init=True):
    """LSQ quantiser on activation rows (bf16 or fp32 [rows, c]); runs the lazy init while `step` may still be 1."""
    qp = qp_of(bits)
    if lazy_init:
        rt.lsq_init_step(x, step, qp)
    numel = x.shape[0] * x.shape[1] if numel is None else numel
    if torch.is_grad_enabled() and (x.requires_grad or step.requires_grad):
        return LsqQuantFn.apply(x, step, qp, codes, numel)
    return rt.lsq_quant(x, step.detach(), qp, codes=codes, numel=numel)


# ---------------------------------------------------------------------------------------------- weight side (torch)
def _grad_scale(x, scale):
    y = x * scale
    return (x - y).detach() + y


def _round_ste(x):
    return (torch.round(x) - x).detach() + x


def init_weight_step_(step, weight, bits):
    """`if step_size_weight == 1: step_size_weight.data = init_step_size(weight, bit)` without the host round trip."""
    with torch.no_grad():
        qp = qp_of(bits)
        rng = weight.detach().abs().max()
        init = 1.0 / (1.0 / rng * qp)
        init = torch.where(rng == 0, torch.ones_like(init), init)
        step.copy_(torch.where(step == 1, init.to(step.dtype), step))


def quantize_weight(weight, step, bits, isint):
    """weight_quant_lsq (quant_noise_utils.py:160-179): returns (quantised weight, w_scale)."""
    qp = qp_of(bits)
    s = _grad_scale(step, 1.0 / math.sqrt(qp * weight.numel()))
    q = _round_ste(torch.clamp(weight / s, -qp, qp))
    if isint:
        return q * s / s.detach(), (1.0 / step).detach()
    return q * s, 1.0


def add_weight_noise(wq, n_scale, unit=None):
    """add_noise (quant_noise_utils.py:92-98): w + (max - min) * n_scale * N(0, 1); `unit` injects the normal draw."""
    if n_scale == 0:
        return wq
    unit = torch.randn_like(wq) if unit is None else unit.to(wq)
    return wq + (wq.max() - wq.min()) * n_scale * unit


def rescale_step_sizes(module, weight_bit_old, input_bit_old, output_bit_old):
    """update_step_size (layers_utils_lsq.py:11-28): a bit-width change by d bits divides the step size by 2^d."""
    for name, new, old in (("weight", module.weight_bit, weight_bit_old), ("input", module.input_bit, input_bit_old),
                           ("output", module.output_bit, output_bit_old)):
        if new != old:
            getattr(module, f"step_size_{name}").data /= 2 ** (new - old)

"""ORACLE — test infrastructure only (imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs; never by the product path).

Plain PyTorch fp32 CPU restatement of the reference's cim_layers arithmetic, functional over an explicit parameter
dict `p` (weight, bias, step_size_input/output/weight, adc_gain) and a config dict `c`:
  LSQ quantisers        cim_layers/quant_noise_utils.py:51-69,92-120,160-179
  lazy step-size init   cim_layers/layers_utils_lsq.py:31-82
  Conv2d/Linear_qn_lsq  cim_layers/layers_qn_lsq.py:98-121,193-216
  ADDA layer            cim_layers/layers_qn_lsq_adda_cim.py:430-522 with layers_utils_adda.py:12-20,40-49,108-185
Differentiable (torch autograd), so gradient parity of the product's backward is taken against it as well.

Pinning: oracle/gen_golden.py runs the REAL reference layers in the build container on seeded inputs (weight noise
injected by monkey-patching torch.randn_like, SURVEY §8c trap iv), asserts this restatement reproduces outputs and
the lazily initialised parameters, and commits them as tests/golden/cim_*.pt.
"""
import math

import torch
import torch.nn.functional as F


def qp_of(bits):
    return 2 ** (bits - 1) - 1


def grad_scale(x, scale):
    y = x * scale
    return (x - y).detach() + y


def round_pass(x):
    return (torch.round(x) - x).detach() + x


def clamp_pass(x, lo, hi):
    return (torch.clamp(x, lo, hi) - x).detach() + x


def init_step(x, bits):
    rng = x.detach().abs().max()
    if rng == 0:
        return torch.tensor(1.0)
    return 1 / (1 / rng * qp_of(bits))


def quant_lsq(x, bits, step, isint):
    qp = qp_of(bits)
    scale = (1 / step).detach()
    s = grad_scale(step, 1.0 / ((qp * x.numel()) ** 0.5))
    q = round_pass(torch.clamp(x / s, -qp, qp))
    if isint:
        return q * s / s.detach(), scale
    return q * s, 1.0


def codes(x, bits, step):
    """The integer codes the quantiser assigns (what `LSQ integer codes bit-exact` is about): decided with the value
    of the grad-scaled step, fl(fl(s - s*g) + s*g), exactly as data_quant_lsq does."""
    qp = qp_of(bits)
    s = grad_scale(step, 1.0 / ((qp * x.numel()) ** 0.5))
    return torch.round(torch.clamp(x / s, -qp, qp))


def lazy_init_(p, key, x, bits):
    if p[key].item() == 1:
        p[key].data = torch.as_tensor(init_step(x, bits), dtype=torch.float32).reshape(())


def weight_qn(p, c, isint, noise_unit):
    w, scale = p["weight"], 1.0
    if c.get("weight_quant", True):
        lazy_init_(p, "step_size_weight", p["weight"], c["weight_bit"])
        w, scale = quant_lsq(p["weight"], c["weight_bit"], p["step_size_weight"], isint)
    if c["noise_scale"] != 0:
        w = w + (w.max() - w.min()) * c["noise_scale"] * noise_unit
    return w, scale


def lsq_layer(p, c, x, noise_unit=None, conv=None):
    """Conv2d_qn_lsq / Linear_qn_lsq forward; conv = dict(stride=, padding=) for the convolution."""
    if c.get("input_quant", True):
        lazy_init_(p, "step_size_input", x, c["input_bit"])
        x, _ = quant_lsq(x, c["input_bit"], p["step_size_input"], False)
    w, _ = weight_qn(p, c, False, noise_unit)
    y = F.conv2d(x, w, p.get("bias"), **conv) if conv is not None else F.linear(x, w, p.get("bias"))
    if c.get("output_quant", True):
        lazy_init_(p, "step_size_output", y, c["output_bit"])
        y, _ = quant_lsq(y, c["output_bit"], p["step_size_output"], False)
    return y


# ------------------------------------------------------------------------------------------------------ ADDA
def bit_split(xq, x_bit, slice_bit):
    """[B, ...] codes -> [slices * B, ...] signed digits; straight-through gradient shift / slices per slice."""
    n = int(math.ceil((x_bit - 1) / slice_bit))
    outs = []
    for k in range(n):
        lsb = k * slice_bit
        msb = min(lsb + slice_bit, x_bit - 1)
        v = xq / 2 ** lsb
        shift = (torch.floor(v.abs().detach()) * v.sign().detach() - v).detach() + v
        t = xq / 2 ** msb
        resid = torch.where(t >= 0, t.floor(), t.ceil()) * 2 ** slice_bit
        digit = shift - resid
        outs.append((digit - shift / n).detach() + shift / n)
    return torch.cat(outs, 0)


def adc_scale(p, c):
    g = clamp_pass(p["adc_gain"], c["adc_gain_min"], c["adc_gain_max"])
    if c["adc_adjust_mode"] == 'gain':
        g = round_pass(g)
    else:
        g = 1 / round_pass(1 / g)
    return g * c["adc_gain_1_scale"]


def adda_linear(p, c, x, mapping, noise_unit=None, gain_noise=None, offset_noise=None, conv=None):
    """Linear_lsq_adda_cim.forward on x [B, ..., K] (layers_qn_lsq_adda_cim.py:428-522) or, with conv = dict(kernel_size,
    stride, padding), Conv2d_lsq_adda_cim.forward on x [B, C, H, W] (:198-291): the DAC digits are unfolded with nn.Unfold
    ((c, ky, kx) row order, matching weight.view(Cout, -1)), the block loop runs on the [slices * B, L, C*k*k] matrix and
    the result is folded back to [B, Cout, oh, ow]."""
    R = qp_of(c["adc_bit"])
    slice_bit = c["dac_bit"] - 1
    slices = int(math.ceil((c["input_bit"] - 1) / slice_bit))
    lazy_init_(p, "step_size_input", x, c["input_bit"])
    xq, in_scale = quant_lsq(x, c["input_bit"], p["step_size_input"], True)
    xs = bit_split(xq, c["input_bit"], slice_bit)
    w, w_scale = weight_qn(p, c, True, noise_unit)
    if conv is not None:
        B, _, H, W = x.shape
        k, st, pd = conv["kernel_size"], conv["stride"], conv["padding"]
        oh, ow = (H + 2 * pd - k) // st + 1, (W + 2 * pd - k) // st + 1
        xs = torch.nn.functional.unfold(xs, kernel_size=k, padding=pd, stride=st).transpose(1, 2)   # [slices*B, L, C*k*k]
        w = w.reshape(w.shape[0], -1)
        x = xs[:B]   # only its leading shape is used below
    w2 = w.transpose(0, 1)
    out = torch.zeros(x.shape[:-1] + (w.shape[0],), dtype=x.dtype)
    alpha = adc_scale(p, c)
    for info in mapping.values():
        r0, rn, c0, cn = info["start_row"], info["row_num"], info["start_col"], info["col_num"]
        part = torch.matmul(xs[..., r0:r0 + rn], w2[r0:r0 + rn, c0:c0 + cn])
        if p["adc_gain"].item() == c["adc_gain_min"] and part.abs().max() != 0:
            ideal = (R / part.detach().abs().max()) / c["adc_gain_1_scale"]
            p["adc_gain"].data.copy_(torch.clamp(ideal, min=0.8 * c["adc_gain_min"], max=1.2 * c["adc_gain_max"]))
            alpha = adc_scale(p, c)
        a = alpha * part
        if gain_noise is not None:
            noisy = a * (1 + gain_noise[c0:c0 + cn]) + R * offset_noise[c0:c0 + cn]
            a = (noisy - a).detach() + a
        a = round_pass(torch.clamp(a, min=-R - 1, max=R))
        a = a.reshape((slices, a.shape[0] // slices) + a.shape[1:])
        pw = 2.0 ** (slice_bit * torch.arange(slices, dtype=a.dtype))
        out[..., c0:c0 + cn] = out[..., c0:c0 + cn] + torch.einsum('k...,k->...', a, pw)
    if conv is not None:
        out = out.transpose(1, 2).reshape(B, w.shape[0], oh, ow)
    y = out / w_scale / in_scale / alpha
    if p.get("bias") is not None:
        y = y + (p["bias"].view(1, -1, 1, 1) if conv is not None else p["bias"].view(1, -1))
    if c.get("output_quant", True):
        lazy_init_(p, "step_size_output", y, c["output_bit"])
        y, _ = quant_lsq(y, c["output_bit"], p["step_size_output"], False)
    return y

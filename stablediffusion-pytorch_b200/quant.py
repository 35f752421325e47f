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

from . import bank as BK
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


class WeightQnFn(Function):
    """weight_quant_noise (cim_layers/layers_utils_lsq.py:48-64) as two kernels forward and two backward
    (csrc/wquant.cu) instead of ~50 elementwise / reduction torch kernels per layer and step."""

    @staticmethod
    def forward(ctx, weight, step, qp, isint, n_scale, unit, scratch):
        ctx.save_for_backward(weight, step, unit)
        ctx.cfg = (qp, isint, n_scale, scratch)
        return rt.weight_qn(weight, step, qp, isint, n_scale, unit, scratch)

    @staticmethod
    def backward(ctx, g):
        weight, step, unit = ctx.saved_tensors
        qp, isint, n_scale, scratch = ctx.cfg
        dw, ds = rt.weight_qn_bwd(g, weight, step, qp, isint, n_scale, unit, scratch)
        return dw, ds.reshape(step.shape).to(step.dtype), None, None, None, None, None


BANK_DERIVED = True     # False: quantised weights keep the per-leaf packing launches inside a captured step (tests)


def _bank(w):
    """The active weight bank if it holds `w` (a quantised weight WeightSideFn wrote into the bank's buffer)."""
    b = BK.ACTIVE
    return b if b is not None and b.knows(w) else None


class WeightSideFn(Function):
    """weight_quant_noise of ALL quantised leaves of a model as one launch per phase (b200sd_weight_qn_multi): inputs
    (w0, s0, w1, s1, ...), outputs the quantised (+ noisy) weights, views of one flat buffer. Rows whose output received
    no gradient (a leaf the pass did not run) are skipped in the backward pass."""

    @staticmethod
    def forward(ctx, specs, *tensors):
        ws, steps = tensors[0::2], tensors[1::2]
        dev = ws[0].device
        # inside a captured training step the results live in the weight bank's buffer: it then re-derives the packed
        # bf16 forms of all of them with one launch (bank.refresh_derived) instead of one or two per leaf and direction
        bank = BK.ACTIVE if BANK_DERIVED else None
        if bank is not None:
            flat = bank.derived_buffer([w.numel() for w in ws])
        else:
            flat = torch.empty(sum(w.numel() for w in ws), device=dev, dtype=torch.float32)
        draw = [w.numel() if sp["n_scale"] != 0 and sp["unit"] is None else 0 for w, sp in zip(ws, specs)]
        noise = torch.randn(sum(draw), device=dev, dtype=torch.float32) if sum(draw) else None
        rows, outs, units = [], [], []
        o = z = 0
        for w, st, sp, nd in zip(ws, steps, specs, draw):
            n = w.numel()
            out = flat[o:o + n].view(w.shape)
            unit = sp["unit"]
            if nd:
                unit = noise[z:z + n].view(w.shape)
                z += n
            o += n
            rows.append({"w": w, "step": st, "unit": unit if sp["n_scale"] != 0 else None, "out": out, "qp": sp["qp"],
                         "n_scale": sp["n_scale"], "isint": sp["isint"], "scratch": sp["scratch"]})
            outs.append(out)
            units.append(rows[-1]["unit"])
        rt.weight_qn_multi(rows, backward=False)
        if bank is not None:
            bank.refresh_derived()
        ctx.save_for_backward(*tensors, *units)
        ctx.specs = specs
        ctx.set_materialize_grads(False)
        return tuple(outs)

    @staticmethod
    def backward(ctx, *gouts):
        k = len(ctx.specs)
        saved = ctx.saved_tensors
        tensors, units = saved[:2 * k], saved[2 * k:]
        ws, steps = tensors[0::2], tensors[1::2]
        live = [i for i, g in enumerate(gouts) if g is not None]
        grads = [None] * (2 * k)
        if live:
            dev = ws[0].device
            dflat = torch.empty(sum(ws[i].numel() for i in live), device=dev, dtype=torch.float32)
            dsteps = torch.empty(len(live), device=dev, dtype=torch.float32)
            rows = []
            o = 0
            for j, i in enumerate(live):
                sp, w, g = ctx.specs[i], ws[i], gouts[i]
                if g.dtype != torch.float32 or not g.is_contiguous():
                    g = g.float().contiguous()
                dw = dflat[o:o + w.numel()].view(w.shape)
                o += w.numel()
                rows.append({"w": w, "step": steps[i], "unit": units[i], "out": dw, "gout": g, "dstep": dsteps[j:j + 1],
                             "qp": sp["qp"], "n_scale": sp["n_scale"], "isint": sp["isint"], "scratch": sp["scratch"]})
                grads[2 * i] = dw
                grads[2 * i + 1] = dsteps[j].reshape(steps[i].shape).to(steps[i].dtype)
            rt.weight_qn_multi(rows, backward=True)
        return (None, *grads)


FUSED_WEIGHT_BANK = True


def prepare_weight_side(root):
    """Quantise (+ noise) the weights of every cim_layers leaf under `root` in one table-driven launch per phase and park
    the results on the leaves (`_wq_pending`), where their next `_w_qn()` picks them up; leaves that bypass
    quantisation, keep a non-fp32 / non-CUDA weight or are disabled simply keep the per-layer path. Returns the number
    of leaves served."""
    if not FUSED_WEIGHT_BANK:
        return 0
    leaves, specs, tensors = [], [], []
    for m in root.modules():
        m.__dict__.pop("_wq_pending", None)            # a result the previous pass did not consume
        spec = m._wq_spec() if hasattr(m, "_wq_spec") else None
        if spec is None:
            continue
        w, step = m.weight, m.step_size_weight
        scratch = m.__dict__.get("_wq_scratch")
        if scratch is None or scratch.device != w.device:
            scratch = m.__dict__["_wq_scratch"] = rt.wq_scratch(w.device)
        unit = spec["unit"]
        if spec["n_scale"] != 0 and unit is not None:
            unit = unit.to(w).contiguous()
        leaves.append(m)
        specs.append({"qp": qp_of(spec["bits"]), "isint": spec["isint"], "n_scale": float(spec["n_scale"]),
                      "unit": unit if spec["n_scale"] != 0 else None, "scratch": scratch})
        tensors += [w if w.is_contiguous() else w.contiguous(), step]
    if len(leaves) < 2:
        return 0
    if torch.is_grad_enabled() and any(t.requires_grad for t in tensors):
        outs = WeightSideFn.apply(specs, *tensors)
    else:
        outs = WeightSideFn.forward(_NoCtx(), specs, *[t.detach() for t in tensors])
    for m, sp, out in zip(leaves, specs, outs):
        m.__dict__["_wq_pending"] = (out, sp["isint"])
    return len(leaves)


class _NoCtx:
    """Stand-in for the autograd context when WeightSideFn.forward runs without recording."""

    def save_for_backward(self, *a):
        pass

    def set_materialize_grads(self, v):
        pass


def weight_qn(owner, weight, step, bits, isint, n_scale, unit=None, want_scale=True):
    """(quantised + noisy weight as a differentiable fp32 tensor, w_scale) of a cim_layers leaf `owner`; want_scale
    False skips the 1 / step launch for callers that read the step size on the device themselves."""
    qp = qp_of(bits)
    pend = owner.__dict__.pop("_wq_pending", None)
    if pend is not None and pend[1] == bool(isint) and pend[0].shape == weight.shape:
        return pend[0], ((1.0 / step).detach() if isint and want_scale else 1.0)
    scratch = owner.__dict__.get("_wq_scratch")
    if scratch is None or scratch.device != weight.device:
        scratch = owner.__dict__["_wq_scratch"] = rt.wq_scratch(weight.device)
    if n_scale != 0:
        unit = torch.randn_like(weight) if unit is None else unit.to(weight).contiguous()
    else:
        unit = None
    if torch.is_grad_enabled() and (weight.requires_grad or step.requires_grad):
        w = WeightQnFn.apply(weight, step, qp, isint, float(n_scale), unit, scratch)
    else:
        w = rt.weight_qn(weight.detach(), step, qp, isint, float(n_scale), unit, scratch)
    return w, ((1.0 / step).detach() if isint and want_scale else 1.0)


class AdcScaleFn(Function):
    """get_adc_scale (cim_layers/layers_utils_adda.py:40-49) as one launch; backward = one multiplication."""

    @staticmethod
    def forward(ctx, gain, gmin, gmax, gain_mode, scale1):
        out = rt.adc_scale(gain.detach(), gmin, gmax, gain_mode, scale1)
        ctx.save_for_backward(out)
        ctx.like = (gain.shape, gain.dtype)
        return out[0]

    @staticmethod
    def backward(ctx, g):
        (out,) = ctx.saved_tensors
        return (g * out[1]).reshape(ctx.like[0]).to(ctx.like[1]), None, None, None, None


class AddaOutFn(Function):
    """`acc / w_scale / in_scale / adc_scale + bias` followed by the LSQ output quantiser (step_out None: none) —
    cim_layers/layers_qn_lsq_adda_cim.py:482-500 — as one launch forward and one backward (csrc/adda_out.cu).
    w_scale = 1 / step_w and in_scale = 1 / step_in are detached in the reference: they receive no gradient here."""

    @staticmethod
    def forward(ctx, acc, alpha, bias, step_out, step_w, step_in, qp, numel, scratch):
        ctx.save_for_backward(acc, alpha, bias, step_out, step_w, step_in)
        ctx.cfg = (qp, numel, scratch)
        return rt.adda_out(acc, bias, step_w, step_in, alpha, step_out, qp, numel, 1 if step_out is not None else 2)

    @staticmethod
    def backward(ctx, g):
        acc, alpha, bias, step_out, step_w, step_in = ctx.saved_tensors
        qp, numel, scratch = ctx.cfg
        if g.dtype != torch.bfloat16 or g.stride(1) != 1 or g.stride(0) % 8:
            g = g.to(torch.bfloat16).contiguous()
        want_b = bias is not None and ctx.needs_input_grad[2]
        dacc, dy, res = rt.adda_out_bwd(g, acc, bias, step_w, step_in, alpha, step_out, qp, numel, scratch, want_b)
        dbias = rt.colsum(dy, 1, dy.shape[0], dy.shape[1]).to(bias.dtype) if want_b else None
        dstep = res[0].reshape(step_out.shape).to(step_out.dtype) if step_out is not None else None
        return dacc, res[1].reshape(alpha.shape).to(alpha.dtype), dbias, dstep, None, None, None, None, None


class AddaLayerFn(Function):
    """A compute-in-memory layer whose whole weight is ONE mapped block (every Linear of the DiT-9L but the MLP's
    second), end to end: DAC digits -> block product (tcgen05 GEMM, fp32) -> ADC + recombination + division chain + bias +
    output quantiser in one pass (b200sd_adda_tail). The integer accumulation and its gradient never reach memory, and
    the fp32 block product is KEPT for the backward pass instead of being recomputed (one GEMM less per layer; 8 bytes
    per output element and DAC slice held between forward and backward). Same arithmetic as AddaFn + AddaOutFn."""

    @staticmethod
    def forward(ctx, x, step_in, w, alpha, bias, step_out, step_w, cfg, qp_out, numel_out, scratch):
        M, N = x.shape[0], w.shape[0]
        xs, ws = x.detach(), w.detach()
        digits = AddaFn._digits(xs, step_in.detach(), cfg)
        bk = None if cfg.get("split_w") else _bank(ws)
        wp = bk.linear(ws) if bk else ops.pack_linear_weight(ws)
        part = rt.linear(digits, wp, out_dtype=torch.float32)
        if cfg.get("split_w"):
            rt.linear(digits, ops.pack_linear_weight(ws - wp[:, :ws.shape[1]].float()), out=part, residual=part)
        out = rt.adda_tail(part, M, N, bias, step_w, step_in, alpha.detach().reshape(1).float(), step_out, qp_out, numel_out,
                           cfg["R"], cfg["slices"], cfg["slice_bit"], cfg["gain_noise"], cfg["offset_noise"])
        ctx.save_for_backward(x, step_in, w, alpha, bias, step_out, step_w, digits, part)
        ctx.cfg = (cfg, qp_out, numel_out, scratch)
        return out

    @staticmethod
    def backward(ctx, g):
        x, step_in, w, alpha, bias, step_out, step_w, digits, part = ctx.saved_tensors
        cfg, qp_out, numel_out, scratch = ctx.cfg
        M, K = x.shape
        N = w.shape[0]
        S, sb = cfg["slices"], cfg["slice_bit"]
        if g.dtype != torch.bfloat16 or g.stride(1) != 1 or g.stride(0) % 8:
            g = g.to(torch.bfloat16).contiguous()
        want_b = bias is not None and ctx.needs_input_grad[4]
        dp, dy, res = rt.adda_tail_bwd(g, part, M, N, bias, step_w, step_in, alpha.detach().reshape(1).float(), step_out,
                                       qp_out, numel_out, cfg["R"], S, sb, scratch, want_b, cfg["gain_noise"],
                                       cfg["offset_noise"])
        wd = w.detach()
        kp = (K + 63) // 64 * 64
        with rt.Side(True, x.device) as side:      # weight and bias gradients: a branch beside the input-gradient chain
            dwp = rt.wgrad([rt.make_src(digits, 1, 1, S * M, K, ld=K)], [(0, 0, 0, 0, kp // 64)],
                           rt.make_src(dp, 1, 1, S * M, N, ld=dp.stride(0)), N, 1, 1, S * M)
            dbias = rt.colsum(dy, 1, M, N).to(bias.dtype) if want_b else None
        ddigits = torch.empty(S * M, K, device=x.device, dtype=torch.bfloat16)
        bk = _bank(wd)
        rt.linear(dp, bk.linear_t(wd) if bk else ops.pack_linear_weight(wd.t()), out=ddigits)
        dxq = rt.bit_split_bwd(ddigits, M, K, S, sb)
        dx, ds = rt.lsq_quant_bwd(dxq, x, step_in.detach(), cfg["qp_in"], codes=True, numel=cfg["numel_in"])
        side.join()
        dstep_out = res[0].reshape(step_out.shape).to(step_out.dtype) if step_out is not None else None
        return (dx if ctx.needs_input_grad[0] else None), ds.reshape(step_in.shape), dwp[:, :K], \
            res[1].reshape(alpha.shape).to(alpha.dtype), dbias, dstep_out, None, None, None, None, None


def adda_layer(owner, x, step_in, w, alpha, bias, step_out, step_w, cfg, bits_out, numel_out):
    """Single-block compute-in-memory layer on bf16 rows [M, K] -> bf16 rows [M, N] (see AddaLayerFn)."""
    qp = qp_of(bits_out) if step_out is not None else 0.0
    scratch = owner.__dict__.get("_ao_scratch")
    if scratch is None or scratch.device != x.device:
        scratch = owner.__dict__["_ao_scratch"] = rt.adda_out_scratch(x.device)
    if bias is not None and bias.dtype != torch.float32:
        bias = bias.float()
    if torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in (x, step_in, w, alpha, bias, step_out)):
        return AddaLayerFn.apply(x, step_in, w, alpha, bias, step_out, step_w, cfg, qp, numel_out, scratch)
    return AddaLayerFn.forward(_NoCtx(), x, step_in, w, alpha, bias, step_out, step_w, cfg, qp, numel_out, scratch)


def adda_out(owner, acc, alpha, bias, step_out, step_w, step_in, bits, numel):
    """Output side of a compute-in-memory leaf `owner`: bf16 rows [M, Np]."""
    qp = qp_of(bits) if step_out is not None else 0.0
    scratch = owner.__dict__.get("_ao_scratch")
    if scratch is None or scratch.device != acc.device:
        scratch = owner.__dict__["_ao_scratch"] = rt.adda_out_scratch(acc.device)
    if bias is not None and bias.dtype != torch.float32:
        bias = bias.float()
    if torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in (acc, alpha, bias, step_out)):
        return AddaOutFn.apply(acc, alpha, bias, step_out, step_w, step_in, qp, numel, scratch)
    return rt.adda_out(acc, bias, step_w, step_in, alpha, step_out, qp, numel, 1 if step_out is not None else 2)


def rescale_step_sizes(module, weight_bit_old, input_bit_old, output_bit_old):
    """update_step_size (layers_utils_lsq.py:11-28): a bit-width change by d bits divides the step size by 2^d."""
    for name, new, old in (("weight", module.weight_bit, weight_bit_old), ("input", module.input_bit, input_bit_old),
                           ("output", module.output_bit, output_bit_old)):
        if new != old:
            getattr(module, f"step_size_{name}").data /= 2 ** (new - old)


# ---------------------------------------------------------------------------------------------- ADDA block products
from . import ops  # noqa: E402


def _covers(ranges, width):
    """True when the (start, length) ranges tile [0, width) without a gap (overlaps / repeats allowed)."""
    end = 0
    for a, n in sorted(set(ranges)):
        if a > end:
            return False
        end = max(end, a + n)
    return end >= width


class AddaFn(Function):
    """Integer accumulation of a compute-in-memory layer (cim_layers/layers_qn_lsq_adda_cim.py:430-480):
    acc[M, N] = sum over mapped blocks and DAC slices of 2^(k*slice_bit) * ADC(alpha * digits_k[:, rows] @ W[cols, rows]^T).
    x: bf16 activation rows; w: fp32 [N, K] weight codes (+ noise), differentiable; alpha: 0-dim ADC scale.
    Backward recomputes the block products (no fp32 block products are kept between forward and backward; the bf16
    digits and the packed weight blocks are):
    per block one ADC-backward kernel, a dgrad GEMM (accumulated over column blocks through the residual epilogue) and a
    wgrad GEMM; then bit-split and LSQ-code backward."""

    @staticmethod
    def forward(ctx, x, step_in, w, alpha, cfg):
        ctx.save_for_backward(x, step_in, w, alpha)
        ctx.cfg = cfg
        ctx.keep = {}      # DAC digits and packed weight blocks of this forward, reused by its backward
        return AddaFn._run(x, step_in.detach(), w.detach(), alpha.detach(), cfg, ctx.keep)

    @staticmethod
    def _digits(x, step_in, cfg):
        return rt.bit_split(x, cfg["slices"], cfg["slice_bit"], cfg["input_bit"], step=step_in, qp=cfg["qp_in"],
                            numel=cfg["numel_in"])

    @staticmethod
    def _run(x, step_in, w, alpha, cfg, keep=None):
        M, N = x.shape[0], w.shape[0]
        digits = AddaFn._digits(x, step_in, cfg)
        if keep is not None:
            keep["digits"], keep["packed"] = digits, []
        a1 = alpha.reshape(1).float()
        # the first block of a column range stores, later row blocks accumulate; a zero fill only when the blocks
        # leave columns uncovered (padding columns of a layer whose width is not a multiple of 8)
        covered = _covers(((c0, cn) for _, _, c0, cn in cfg["blocks"]), N)
        out = (torch.empty if covered else torch.zeros)(M, N, device=x.device, dtype=torch.float32)
        seen = set()
        bk = None if cfg.get("split_w") else _bank(w)
        for r0, rn, c0, cn in cfg["blocks"]:
            wb = w[c0:c0 + cn, r0:r0 + rn]
            wp = bk.block(w, c0, cn, r0, rn) if bk else ops.pack_linear_weight(wb)
            if keep is not None:
                keep["packed"].append(wp)
            part = rt.linear(digits[:, r0:r0 + rn], wp, out_dtype=torch.float32)
            if cfg.get("split_w"):
                # precise parity mode: the noisy (non-integer) weight codes as a (hi, lo) pair of bf16 operands — the
                # digits are exact, so the analogue partial sums carry 16 significand bits of the weights instead of 8
                # and the ADC codes stop flipping on the operand format (the fast path: 2-6 % of output codes)
                lo = ops.pack_linear_weight(wb - wp[:, :rn].float())
                rt.linear(digits[:, r0:r0 + rn], lo, out=part, residual=part)
            rt.adc_combine(part, out, M, cn, c0, cfg["slices"], cfg["slice_bit"], a1, cfg["R"], cfg["gain_noise"],
                           cfg["offset_noise"], accumulate=not covered or c0 in seen)
            seen.add(c0)
        return out

    @staticmethod
    def backward(ctx, dacc):
        x, step_in, w, alpha = ctx.saved_tensors
        cfg = ctx.cfg
        M, K = x.shape
        N = w.shape[0]
        S, sb = cfg["slices"], cfg["slice_bit"]
        dacc = dacc.float().contiguous()
        wd = w.detach()
        a1 = alpha.detach().reshape(1).float()
        keep = ctx.keep
        digits = keep["digits"] if "digits" in keep else AddaFn._digits(x, step_in.detach(), cfg)
        packed = keep.get("packed")
        blocks = cfg["blocks"]
        rows_covered = _covers(((r0, rn) for r0, rn, _, _ in blocks), K)
        ddigits = (torch.empty if rows_covered else torch.zeros)(S * M, K, device=x.device, dtype=torch.bfloat16)
        single = len(blocks) == 1 and blocks[0][1] == K and blocks[0][3] == N
        dw = None if single else torch.zeros(N, K, device=x.device, dtype=torch.float32)
        das = []
        seen = set()
        bk = _bank(wd)
        for bi, (r0, rn, c0, cn) in enumerate(blocks):
            wb = wd[c0:c0 + cn, r0:r0 + rn]
            part = rt.linear(digits[:, r0:r0 + rn], packed[bi] if packed else ops.pack_linear_weight(wb),
                             out_dtype=torch.float32)
            dp, da = rt.adc_bwd(part, dacc, M, cn, c0, S, sb, a1, cfg["R"], cfg["gain_noise"], cfg["offset_noise"])
            das.append(da)
            view = ddigits[:, r0:r0 + rn]
            rt.linear(dp[:, :cn], bk.block_t(wd, c0, cn, r0, rn) if bk else ops.pack_linear_weight(wb.t()), out=view,
                      residual=view if r0 in seen else None)
            seen.add(r0)
            kp = (rn + 63) // 64 * 64
            dwp = rt.wgrad([rt.make_src(digits[:, r0:r0 + rn], 1, 1, S * M, rn, ld=K)], [(0, 0, 0, 0, kp // 64)],
                           rt.make_src(dp, 1, 1, S * M, cn, ld=dp.stride(0)), cn, 1, 1, S * M)
            if single:
                dw = dwp[:, :rn]
            else:
                dw[c0:c0 + cn, r0:r0 + rn] = dwp[:, :rn]
        dalpha = das[0] if len(das) == 1 else torch.cat(das).sum()
        keep.clear()
        dxq = rt.bit_split_bwd(ddigits, M, K, S, sb)
        dx, ds = rt.lsq_quant_bwd(dxq, x, step_in.detach(), cfg["qp_in"], codes=True, numel=cfg["numel_in"])
        return (dx if ctx.needs_input_grad[0] else None), ds.reshape(step_in.shape), dw, \
            dalpha.reshape(alpha.shape).to(alpha.dtype), None

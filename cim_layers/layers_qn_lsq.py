"""Drop-in for the reference's cim_layers/layers_qn_lsq.py: `Conv2d_qn_lsq` / `Linear_qn_lsq` —
    y = Q_out( conv_or_linear( Q_in(x), W_q + noise, bias ) )
with LSQ (learned step size) quantisers, weight noise proportional to the quantised weight range, a `use_FP` bypass
and `update_para` bit-width changes. Same constructor keywords, parameter names (`step_size_input/output/weight`,
checkpoints interchange) and lazy step-size initialisation as the reference (:17-121,126-216); activations are
quantised by CUDA kernels, the product runs on the tcgen05 implicit GEMM, backward included."""
import torch
import torch.nn as nn

import sdb200  # noqa: F401
from sdb200 import autograd as AG
from sdb200 import engine as E
from sdb200 import ops
from sdb200 import quant as Q
from sdb200 import runtime as rt


FUSED_WEIGHT_SIDE = True   # tests flip this to compare the fused kernels with the op-by-op torch formulation


class _LsqMixin:
    def _init_lsq(self, weight_bit, input_bit, output_bit, noise_scale, input_quant, output_quant, weight_quant,
                  gain_noise_scale, offset_noise_scale):
        self.use_FP = False
        self.weight_bit = weight_bit
        self.input_bit = input_bit
        self.output_bit = output_bit
        self.noise_scale = noise_scale
        self.gain_noise_scale = gain_noise_scale
        self.offset_noise_scale = offset_noise_scale
        self.input_quant = input_quant
        self.output_quant = output_quant
        self.weight_quant = weight_quant
        self.step_size_input = nn.Parameter(torch.tensor(1.0))
        self.step_size_output = nn.Parameter(torch.tensor(1.0))
        self.step_size_weight = nn.Parameter(torch.tensor(1.0))
        self.weight_noise_unit = None   # tests inject the N(0,1) draw of the weight noise here

    def update_para(self, use_FP=False, weight_bit=None, input_bit=None, output_bit=None, noise_scale=None,
                    gain_noise_scale=0, offset_noise_scale=0):
        old = (self.weight_bit, self.input_bit, self.output_bit)
        self.use_FP = use_FP
        if weight_bit is not None:
            self.weight_bit = weight_bit
        if input_bit is not None:
            self.input_bit = input_bit
        if output_bit is not None:
            self.output_bit = output_bit
        if noise_scale is not None:
            self.noise_scale = noise_scale
        if gain_noise_scale is not None:
            self.gain_noise_scale = gain_noise_scale
        if offset_noise_scale is not None:
            self.offset_noise_scale = offset_noise_scale
        Q.rescale_step_sizes(self, *old)

    def _linear(self, rows, w, act=None):
        """rows @ w^T + bias with an explicit (already quantised) weight tensor."""
        if E.training():
            o, _ = AG.LinearFn.apply(rows, w, self.bias, None, None, False, act)
            return o
        return rt.linear(rows, ops.pack_linear_weight(w), bias=None if self.bias is None else self.bias.detach().float(),
                         act=act)

    # ---- the three quantisers ---------------------------------------------------------------------------------
    def _needs_lazy_init(self, which):
        """True while step_size_<which> may still be the untouched 1.0 (layers_utils_lsq.py:40-42,57-59,74-76 test it
        on the host on every forward). Here the test runs on the device (absmax + conditional init: two launches); once a
        forward outside stream capture has SEEN the step differ from 1 the two launches are dropped for good — 384 launches
        per training step of the DiT-9L otherwise. Loading a state_dict re-arms the check."""
        return which not in self.__dict__.setdefault("_lsq_ready", set())

    def _note_lazy_init(self, which, step):
        if which not in self.__dict__["_lsq_ready"] and not torch.cuda.is_current_stream_capturing() \
                and step.item() != 1.0:
            self.__dict__["_lsq_ready"].add(which)

    def _q_in(self, rows, numel, isint=False):
        if not self.input_quant:
            return rows
        lazy = self._needs_lazy_init("input")
        out = Q.quantize_rows(rows, self.step_size_input, self.input_bit, codes=isint, numel=numel, lazy_init=lazy)
        if lazy:
            self._note_lazy_init("input", self.step_size_input)
        return out

    def _q_out(self, rows, numel):
        if not self.output_quant:
            return rows
        lazy = self._needs_lazy_init("output")
        out = Q.quantize_rows(rows, self.step_size_output, self.output_bit, numel=numel, lazy_init=lazy)
        if lazy:
            self._note_lazy_init("output", self.step_size_output)
        return out

    def _rearm_lazy_init(self):
        self.__dict__.pop("_lsq_ready", None)     # loaded step sizes may be the untouched 1.0 again
        self.__dict__["_gain_init_pending"] = True

    # ---- precise parity mode (sdb200.engine.set_parity_mode("precise"), inference): the product in the CODE domain
    def _exact_mode(self):
        return E.precise() and self.input_quant and self.weight_quant and self.weight.dtype == torch.float32

    def _eff_step(self, step, bits, numel):
        """VALUE of the grad-scaled step the reference quantises with (quant_noise_utils.py:51-58), fp32 0-dim."""
        return Q._grad_scale(step.detach().float(), rt.lsq_grad_scale(Q.qp_of(bits), numel)).reshape(())

    def _exact_weights(self):
        """Operand list [(bf16-exact weight tensor, fp32 scalar factor or None)] whose products with the integer input
        codes sum to the reference's fp32 product up to fp32 summation order: without weight noise the integer weight
        codes (times s_x' s_w' afterwards); with noise the fp32 noisy weight as a (hi, lo) pair of bf16 tensors — 16
        significand bits instead of 8, so a pre-quantisation value no longer crosses an output rounding boundary
        because of the operand format (the fast path's 2-10 % of output codes off by one)."""
        if self.noise_scale == 0:
            wq, _ = self._w_qn(isint=True)
            return [torch.round(wq.detach())], self._eff_step(self.step_size_weight, self.weight_bit, self.weight.numel())
        w, _ = self._w_qn()
        w = w.detach()
        hi = w.to(torch.bfloat16).float()
        return [hi, (w - hi).to(torch.bfloat16).float()], None

    _WQ_ISINT = False      # the ADDA classes quantise their weights to integer codes (isint=True)

    def _wq_spec(self):
        """What sdb200.quant.prepare_weight_side needs to serve this leaf from the model-wide launch, or None when the
        leaf keeps the per-layer path (bypassed / unquantised / weight not an fp32 CUDA tensor)."""
        if getattr(self, "use_FP", False) or not self.weight_quant or not FUSED_WEIGHT_SIDE:
            return None
        if not self.weight.is_cuda or self.weight.dtype != torch.float32:
            return None
        return {"bits": self.weight_bit, "isint": self._WQ_ISINT, "n_scale": self.noise_scale,
                "unit": self.weight_noise_unit}

    def _w_qn(self, isint=False, want_scale=True):
        """(quantised + noisy weight as a differentiable fp32 tensor, w_scale)."""
        w, scale = self.weight, 1.0
        if self.weight_quant:
            if FUSED_WEIGHT_SIDE and self.weight.is_cuda and self.weight.dtype == torch.float32:
                # lazy init + LSQ quantiser + noise: two launches forward, two backward (csrc/wquant.cu)
                return Q.weight_qn(self, self.weight, self.step_size_weight, self.weight_bit, isint, self.noise_scale,
                                   self.weight_noise_unit, want_scale=want_scale)
            Q.init_weight_step_(self.step_size_weight, self.weight, self.weight_bit)
            w, scale = Q.quantize_weight(self.weight, self.step_size_weight, self.weight_bit, isint)
        return Q.add_weight_noise(w, self.noise_scale, self.weight_noise_unit), scale


class Conv2d_qn_lsq(nn.Conv2d, _LsqMixin):
    def __init__(self, in_channels, out_channels, kernel_size, stride, padding, groups, weight_bit, input_bit,
                 output_bit, noise_scale, input_quant=True, output_quant=True, weight_quant=True, gain_noise_scale=0,
                 offset_noise_scale=0, bias=True):
        super().__init__(in_channels=in_channels, out_channels=out_channels, kernel_size=kernel_size, stride=stride,
                         padding=padding, groups=groups, bias=bias)
        self._init_lsq(weight_bit, input_bit, output_bit, noise_scale, input_quant, output_quant, weight_quant,
                       gain_noise_scale, offset_noise_scale)

    def _conv(self, a, x_rows, w):
        """Convolution of NHWC rows with an explicit (already quantised) weight tensor -> Act."""
        k, s, p = self.kernel_size, self.stride, self.padding
        if self.groups != 1 or self.dilation != (1, 1):
            raise rt.B200Error("Conv2d_qn_lsq: groups/dilation are not supported")
        cout = self.out_channels
        same = k in ((3, 3), (1, 1)) and s == (1, 1) and p == (k[0] // 2, k[0] // 2)
        down = k == (4, 4) and s == (2, 2) and p == (1, 1)
        if not (same or down):
            raise rt.B200Error(f"Conv2d_qn_lsq: kernel {k} stride {s} padding {p} has no CUDA path")
        if E.training():
            if same:
                o, _ = AG.ConvSameFn.apply(x_rows, w, self.bias, None, None, None, None, (a.n, a.h, a.w, a.c, 0), False,
                                           False)
                return E.Act(o, a.n, a.h, a.w, cout)
            o, _ = AG.ConvDownFn.apply(x_rows, w, self.bias, (a.n, a.h, a.w, a.c), False)
            return E.Act(o, a.n, a.h // 2, a.w // 2, cout)
        wmat = ops.pack_conv_weight(w)
        bias = None if self.bias is None else self.bias.detach().float()
        if same:
            o = torch.empty(a.n * a.hw, (cout + 7) // 8 * 8, device=x_rows.device, dtype=torch.bfloat16)
            ops.conv_same(x_rows, a.n, a.h, a.w, a.c, wmat, cout, k[0], bias=bias, ld_x=x_rows.stride(0), out=o)
            return E.Act(o, a.n, a.h, a.w, cout)
        o = E.Act.empty(a.n, a.h // 2, a.w // 2, cout, x_rows.device)
        ops.conv_down4x4(x_rows, a.n, a.h, a.w, a.c, wmat, cout, bias=bias, ld_x=x_rows.stride(0), out=o.t)
        return o

    def _conv_exact(self, a):
        """Precise parity mode: integer input codes against exact weight operands, fp32 accumulation, the steps applied
        afterwards (see _exact_weights). 3x3 / 1x1 'same' convolutions with >= 8-aligned output channels; None otherwise."""
        k, s, p = self.kernel_size, self.stride, self.padding
        cout = self.out_channels
        if not (k in ((3, 3), (1, 1)) and s == (1, 1) and p == (k[0] // 2, k[0] // 2) and self.groups == 1 and cout % 8 == 0):
            return None
        codes = self._q_in(a.t, a.n * a.hw * a.c, isint=True)
        ws, sw = self._exact_weights()
        acc = None
        for wpart in ws:
            wm = ops.pack_conv_weight(wpart)
            if acc is None:
                acc = torch.empty(a.n * a.hw, cout, device=codes.device, dtype=torch.float32)
                ops.conv_same(codes, a.n, a.h, a.w, a.c, wm, cout, k[0], ld_x=codes.stride(0), out=acc)
            else:
                ops.conv_same(codes, a.n, a.h, a.w, a.c, wm, cout, k[0], ld_x=codes.stride(0), out=acc, residual=acc)
        sx = self._eff_step(self.step_size_input, self.input_bit, a.n * a.hw * a.c)
        y = acc * (sx * sw if sw is not None else sx)
        if self.bias is not None:
            y = y + self.bias.detach().float()
        return E.Act(self._q_out(y.contiguous(), a.n * a.hw * cout), a.n, a.h, a.w, cout)

    def forward_act(self, a):
        """Engine entry: NHWC bf16 Act -> Act. Channel counts that are not multiples of 8 (the DiT mask conv: 18 -> 3)
        travel in zero-padded rows; zeros quantise to code 0 and do not move the lazily initialised step sizes."""
        if self.use_FP:
            return self._conv(a, a.t, self.weight)
        cp = (a.c + 7) // 8 * 8
        if a.t.shape[1] != cp:
            raise rt.B200Error("Conv2d_qn_lsq: input rows must be packed to the channel count rounded up to 8")
        if self._exact_mode():
            y = self._conv_exact(a)
            if y is not None:
                return y
        xq = self._q_in(a.t, a.n * a.hw * a.c)
        w, _ = self._w_qn()
        y = self._conv(a, xq, w)
        op = (y.c + 7) // 8 * 8
        t = y.t
        if y.c % 8:  # the GEMM leaves the padding columns of its output unwritten: make them zeros
            t = torch.nn.functional.pad(t[:, :y.c], (0, op - y.c))
        elif t.shape[1] != op or not t.is_contiguous():
            t = t[:, :op].contiguous()
        y = E.Act(self._q_out(t, y.n * y.hw * y.c), y.n, y.h, y.w, y.c)
        return y

    def forward(self, x):
        if not x.is_cuda:
            raise rt.B200Error("Conv2d_qn_lsq.forward: CUDA tensors required (no CPU implementation in this package)")
        with torch.autocast("cuda", enabled=False):
            out = self.forward_act(E.Act.from_nchw(x)).to_nchw()
        return out.to(x.dtype)


class Linear_qn_lsq(nn.Linear, _LsqMixin):
    def __init__(self, in_features, out_features, weight_bit, input_bit, output_bit, noise_scale, input_quant=True,
                 output_quant=True, weight_quant=True, gain_noise_scale=0, offset_noise_scale=0, bias=False):
        super().__init__(in_features, out_features, bias)
        self._init_lsq(weight_bit, input_bit, output_bit, noise_scale, input_quant, output_quant, weight_quant,
                       gain_noise_scale, offset_noise_scale)

    def forward_rows(self, rows, act=None):
        """Engine entry: bf16 rows [M, >= in_features] -> bf16 rows [M, out_features] (act applied after Q_out).
        in_features that are not a multiple of 8 (the DiT patch embedding: 28) are served on rows zero-padded to the
        next multiple of 8: zeros quantise to code 0 and meet zero weight columns."""
        K, N = self.in_features, self.out_features
        if self.use_FP:
            return self._linear(rows if rows.shape[1] == K or K % 8 else rows[:, :K].contiguous(), self._pad_k(self.weight), act)
        if N % 8:
            raise rt.B200Error("Linear_qn_lsq: out_features must be a multiple of 8")
        Kp = (K + 7) // 8 * 8
        if rows.shape[1] < Kp:
            raise rt.B200Error("Linear_qn_lsq: input rows narrower than in_features rounded up to 8")
        x = rows if rows.shape[1] == Kp else rows[:, :Kp].contiguous()
        if self._exact_mode():
            codes = self._q_in(x, x.shape[0] * K, isint=True)          # bf16 integers: exact
            ws, sw = self._exact_weights()
            acc = None
            for wpart in ws:      # fp32 accumulation; the second operand of a (hi, lo) pair adds in place
                wp = ops.pack_linear_weight(self._pad_k(wpart))
                acc = rt.linear(codes, wp, out_dtype=torch.float32) if acc is None else \
                    rt.linear(codes, wp, out=acc, residual=acc)
            sx = self._eff_step(self.step_size_input, self.input_bit, x.shape[0] * K)
            y = acc * (sx * sw if sw is not None else sx)
            if self.bias is not None:
                y = y + self.bias.detach().float()
            y = self._q_out(y.contiguous(), y.shape[0] * N)
            if act is not None:
                if act != "relu":
                    raise rt.B200Error("Linear_qn_lsq: only ReLU may follow a quantised Linear")
                y = torch.relu(y)
            return y
        xq = self._q_in(x, x.shape[0] * K)
        w, _ = self._w_qn()
        y = self._linear(xq, self._pad_k(w))
        y = self._q_out(y if y.is_contiguous() else y.contiguous(), y.shape[0] * N)
        if act is not None:
            if act != "relu":
                raise rt.B200Error("Linear_qn_lsq: only ReLU may follow a quantised Linear")
            y = torch.relu(y) if E.training() else rt.copy_act(y, torch.empty_like(y), 0, act="relu")
        return y

    def _pad_k(self, w):
        K = self.in_features
        return w if K % 8 == 0 else torch.nn.functional.pad(w, (0, (K + 7) // 8 * 8 - K))

    def forward(self, x):
        if not x.is_cuda:
            raise rt.B200Error("Linear_qn_lsq.forward: CUDA tensors required (no CPU implementation in this package)")
        lead = x.shape[:-1]
        with torch.autocast("cuda", enabled=False):
            rows = x.reshape(-1, x.shape[-1])
            rows = rows.to(torch.bfloat16) if rows.dtype != torch.bfloat16 else rows
            if rows.shape[1] % 8:
                rows = torch.nn.functional.pad(rows, (0, 8 - rows.shape[1] % 8))
            out = self.forward_rows(rows.contiguous())
        return out.reshape(*lead, self.out_features).to(x.dtype)


def _rearm_on_load(cls):
    """nn.Conv2d / nn.Linear come first in the MRO, so a mixin cannot override _load_from_state_dict: patch the class."""
    base = cls._load_from_state_dict

    def _load_from_state_dict(self, *a, **k):
        self._rearm_lazy_init()
        return base(self, *a, **k)

    cls._load_from_state_dict = _load_from_state_dict
    return cls


_rearm_on_load(Conv2d_qn_lsq)
_rearm_on_load(Linear_qn_lsq)

"""Drop-in for the reference's cim_layers/layers_qn_lsq_adda_cim.py: `Linear_lsq_adda_cim` / `Conv2d_lsq_adda_cim`,
the compute-in-memory model of a layer (reference :25-522):

  input  -> LSQ integer codes -> DAC bit slices (signed digits of dac_bit-1 bits)
  weight -> LSQ integer codes (+ noise proportional to the code range)
  for every mapped weight block (rows <= array height, `module.weight_mapping_info`):
      analogue partial sum of each bit slice against the block  ->  ADC: scale, (gain/offset noise), clamp to
      [-R-1, R], round  ->  slices recombined with weights 2^(k*slice_bit)  ->  accumulated into the output columns
  output -> / (w_scale * in_scale * adc_scale) + bias -> LSQ output quantiser

Here the digits and weight codes are exact in bf16 and every block product runs on the tcgen05 GEMM with fp32
accumulation (partial sums are integers < 2^24: exact), so the ADC codes are bit-exact with the reference's fp32
arithmetic when no weight noise is injected. Step sizes, the ADC gain and its lazy initialisation stay on the device.

The backward pass (progressive quant-noise training, Model_DiT_9L_train.py:610-623) follows the reference's autograd
graph: straight-through rounds, the ADC clamp mask, bit-split gradient 1/(slices * 2^lsb), LSQ step-size gradients,
and the ADC-gain gradient through both the scale and the final division (sdb200/quant.py: AddaFn).

k x k convolutions (the leaves of the ReLU-UNet pipeline, Model_Unet_train.py) follow the reference's own route
(:92-94,198-260): the input is unfolded into rows in nn.Unfold's (c, ky, kx) column order — the K axis the mapped weight
blocks split — by a CUDA im2col kernel (b200sd_im2col, adjoint b200sd_col2im for the backward pass), and the rows run
through the same block loop as a Linear layer. Quantising before or after the unfold is the same thing element by
element; padding taps are code 0; the LSQ gradient scale keeps the ORIGINAL tensor's element count."""
import math

import torch
import torch.nn as nn

import sdb200  # noqa: F401
from sdb200 import engine as E
from sdb200 import ops
from sdb200 import quant as Q
from sdb200 import runtime as rt
from cim_layers.layers_qn_lsq import _LsqMixin, _rearm_on_load


def gen_weight_split_dict(module, weight_block_size):
    """Block list of a layer for an array of `weight_block_size` = [rows, cols] — same keys and fields as the
    reference mapper produces (cim_weight_mapper/weight_process.py:10-40), for callers that do not run the mapper."""
    max_rows, max_cols = weight_block_size
    cols = module.weight.shape[0]
    rows = module.weight.numel() // cols
    info = {}
    for rb in range(rows // max_rows + 1):
        for cb in range(cols // max_cols + 1):
            r0, c0 = rb * max_rows, cb * max_cols
            rn, cn = min(max_rows, rows - r0), min(max_cols, cols - c0)
            if rn > 0 and cn > 0:
                info[f"{rb}_{cb}"] = {"start_row": r0, "start_col": c0, "row_num": rn, "col_num": cn}
    return info


FUSED_SINGLE_BLOCK = True     # single-block layers through AddaLayerFn (tests flip it to compare with the block loop)


class _Im2colFn(torch.autograd.Function):
    """NHWC bf16 rows -> nn.Unfold rows ((c, ky, kx) order), backward = nn.Fold (b200sd_col2im)."""

    @staticmethod
    def forward(ctx, x, geom):
        n, h, w, c, k, stride, pad = geom
        ctx.geom = geom
        ctx.ld = x.shape[1]
        return rt.im2col(x, n, h, w, c, k, stride, pad)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        n, h, w, c, k, stride, pad = ctx.geom
        dx = rt.col2im(g.contiguous(), n, h, w, c, k, stride, pad)
        if dx.shape[1] != ctx.ld:
            dx = torch.nn.functional.pad(dx, (0, ctx.ld - dx.shape[1]))
        return dx, None


class _AddaMixin(_LsqMixin):
    _WQ_ISINT = True

    def _init_adda(self, dac_bit, adc_bit, adc_gain_1_scale, adc_gain_range, adc_adjust_mode, seed):
        self.dac_bit = dac_bit
        self.adc_bit = adc_bit
        self.adc_gain_1_scale = adc_gain_1_scale
        self.slice_bit = dac_bit - 1
        self.bit_slices = int(math.ceil((self.input_bit - 1) / self.slice_bit))
        self.adc_gain_min = float(adc_gain_range[0])
        self.adc_gain_max = float(adc_gain_range[1])
        self.adc_gain = nn.Parameter(torch.tensor(self.adc_gain_min))
        self.adc_adjust_mode = adc_adjust_mode
        self.seed = seed
        self._gen_adc_noise()

    def _gen_adc_noise(self):
        """gen_adc_noise (layers_utils_adda.py:100-105): 1000 gain and offset draws per layer from a generator seeded
        with `seed` (the reference reseeds the GLOBAL generator here — oracle trap (v); a private generator gives the
        same numbers without that side effect)."""
        g = torch.Generator(device=self.weight.device).manual_seed(self.seed)
        self.gain_noise = torch.randn(1000, device=self.weight.device, generator=g) * self.gain_noise_scale
        self.offset_noise = torch.randn(1000, device=self.weight.device, generator=g) * self.offset_noise_scale

    def update_para(self, use_FP=False, weight_bit=None, input_bit=None, output_bit=None, noise_scale=None, adc_bit=None,
                    dac_bit=None, gain_noise_scale=None, offset_noise_scale=None, seed=None):
        old = (self.weight_bit, self.input_bit, self.output_bit)
        adc_old, dac_old = self.adc_bit, self.dac_bit
        self.use_FP = use_FP
        for name, val in (("weight_bit", weight_bit), ("input_bit", input_bit), ("output_bit", output_bit),
                          ("adc_bit", adc_bit), ("dac_bit", dac_bit), ("noise_scale", noise_scale),
                          ("gain_noise_scale", gain_noise_scale), ("offset_noise_scale", offset_noise_scale),
                          ("seed", seed)):
            if val is not None:
                setattr(self, name, val)
        self._gen_adc_noise()
        # update_adc_gain (layers_utils_adda.py:52-73): keep the analogue range matched to the new bit widths
        gain = self.adc_gain.data
        if adc_old != self.adc_bit:
            gain = gain * 2 ** (self.adc_bit - adc_old)
        if dac_old != self.dac_bit:
            gain = gain / 2 ** (self.dac_bit - dac_old)
        if old[0] != self.weight_bit:
            gain = gain / 2 ** (self.weight_bit - old[0])
        self.adc_gain.data = torch.clamp(gain, min=0.8 * self.adc_gain_min, max=1.2 * self.adc_gain_max)
        Q.rescale_step_sizes(self, *old)
        self.slice_bit = self.dac_bit - 1
        self.bit_slices = int(math.ceil((self.input_bit - 1) / self.slice_bit))

    # ---- ADC parameters -----------------------------------------------------------------------------------------
    def _adc_scale(self):
        """get_adc_scale (layers_utils_adda.py:40-49) on the device: clamp (straight-through), round the gain ('gain')
        or the range 1/gain (other modes), times the scale of gain 1."""
        g = self.adc_gain
        if g.is_cuda and g.dtype == torch.float32:
            mode = self.adc_adjust_mode == 'gain'
            if torch.is_grad_enabled() and g.requires_grad:
                return Q.AdcScaleFn.apply(g, self.adc_gain_min, self.adc_gain_max, mode, self.adc_gain_1_scale)
            return rt.adc_scale(g.detach(), self.adc_gain_min, self.adc_gain_max, mode, self.adc_gain_1_scale)[0]
        raise rt.B200Error("ADDA layers: adc_gain must be an fp32 CUDA parameter")

    def _maybe_init_gain(self, partial):
        """init_adc_gain (layers_utils_adda.py:12-20), fired while adc_gain still equals adc_gain_min and the block
        product is not all zero — evaluated on the device (the reference compares on the host per block)."""
        with torch.no_grad():
            amax = partial.abs().max()
            ideal = (2 ** (self.adc_bit - 1) - 1) / amax / self.adc_gain_1_scale
            ideal = torch.clamp(ideal, min=0.8 * self.adc_gain_min, max=1.2 * self.adc_gain_max)
            fire = (self.adc_gain == self.adc_gain_min) & (amax != 0)
            self.adc_gain.copy_(torch.where(fire, ideal.to(self.adc_gain.dtype), self.adc_gain))

    def _cim_rows(self, x, w2d, bias, numel_in):
        """x: bf16 rows [M, Kp] (Kp = in-features rounded up to 8, padding columns zero); w2d: fp32 parameter view
        [N, K] (shape only: the quantised weight is rebuilt from self.weight). Returns bf16 rows [M, Np] with
        Np = N rounded up to 8 (padding columns zero)."""
        M, Kp = x.shape
        N, K = w2d.shape
        Np = (N + 7) // 8 * 8
        if Kp != (K + 7) // 8 * 8:
            raise rt.B200Error("ADDA layers: input rows must be packed to in_features rounded up to 8")
        if self.input_bit > 9 or self.bit_slices * self.slice_bit > 24:
            raise rt.B200Error("ADDA layers: input codes wider than 9 bits are not supported")
        if not self.input_quant:
            raise rt.B200Error("ADDA layers: the DAC needs quantised inputs (input_quant=True)")
        qp_in = Q.qp_of(self.input_bit)
        if self._needs_lazy_init("input"):
            rt.lsq_init_step(x, self.step_size_input, qp_in)
            self._note_lazy_init("input", self.step_size_input)
        w, _ = self._w_qn(isint=True, want_scale=False)   # w_scale = 1 / step_size_weight is read on the device
        w = w.reshape(N, K)
        if Kp != K or Np != N:
            w = torch.nn.functional.pad(w, (0, Kp - K, 0, Np - N))   # zero codes meet zero digits / unused columns
            if bias is not None:
                bias = torch.nn.functional.pad(bias, (0, Np - N))
        mapping = getattr(self, "weight_mapping_info", None) or {"0_0": {"start_row": 0, "start_col": 0, "row_num": K,
                                                                          "col_num": N}}
        blocks = [(i["start_row"], i["row_num"], i["start_col"], i["col_num"]) for i in mapping.values()]
        if any(r0 % 8 for r0, _, _, _ in blocks):
            raise rt.B200Error("ADDA layers: weight blocks must start at a row multiple of 8")
        noisy = self.gain_noise_scale != 0 or self.offset_noise_scale != 0
        if noisy and max(c0 + cn for _, _, c0, cn in blocks) > 1000:
            raise rt.B200Error("ADDA layers: ADC noise tables hold 1000 columns (layers_utils_adda.py:100-105)")
        cfg = {"slices": self.bit_slices, "slice_bit": self.slice_bit, "input_bit": self.input_bit, "qp_in": qp_in,
               "numel_in": numel_in, "R": float(2 ** (self.adc_bit - 1) - 1), "blocks": blocks,
               "split_w": E.precise() and self.noise_scale != 0,
               "gain_noise": self.gain_noise.to(x.device, torch.float32) if noisy else None,
               "offset_noise": self.offset_noise.to(x.device, torch.float32) if noisy else None}
        if self.__dict__.get("_gain_init_pending", True):
            # first forward: the analogue range of the first block decides the ADC gain (init_adc_gain)
            with torch.no_grad():
                r0, rn, c0, cn = blocks[0]
                digits = Q.AddaFn._digits(x, self.step_size_input.detach(), cfg)
                first = rt.linear(digits[:, r0:r0 + rn], ops.pack_linear_weight(w.detach()[c0:c0 + cn, r0:r0 + rn]),
                                  out_dtype=torch.float32)
                self._maybe_init_gain(first)
            self.__dict__["_gain_init_pending"] = False
        alpha = self._adc_scale()
        step_w = self.step_size_weight if self.weight_quant else None
        step_out = self.step_size_output if self.output_quant else None
        single = len(blocks) == 1 and blocks[0] == (0, Kp, 0, Np) and Np % 8 == 0 and Kp == K and Np == N
        if single and FUSED_SINGLE_BLOCK and not (step_out is not None and self._needs_lazy_init("output")):
            # the whole weight is one mapped block: digits -> GEMM -> one pass (csrc/adda_out.cu: b200sd_adda_tail)
            if bias is not None and bias.dtype != torch.float32:
                bias = bias.float()
            return Q.adda_layer(self, x, self.step_size_input, w, alpha, bias, step_out, step_w, cfg, self.output_bit, M * N)
        if torch.is_grad_enabled() and any(t.requires_grad for t in (x, w, alpha, self.step_size_input)):
            acc = Q.AddaFn.apply(x, self.step_size_input, w, alpha, cfg)
        else:
            acc = Q.AddaFn._run(x, self.step_size_input.detach(), w.detach(), alpha.detach(), cfg)
        # / w_scale / in_scale / adc_scale + bias and the output quantiser: one launch (csrc/adda_out.cu)
        if bias is not None and bias.dtype != torch.float32:
            bias = bias.float()
        if step_out is not None and self._needs_lazy_init("output"):
            y = rt.adda_out(acc.detach(), None if bias is None else bias.detach(), step_w, self.step_size_input,
                            alpha.detach(), None, 0.0, M * N, 0)
            rt.lsq_init_step(y, step_out, Q.qp_of(self.output_bit))
            self._note_lazy_init("output", step_out)
        return Q.adda_out(self, acc, alpha, bias, step_out, step_w, self.step_size_input, self.output_bit, M * N)

    def _rows_in(self, rows, K):
        Kp = (K + 7) // 8 * 8
        if rows.shape[1] < Kp:
            raise rt.B200Error("ADDA layers: input rows narrower than in_features rounded up to 8")
        return rows if rows.shape[1] == Kp else rows[:, :Kp].contiguous()



class Linear_lsq_adda_cim(nn.Linear, _AddaMixin):
    def __init__(self, in_features, out_features, weight_bit, input_bit, output_bit, noise_scale, dac_bit, adc_bit,
                 adc_gain_1_scale, adc_gain_range, adc_adjust_mode='gain', input_quant=True, output_quant=True,
                 weight_quant=True, gain_noise_scale=0, offset_noise_scale=0, seed=0, bias=True):
        super().__init__(in_features, out_features, bias)
        self._init_lsq(weight_bit, input_bit, output_bit, noise_scale, input_quant, output_quant, weight_quant,
                       gain_noise_scale, offset_noise_scale)
        self._init_adda(dac_bit, adc_bit, adc_gain_1_scale, adc_gain_range, adc_adjust_mode, seed)

    def forward_rows(self, rows, act=None):
        K, N = self.in_features, self.out_features
        if self.use_FP:
            return self._linear(rows, self.weight, act)
        if N % 8:
            raise rt.B200Error("Linear_lsq_adda_cim: out_features must be a multiple of 8")
        x = self._rows_in(rows, K)
        y = self._cim_rows(x, self.weight, self.bias, x.shape[0] * K)
        if act is not None:
            if act != "relu":
                raise rt.B200Error("Linear_lsq_adda_cim: only ReLU may follow a quantised Linear")
            y = torch.relu(y) if E.training() else rt.copy_act(y, torch.empty_like(y), 0, act="relu")
        return y

    def forward(self, x):
        if not x.is_cuda:
            raise rt.B200Error("Linear_lsq_adda_cim.forward: CUDA tensors required (no CPU implementation in this package)")
        lead = x.shape[:-1]
        with torch.autocast("cuda", enabled=False):
            rows = x.reshape(-1, x.shape[-1])
            rows = rows.to(torch.bfloat16) if rows.dtype != torch.bfloat16 else rows
            if rows.shape[1] % 8:
                rows = torch.nn.functional.pad(rows, (0, 8 - rows.shape[1] % 8))
            out = self.forward_rows(rows.contiguous())
        return out.reshape(*lead, self.out_features).to(x.dtype)


class Conv2d_lsq_adda_cim(nn.Conv2d, _AddaMixin):
    def __init__(self, in_channels, out_channels, kernel_size, stride, padding, groups, weight_bit, input_bit,
                 output_bit, noise_scale, dac_bit, adc_bit, adc_gain_1_scale, adc_gain_range, adc_adjust_mode='gain',
                 input_quant=True, output_quant=True, weight_quant=True, gain_noise_scale=0, offset_noise_scale=0,
                 seed=0, bias=True):
        super().__init__(in_channels=in_channels, out_channels=out_channels, kernel_size=kernel_size, stride=stride,
                         padding=padding, groups=groups, bias=bias)
        self._init_lsq(weight_bit, input_bit, output_bit, noise_scale, input_quant, output_quant, weight_quant,
                       gain_noise_scale, offset_noise_scale)
        self._init_adda(dac_bit, adc_bit, adc_gain_1_scale, adc_gain_range, adc_adjust_mode, seed)

    def forward_act(self, a):
        if self.use_FP:
            raise rt.B200Error("Conv2d_lsq_adda_cim: use_FP bypass is served by Conv2d_qn_lsq / nn.Conv2d")
        k, st, pd = self.kernel_size, self.stride, self.padding
        if isinstance(pd, str) or self.groups != 1 or self.dilation != (1, 1) or k[0] != k[1] or st[0] != st[1] \
                or pd[0] != pd[1]:
            raise rt.B200Error("Conv2d_lsq_adda_cim: square kernels with symmetric stride / padding, groups = 1 only")
        if k == (1, 1) and st == (1, 1) and pd == (0, 0):
            y = self._cim_rows(self._rows_in(a.t, a.c), self.weight.reshape(self.out_channels, -1), self.bias,
                               a.n * a.hw * a.c)
            return E.Act(y, a.n, a.h, a.w, self.out_channels)
        geom = (a.n, a.h, a.w, a.c, k[0], st[0], pd[0])
        oh, ow = (a.h + 2 * pd[0] - k[0]) // st[0] + 1, (a.w + 2 * pd[0] - k[0]) // st[0] + 1
        if torch.is_grad_enabled() and a.t.requires_grad:
            rows = _Im2colFn.apply(a.t, geom)
        else:
            rows = rt.im2col(a.t, *geom)
        y = self._cim_rows(rows, self.weight.reshape(self.out_channels, -1), self.bias, a.n * a.hw * a.c)
        return E.Act(y, a.n, oh, ow, self.out_channels)

    def forward(self, x):
        if not x.is_cuda:
            raise rt.B200Error("Conv2d_lsq_adda_cim.forward: CUDA tensors required (no CPU implementation in this package)")
        with torch.autocast("cuda", enabled=False):
            out = self.forward_act(E.Act.from_nchw(x)).to_nchw()
        return out.to(x.dtype)


_rearm_on_load(Linear_lsq_adda_cim)
_rearm_on_load(Conv2d_lsq_adda_cim)

"""Enhance wrappers of the CIM quant-noise pipeline (reference cim_qn_train/layers_enhance.py) on the sm_100a engine.

The reference wraps every quantised Linear of the DiT in two ways (cim_qn_train/progressive_qn_train.py:884-950):
`EnhanceBranchLinear_LoR` adds a low-rank, grouped side branch blended in through a learned sigmoid factor
(layers_enhance.py:556-621), `EnhanceLayerLinear` appends a grouped Linear (:413-436); both are built from
`GroupedLinear` (:324-365), a block-diagonal Linear evaluated with an einsum. Here a GroupedLinear is ONE tcgen05 GEMM
over the block-diagonal dense form of its (groups, out_group, in_group) parameter — assembled with differentiable torch
ops, so autograd hands the per-group gradient back — and the wrappers expose the engine's leaf protocol
(`forward_rows`: bf16 token rows in, bf16 rows out), so a DiT whose leaves were swapped for them keeps running on the
fused path (sdb200.engine.linear). Module names, parameter shapes, initialisation and the group/rank search
(`analyze_*`) follow the reference, so state_dicts are interchangeable.

Built: the Linear family (what the DiT-9L pipeline instantiates for its 95 Linear leaves), EnhanceLayerConv2d (any
original conv the engine or a cim_layers drop-in runs) and EnhanceBranchConv2d_LoR around any square-kernel original
conv (the DiT's 1x1 mask conv; the ReLU-UNet's 3x3 and 4x4-stride-2 leaves with the reference's padded-1x1 frame,
average pooling and top-left crop).
"""
import math

import torch
import torch.nn as nn

from sdb200 import autograd as AG
from sdb200 import engine as E
from sdb200 import ops
from sdb200 import runtime as rt


# ------------------------------------------------------------------------------------------------ group / rank search
def _divisors(x):
    small = [d for d in range(1, math.isqrt(x) + 1) if x % d == 0]
    return sorted(set(small + [x // d for d in small]))


def _nearest(target, candidates):
    """First candidate (in the given order) at minimal distance from target (layers_enhance.py:22-24)."""
    return min(candidates, key=lambda d: abs(d - target))


def ideal_hidden_features(F_in, F_out, ops_factor):
    """Rank h at which an ungrouped two-layer branch costs ops_factor^(1/3) of the Linear (layers_enhance.py:43-50)."""
    return (ops_factor ** (1 / 3) * F_in * F_out) / (F_in + F_out)


def ideal_hidden_channels(C_in, C_out, k_h, k_w, ops_factor):
    """Conv counterpart (layers_enhance.py:27-40)."""
    return max(1, round(ops_factor ** (1 / 3) * C_in * C_out * (k_h * k_w) / (C_in + C_out)))


def _search_lora(n_in, n_out, ops_orig, hid_ideal, ops_factor, key):
    """Joint search of (group1, group2, hidden) shared by the Linear and Conv variants (layers_enhance.py:53-131,
    134-207): over hidden sizes in [0.1, 10] x ideal, the groups closest to 1 / ops_factor^(1/3) among the common
    divisors; lexicographic minimum of (|cost ratio - ops_factor|, imbalance of the three compression shares), first
    hidden size winning ties."""
    base = ops_factor ** (1 / 3)
    d_in, d_out = set(_divisors(n_in)), set(_divisors(n_out))
    best = None
    for h in range(max(1, int(hid_ideal * 0.1)), max(1, int(hid_ideal * 10)) + 1):
        d_h = _divisors(h)
        g1 = _nearest(1 / base, [d for d in d_h if d in d_in])
        g2 = _nearest(1 / base, [d for d in d_h if d in d_out])
        factor = (n_in * h // g1 + h * n_out // g2) / ops_orig
        shares = ((n_in * h + h * n_out) / ops_orig,
                  (n_in * hid_ideal / g1 + hid_ideal * n_out) / ops_orig,
                  (n_in * hid_ideal + hid_ideal * n_out / g2) / ops_orig)
        score = (abs(factor - ops_factor),
                 math.sqrt((shares[0] - base) ** 2 + (shares[1] - base) ** 2 + (shares[2] - base) ** 2))
        if best is None or score < best[0]:
            best = (score, {"group1": g1, "group2": g2, key: h, "estimated_ops_factor": factor})
    return None if best is None else best[1]


def analyze_lora_linear_ops(in_features, out_features, ops_factor):
    hid = ideal_hidden_features(in_features, out_features, ops_factor)
    cfg = _search_lora(in_features, out_features, in_features * out_features, hid, ops_factor, "hidden_features")
    if cfg is None:
        h = max(1, round(hid))
        cfg = {"group1": 1, "group2": 1, "hidden_features": h,
               "estimated_ops_factor": (in_features * h + h * out_features) / (in_features * out_features)}
    return cfg


def analyze_lora_conv_ops(in_channels, out_channels, kh, kw, ops_factor):
    ops_orig = in_channels * out_channels * kh * kw
    hid = ideal_hidden_channels(in_channels, out_channels, kh, kw, ops_factor)
    cfg = _search_lora(in_channels, out_channels, ops_orig, hid, ops_factor, "hidden_channels")
    if cfg is None:
        h = max(1, int(hid))
        cfg = {"group1": 1, "group2": 1, "hidden_channels": h,
               "estimated_ops_factor": (in_channels * h + h * out_channels) / ops_orig}
    return cfg


def _smallest_group(width, ops_orig, ops_factor, fallback):
    """Smallest divisor g of `width` whose grouped width x width map costs at most ops_factor of the original layer."""
    for g in _divisors(width):
        ratio = (width * width // g) / ops_orig
        if ratio <= ops_factor:
            return {"group": g, "estimated_ops_factor": ratio}
    return {"group": width, "estimated_ops_factor": fallback}


def analyze_enhance_conv_ops(in_channels, out_channels, kh, kw, ops_factor):
    """layers_enhance.py:210-235."""
    return _smallest_group(out_channels, in_channels * out_channels * kh * kw, ops_factor, 1.0 / (kh * kw))


def analyze_enhance_linear_ops(in_features, out_features, ops_factor):
    """layers_enhance.py:238-261."""
    return _smallest_group(out_features, in_features * out_features, ops_factor, 1.0)


# ------------------------------------------------------------------------------------------------ modules
def _pad8(n):
    return (n + 7) // 8 * 8


def _as_rows(x):
    rows = x.reshape(-1, x.shape[-1])
    rows = rows if rows.dtype == torch.bfloat16 else rows.to(torch.bfloat16)
    if rows.shape[1] % 8:
        rows = torch.nn.functional.pad(rows, (0, 8 - rows.shape[1] % 8))
    return rows.contiguous()


class _RowsModule(nn.Module):
    """forward(x) of a leaf that implements forward_rows: (..., in) -> (..., out) in x's dtype; CUDA only."""

    def forward(self, x):
        if not x.is_cuda:
            raise rt.B200Error(f"{type(self).__name__}.forward: CUDA tensors required (this package has no CPU "
                               "implementation; the reference itself is the CPU path)")
        with torch.autocast("cuda", enabled=False):
            out = self.forward_rows(_as_rows(x))
        return out[:, :self.out_features].reshape(*x.shape[:-1], self.out_features).to(x.dtype)


class GroupedLinear(_RowsModule):
    """Block-diagonal Linear: `groups` independent (in/groups -> out/groups) maps (layers_enhance.py:324-365). Parameters
    keep the reference's shapes — weight (groups, out_group, in_group) initialised to truncated identities, bias
    (groups, out_group) zeros — the product runs as one GEMM over the dense block-diagonal matrix."""

    def __init__(self, in_features, out_features, groups, dtype=torch.float32, device=None):
        super().__init__()
        assert in_features % groups == 0, "in_features must be divisible by groups"
        assert out_features % groups == 0, "out_features must be divisible by groups"
        self.in_features, self.out_features, self.groups = in_features, out_features, groups
        self.in_group, self.out_group = in_features // groups, out_features // groups
        eye = torch.eye(self.out_group, self.in_group, device=device, dtype=dtype)
        self.weight = nn.Parameter(eye.expand(groups, -1, -1).clone())
        self.bias = nn.Parameter(torch.zeros(groups, self.out_group, device=device, dtype=dtype))

    def dense_weight(self):
        """[out_features, in_features rounded up to 8] block-diagonal matrix, differentiable w.r.t. `weight`."""
        return _block_diag_dense(self, self.weight, self.groups, self.out_group, self.in_group)

    def forward_rows(self, rows, act=None):
        """bf16 rows [M, >= in_features rounded up to 8] (zeros beyond in_features) -> bf16 rows [M, out_features rounded
        up to 8], padding columns zero."""
        return _grouped_rows(self, rows, self.weight, self.bias, self.out_features, act)


def _dense_rows(rows, w_dense, bias, n_out, act=None):
    """rows [M, Kp] (zeros beyond the true K) @ w_dense[n_out, Kp]^T + bias -> [M, n_out rounded up to 8], padding
    columns zero; one tcgen05 GEMM (autograd Function in training mode)."""
    if rows.shape[1] < w_dense.shape[1]:
        raise rt.B200Error("grouped map: input rows narrower than the channel count rounded up to 8")
    x = rows if rows.shape[1] == w_dense.shape[1] else rows[:, :w_dense.shape[1]].contiguous()
    fuse = act if n_out % 8 == 0 else None   # (a fused ReLU's backward wants an unpadded, contiguous output)
    if E.training():
        y, _ = AG.LinearFn.apply(x, w_dense, bias, None, None, False, fuse)
    else:
        y = rt.linear(x, ops.pack_linear_weight(w_dense.detach()), bias=None if bias is None else bias.detach().float(),
                      act=fuse)
    if n_out % 8:
        y = torch.nn.functional.pad(y[:, :n_out], (0, _pad8(n_out) - n_out))
    if act is not None and fuse is None:
        if act != "relu":
            raise rt.B200Error("grouped map: only ReLU may follow")
        y = torch.relu(y)
    return y


class _GroupedFn(torch.autograd.Function):
    """y = x @ blockdiag(W)^T + b for a (g, og, ig) parameter W: the bf16 dense matrix and its transpose live in
    persistent zero-initialised buffers of `owner` whose diagonal blocks one launch rewrites (b200sd_blockdiag_pack);
    the backward reads the parameter gradient back out of the dense wgrad (b200sd_blockdiag_gather)."""

    @staticmethod
    def forward(ctx, x, weight, bias, owner, act):
        g, og, ig = weight.shape
        n_out = g * og
        dense, dense_t = _dense_buffers(owner, weight)
        rt.blockdiag_pack(weight.detach(), dense, dense_t)
        out = torch.empty(x.shape[0], _pad8(n_out), device=x.device, dtype=torch.bfloat16)[:, :n_out]
        rt.linear(x, dense, bias=None if bias is None else bias.detach().reshape(-1).float(), out=out, act=act)
        ctx.save_for_backward(x, weight, out if act else None)
        ctx.owner, ctx.bias_like = owner, None if bias is None else (bias.shape, bias.dtype)
        return out

    @staticmethod
    def backward(ctx, dy):
        x, weight, y_act = ctx.saved_tensors
        g, og, ig = weight.shape
        n_out, kp = g * og, x.shape[1]
        M = x.shape[0]
        dy = AG._rows(dy, n_out)
        if y_act is not None:
            dy = rt.relu_bwd(dy, y_act)
        dx = dw = db = None
        want_b = ctx.bias_like is not None and ctx.needs_input_grad[2]
        # parameter gradients as a branch beside the input-gradient GEMM (runtime.Side)
        with rt.Side(ctx.needs_input_grad[0] and (ctx.needs_input_grad[1] or want_b), x.device) as side:
            if ctx.needs_input_grad[1]:
                k64 = (kp + 63) // 64 * 64
                dwp = rt.wgrad([rt.make_src(x, 1, 1, M, kp, ld=x.stride(0))], [(0, 0, 0, 0, k64 // 64)],
                               rt.make_src(dy, 1, 1, M, n_out, ld=dy.stride(0)), n_out, 1, 1, M)
                dw = rt.blockdiag_gather(dwp, g, og, ig).to(weight.dtype)
            if want_b:
                db = rt.colsum(dy, 1, M, n_out) if n_out % 8 == 0 else dy.float().sum(0)
                db = db.reshape(ctx.bias_like[0]).to(ctx.bias_like[1])
        if ctx.needs_input_grad[0]:
            _, dense_t = _dense_buffers(ctx.owner, weight)      # written by this step's forward
            dx = torch.empty(M, kp, device=x.device, dtype=torch.bfloat16)
            rt.linear(dy, dense_t, out=dx)
        side.join()
        return dx, dw, db, None, None


def _dense_buffers(owner, weight):
    """Persistent bf16 (dense [g*og, pad64(pad8(g*ig))], dense_t [pad8(g*ig), pad64(g*og)]) of a grouped map, zero
    outside the diagonal blocks (only those are ever rewritten)."""
    g, og, ig = weight.shape
    key = (weight.device, g, og, ig)
    ent = owner.__dict__.get("_dense_bf16")
    if ent is None or ent[0] != key:
        kp = _pad8(g * ig)
        dense = torch.zeros(g * og, (kp + 63) // 64 * 64, device=weight.device, dtype=torch.bfloat16)
        dense_t = torch.zeros(kp, (g * og + 63) // 64 * 64, device=weight.device, dtype=torch.bfloat16)
        ent = owner.__dict__["_dense_bf16"] = (key, dense, dense_t)
    return ent[1], ent[2]


def _grouped_rows(owner, rows, weight3, bias, n_out, act=None):
    """rows [M, >= pad8(K)] (zeros beyond the true K) through the block-diagonal map of the (g, og, ig) parameter
    `weight3` -> [M, n_out rounded up to 8], padding columns zero; one scatter launch + one tcgen05 GEMM."""
    g, og, ig = weight3.shape
    kp = _pad8(g * ig)
    if weight3.dtype != torch.float32 or not weight3.is_cuda:
        return _dense_rows(rows, _block_diag_dense(owner, weight3, g, og, ig), None if bias is None else bias.reshape(-1),
                           n_out, act)
    if rows.shape[1] < kp:
        raise rt.B200Error("grouped map: input rows narrower than the channel count rounded up to 8")
    x = rows if rows.shape[1] == kp else rows[:, :kp].contiguous()
    fuse = act if n_out % 8 == 0 else None   # (a fused ReLU's backward wants an unpadded, contiguous output)
    if E.training():
        y = _GroupedFn.apply(x, weight3, bias, owner, fuse)
    else:
        dense, _ = _dense_buffers(owner, weight3)
        rt.blockdiag_pack(weight3.detach(), dense)
        y = rt.linear(x, dense, bias=None if bias is None else bias.detach().reshape(-1).float(), act=fuse)
    if n_out % 8:
        y = torch.nn.functional.pad(y[:, :n_out], (0, _pad8(n_out) - n_out))
    if act is not None and fuse is None:
        if act != "relu":
            raise rt.B200Error("grouped map: only ReLU may follow")
        y = torch.relu(y)
    return y


def _conv_rows(conv, rows, act=None):
    """A grouped 1x1 nn.Conv2d on pixel rows (its [co, ci/g, 1, 1] weight is the (g, co/g, ci/g) block parameter)."""
    g = conv.groups
    return _grouped_rows(conv, rows, conv.weight.view(g, conv.out_channels // g, conv.in_channels // g), conv.bias,
                         conv.out_channels, act)


def _block_diag_dense(owner, weight, g, og, ig):
    """Dense [g*og, g*ig rounded up to 8] block-diagonal matrix of a (g, og, ig)-shaped parameter as ONE scatter into a
    zero matrix through a cached flat index (torch.block_diag costs a kernel per block: 96 for the DiT's widest enhance
    layer); differentiable, the backward is one gather."""
    kp = _pad8(g * ig)
    cache = owner.__dict__.get("_dense_index")
    if cache is None or cache.device != weight.device:
        gi = torch.arange(g, device=weight.device).view(g, 1, 1)
        r = torch.arange(og, device=weight.device).view(1, og, 1)
        c = torch.arange(ig, device=weight.device).view(1, 1, ig)
        cache = ((gi * og + r) * kp + gi * ig + c).reshape(-1)
        owner.__dict__["_dense_index"] = cache
    flat = torch.zeros(g * og * kp, device=weight.device, dtype=torch.float32)
    return flat.index_put((cache,), weight.float().reshape(-1)).view(g * og, kp)


def _conv1x1_dense(conv):
    """Block-diagonal dense [out, in rounded up to 8] matrix of a grouped 1x1 nn.Conv2d, differentiable."""
    g, co, ci = conv.groups, conv.out_channels, conv.in_channels
    return _block_diag_dense(conv, conv.weight, g, co // g, ci // g)


def _clean_rows(a):
    """Rows of an Act with the columns beyond its channel count zeroed (a producer may leave them unwritten)."""
    cp = _pad8(a.c)
    if a.c % 8 == 0:
        return a.t if a.t.shape[1] == cp else a.t[:, :cp].contiguous()
    return torch.nn.functional.pad(a.t[:, :a.c], (0, cp - a.c))


class _ActModule(nn.Module):
    """forward(x) of a leaf that implements forward_act: NCHW in, NCHW out in x's dtype; CUDA only."""

    def forward(self, x):
        if not x.is_cuda:
            raise rt.B200Error(f"{type(self).__name__}.forward: CUDA tensors required (this package has no CPU "
                               "implementation; the reference itself is the CPU path)")
        with torch.autocast("cuda", enabled=False):
            out = self.forward_act(E.Act.from_nchw(x)).to_nchw()
        return out.to(x.dtype)


class EnhanceLayerConv2d(_ActModule):
    """original conv followed by a grouped 1x1 convolution (out -> out) initialised to the identity
    (layers_enhance.py:368-410). The original conv runs on its own path (a cim_layers drop-in or a plain nn.Conv2d the
    engine supports); the grouped 1x1 map is one GEMM over the pixel rows."""

    def __init__(self, original_conv, ops_factor=0.05):
        super().__init__()
        self.original_conv = original_conv
        co, ci = original_conv.out_channels, original_conv.in_channels
        kh, kw = original_conv.kernel_size
        w = original_conv.weight
        self.groups = analyze_enhance_conv_ops(ci, co, kh, kw, ops_factor)["group"]
        self.conv_enhance = nn.Conv2d(co, co, kernel_size=1, stride=1, padding=0, groups=self.groups, device=w.device,
                                      dtype=w.dtype)
        self.conv_enhance.layer_flag, self.conv_enhance.distance = 'enhance_layer', 1
        with torch.no_grad():
            eye = torch.eye(co // self.groups, dtype=w.dtype, device=w.device)
            self.conv_enhance.weight.copy_(eye.repeat(self.groups, 1).view(co, co // self.groups, 1, 1))
            self.conv_enhance.bias.fill_(0)

    def forward_act(self, a):
        y = E.conv2d(a, self.original_conv)
        rows = _conv_rows(self.conv_enhance, _clean_rows(y))
        return E.Act(rows, y.n, y.h, y.w, y.c)


class EnhanceBranchConv2d_LoR(_ActModule):
    """original conv blended with a two-layer grouped 1x1 low-rank branch (layers_enhance.py:439-535).

    The two branch convolutions are 1x1 but carry the ORIGINAL conv's padding (:467,:473), so for a padded original the
    branch output is an (H + 4p) x (W + 4p) map whose interior is the per-pixel map z = W2 act(W1 x + b1) + b2 shifted by
    2p, framed by a ring of the constant W2 act(b1) + b2 (pixels whose first conv saw only padding) and an outer ring of
    b2 (pixels whose second conv saw only padding); a strided original average-pools that map by its stride (:484-487);
    the reference then crops from the TOP-LEFT corner (or zero-pads around) to the original's output size (:520-531).
    Here z is two grouped GEMMs on the pixel rows (one tcgen05 GEMM each over the block-diagonal dense weights); the
    frame, the pooling and the crop are assembled exactly as described."""

    def __init__(self, original_conv, ops_factor=0.05, relu=False, sigmoid=True):
        super().__init__()
        kh, kw = original_conv.kernel_size
        st, pd = tuple(original_conv.stride), original_conv.padding
        if isinstance(pd, str) or st[0] != st[1] or pd[0] != pd[1]:
            raise NotImplementedError("EnhanceBranchConv2d_LoR: symmetric integer stride / padding only")
        self.original_conv = original_conv
        ci, co = original_conv.in_channels, original_conv.out_channels
        self.stride, self.padding = original_conv.stride, original_conv.padding
        self.relu, self.sigmoid = relu, sigmoid
        w = original_conv.weight
        cfg = analyze_lora_conv_ops(ci, co, kh, kw, ops_factor)
        hidden = cfg["hidden_channels"]
        self.enhance_branch_1 = nn.Conv2d(ci, hidden, kernel_size=1, stride=1, padding=self.padding,
                                          groups=cfg["group1"], dtype=w.dtype, device=w.device)
        self.enhance_branch_2 = nn.Conv2d(hidden, co, kernel_size=1, stride=1, padding=self.padding,
                                          groups=cfg["group2"], dtype=w.dtype, device=w.device)
        self.enhance_branch_1.layer_flag = self.enhance_branch_2.layer_flag = 'enhance_branch'
        self.pool = nn.AvgPool2d(kernel_size=self.stride, stride=self.stride, padding=0) if any(v != 1 for v in st) else None
        self.blend_factor = nn.Parameter(torch.tensor(-6.9068 if sigmoid else 0.0))
        self.std_ori = self.mean_ori = self.std_enhance = self.mean_enhance = None

    merge_feature_map = None   # bound below (shared with the Linear wrapper)

    def forward_act(self, a):
        y = E.conv2d(a, self.original_conv)
        b1, b2 = self.enhance_branch_1, self.enhance_branch_2
        enh = _conv_rows(b1, _clean_rows(a), act="relu" if self.relu else None)
        enh = _conv_rows(b2, enh)
        p, st = self.padding[0], self.stride[0]
        if p != 0 or st != 1 or (y.h, y.w) != (a.h, a.w):
            enh = self._frame(enh[:, :y.c], a, y, p, st)
        return E.Act(self._merge_rows(y.t, enh, y.c), y.n, y.h, y.w, y.c)


def _frame(self, z, a, y, p, st):
    """Branch output of a padded / strided original conv from the per-pixel map z [n*h*w, co] (see the class docstring)."""
    b1, b2 = self.enhance_branch_1, self.enhance_branch_2
    co, n, h, w = y.c, a.n, a.h, a.w
    f32 = torch.float32
    bias2 = b2.bias.to(f32) if b2.bias is not None else torch.zeros(co, device=z.device, dtype=f32)
    bias1 = b1.bias.to(f32) if b1.bias is not None else torch.zeros(b1.out_channels, device=z.device, dtype=f32)
    hid = torch.relu(bias1) if self.relu else bias1
    ring = torch.nn.functional.linear(hid, _conv1x1_dense(b2).to(f32)[:, :b2.in_channels], bias2)   # W2 act(b1) + b2
    e = bias2.view(1, 1, 1, co).expand(n, h + 4 * p, w + 4 * p, co).clone()
    if p > 0:
        e[:, p:p + h + 2 * p, p:p + w + 2 * p, :] = ring
    e[:, 2 * p:2 * p + h, 2 * p:2 * p + w, :] = z.to(f32).reshape(n, h, w, co)
    if st != 1:
        e = torch.nn.functional.avg_pool2d(e.permute(0, 3, 1, 2), kernel_size=st, stride=st, padding=0).permute(0, 2, 3, 1)
    dy, dx = y.h - e.shape[1], y.w - e.shape[2]
    if dy != 0 or dx != 0:
        if dy < 0 or dx < 0:
            e = e[:, :y.h, :y.w, :]
        elif dy > 0 or dx > 0:
            e = torch.nn.functional.pad(e, (0, 0, dx // 2, dx - dx // 2, dy // 2, dy - dy // 2))
    return e.reshape(n * y.h * y.w, co).to(torch.bfloat16)


EnhanceBranchConv2d_LoR._frame = _frame


def _leaf_rows(mod, rows):
    """The wrapped Linear (plain nn.Linear or a cim_layers drop-in) on bf16 rows, through the engine's leaf dispatch."""
    return E.linear(rows, mod)


class EnhanceLayerLinear(_RowsModule):
    """original Linear followed by a GroupedLinear(out, out) initialised to the identity (layers_enhance.py:413-436)."""

    def __init__(self, original_linear, ops_factor=0.05):
        super().__init__()
        self.original_linear = original_linear
        self.in_features, self.out_features = original_linear.in_features, original_linear.out_features
        w = original_linear.weight
        cfg = analyze_enhance_linear_ops(self.in_features, self.out_features, ops_factor)
        self.fc1 = GroupedLinear(self.out_features, self.out_features, cfg["group"], dtype=w.dtype, device=w.device)

    def forward_rows(self, rows, act=None):
        return self.fc1.forward_rows(_leaf_rows(self.original_linear, rows), act=act)


class EnhanceBranchLinear_LoR(_RowsModule):
    """original Linear blended with a two-layer grouped low-rank branch (layers_enhance.py:556-621):
    out = orig * (1 - s) + branch_2(relu?(branch_1(x))) * s,  s = sigmoid(blend_factor) (or blend_factor itself).
    The blend starts at sigmoid(-6.9068) ~ 1e-3. std/mean of both operands are kept as the reference does (:605-608;
    progressive_qn_train.py:993-996 reads them)."""

    def __init__(self, original_linear, ops_factor=0.05, relu=False, sigmoid=True):
        super().__init__()
        self.original_linear = original_linear
        self.in_features, self.out_features = original_linear.in_features, original_linear.out_features
        self.relu, self.sigmoid = relu, sigmoid
        w = original_linear.weight
        cfg = analyze_lora_linear_ops(self.in_features, self.out_features, ops_factor)
        hidden = cfg["hidden_features"]
        self.enhance_branch_1 = GroupedLinear(self.in_features, hidden, cfg["group1"], dtype=w.dtype, device=w.device)
        self.enhance_branch_2 = GroupedLinear(hidden, self.out_features, cfg["group2"], dtype=w.dtype, device=w.device)
        self.blend_factor = nn.Parameter(torch.tensor(-6.9068 if sigmoid else 0.0))
        for br in (self.enhance_branch_1, self.enhance_branch_2):
            br.layer_flag, br.distance = 'enhance_branch', 1
        self.std_ori = self.mean_ori = self.std_enhance = self.mean_enhance = None

    def merge_feature_map(self, original_output, enhance_output):
        s = torch.sigmoid(self.blend_factor) if self.sigmoid else self.blend_factor
        # orig * (1 - s) + enh * s as one fused kernel (forward and backward): orig + s * (enh - orig)
        combined = torch.lerp(original_output, enhance_output, s.to(original_output.dtype))
        with torch.no_grad():   # one fused reduction per operand (the reference: four separate ones, :605-608)
            self.std_ori, self.mean_ori = torch.std_mean(original_output)
            self.std_enhance, self.mean_enhance = torch.std_mean(enhance_output)
        return combined

    def _merge_rows(self, orig, enh, n, act=None):
        """merge_feature_map on bf16 rows -> bf16 rows [M, n rounded up to 8]: one launch (blend + the four moments the
        trainer logs; backward one launch) when the rows are 8-column aligned, the torch formulation otherwise."""
        if act not in (None, "relu"):
            raise rt.B200Error(f"{type(self).__name__}: only ReLU may follow the wrapped layer")
        fused = n % 8 == 0 and orig.dtype == enh.dtype == torch.bfloat16 and orig.stride(1) == enh.stride(1) == 1 \
            and orig.stride(0) % 8 == 0 and enh.stride(0) % 8 == 0 and self.blend_factor.dtype == torch.float32
        if fused:
            if torch.is_grad_enabled() and any(t.requires_grad for t in (orig, enh, self.blend_factor)):
                return _BlendFn.apply(orig, enh, self.blend_factor, self, act == "relu", n)
            out, stats = rt.blend(orig, enh, self.blend_factor, self.sigmoid, act == "relu", n)
            self.std_ori, self.mean_ori, self.std_enhance, self.mean_enhance = stats.unbind(0)
            return out
        out = self.merge_feature_map(orig[:, :n], enh[:, :n]).to(torch.bfloat16)
        if act is not None:
            out = torch.relu(out)
        if n % 8:
            out = torch.nn.functional.pad(out, (0, _pad8(n) - n))
        return out.contiguous()

    def forward_rows(self, rows, act=None):
        orig = _leaf_rows(self.original_linear, rows)
        enh = self.enhance_branch_1.forward_rows(rows, act="relu" if self.relu else None)
        enh = self.enhance_branch_2.forward_rows(enh)
        return self._merge_rows(orig, enh, self.out_features, act)


class _BlendFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, orig, enh, factor, owner, relu, n):
        out, stats = rt.blend(orig, enh, factor, owner.sigmoid, relu, n)
        owner.std_ori, owner.mean_ori, owner.std_enhance, owner.mean_enhance = stats.unbind(0)
        ctx.save_for_backward(orig, enh, factor, out if relu else None)
        ctx.cfg = (owner.sigmoid, n)
        return out

    @staticmethod
    def backward(ctx, g):
        orig, enh, factor, out = ctx.saved_tensors
        sigmoid, n = ctx.cfg
        d_orig, d_enh, d_f = rt.blend_bwd(AG._rows(g, n), orig, enh, out, factor, sigmoid, n)
        if orig.shape[1] != n:
            d_orig = torch.nn.functional.pad(d_orig, (0, orig.shape[1] - n))
        if enh.shape[1] != n:
            d_enh = torch.nn.functional.pad(d_enh, (0, enh.shape[1] - n))
        return d_orig, d_enh, d_f.reshape(factor.shape).to(factor.dtype), None, None, None


EnhanceBranchConv2d_LoR.merge_feature_map = EnhanceBranchLinear_LoR.merge_feature_map
EnhanceBranchConv2d_LoR._merge_rows = EnhanceBranchLinear_LoR._merge_rows


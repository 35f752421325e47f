"""Fused execution engine: runs the reference's block structure on the sm_100a kernels.

The engine never owns a model definition. It reads the LIVE torch module tree every call (leaf types, channel
counts, parameters) so that callers which swap leaves — the cim_qn_train module surgery, SURVEY §7 "the callers do
module surgery" — change what executes. Leaves of exact type nn.Conv2d / nn.Linear / nn.GroupNorm /
nn.MultiheadAttention run on the fused kernels; a leaf exposing `forward_act(act)` (the cim_layers drop-ins)
supplies its own CUDA path; anything else is an error (there is no PyTorch fallback).

Activations travel as `Act`: NHWC bf16, a 2-D [n*h*w, ld] tensor — the conv "pixel x channel" matrix and the
attention (B, L, C) token matrix are the same memory, so no transposes exist anywhere on the path
(the reference does reshape/transpose round trips around every attention, blocks.py:125-131).
"""
import weakref

import torch
import torch.nn as nn

from . import autograd as AG
from . import ops
from . import runtime as rt


def training():
    """Autograd is recording: leaves run through the autograd.Function wrappers (autograd.py), whose backward is
    also on the sm_100a kernels. Inference (torch.no_grad) takes the direct path with cached packed weights."""
    return torch.is_grad_enabled()


def _st(stats_t, p_c):
    """(stats tensor from an autograd Function, or empty) -> Act.stats list."""
    if stats_t is None or stats_t.numel() == 0:
        return None
    return [(stats_t, stats_t.shape[1], stats_t.shape[2], p_c)]


FUSE_GN_STATS = True  # producers leave partial GroupNorm statistics for their consumer (tests flip this)

# "fast": bf16 activations / weights everywhere (what bench.py measures). "precise": fp32 residual stream + split bf16
# operands (precise.py) — whole-network noise prediction within 1e-2 (measured ~2e-3) of the reference's fp32 output,
# inference only. Chosen per process (B200SD_PARITY_MODE) or with set_parity_mode().
import os as _os
PARITY_MODE = _os.environ.get("B200SD_PARITY_MODE", "fast")


def set_parity_mode(mode):
    global PARITY_MODE
    if mode not in ("fast", "precise"):
        raise ValueError("parity mode must be 'fast' or 'precise'")
    PARITY_MODE = mode


def precise():
    """Precise mode applies to inference; autograd-recording passes always run the fast (bf16) training path."""
    return PARITY_MODE == "precise" and not torch.is_grad_enabled()


def _P():
    from . import precise as P
    return P


class Act:
    """stats: None, or a list of (fp32 [n, P, ld, 2] partial sums, P, ld, channels) segments covering the channels in
    order — written by the igemm epilogue(s) that produced this tensor, consumed by the next GroupNorm."""
    __slots__ = ("t", "n", "h", "w", "c", "stats")

    def __init__(self, t, n, h, w, c, stats=None):
        self.t, self.n, self.h, self.w, self.c, self.stats = t, n, h, w, c, stats

    @property
    def ld(self):
        return self.t.stride(0)

    @property
    def hw(self):
        return self.h * self.w

    @staticmethod
    def empty(n, h, w, c, device):
        return Act(torch.empty(n * h * w, c, device=device, dtype=torch.bfloat16), n, h, w, c)

    @staticmethod
    def from_nchw(x):
        n, c, h, w = x.shape
        ld = (c + 7) // 8 * 8
        a = Act(torch.empty(n * h * w, ld, device=x.device, dtype=torch.bfloat16), n, h, w, c)
        if ld != c:
            a.t.zero_()
        if training() and x.requires_grad:
            a.t = AG.FromNchwFn.apply(x, ld)
            return a
        rt.nchw_to_nhwc(x.float(), a.t, 0)
        return a

    def to_nchw(self):
        if training() and self.t.requires_grad:
            return AG.ToNchwFn.apply(self.t, (self.n, self.c, self.h, self.w))
        return rt.nhwc_to_nchw(self.t, self.n, self.c, self.h, self.w, ld=self.ld)


# ------------------------------------------------------------------------------------------ packed-weight cache
class _Cache:
    """Repacked (bf16, K-major) weights per leaf module, invalidated when any source parameter is modified in
    place (tensor._version) or replaced (data_ptr)."""

    def __init__(self):
        self._d = weakref.WeakKeyDictionary()

    def get(self, mod, key, params, build):
        stamp = tuple((p.data_ptr(), p._version) for p in params if p is not None)
        slot = self._d.setdefault(mod, {})
        hit = slot.get(key)
        if hit is not None and hit[0] == stamp:
            return hit[1]
        with torch.no_grad():
            val = build()
        slot[key] = (stamp, val)
        return val


CACHE = _Cache()


def _f32(p):
    return None if p is None else p.detach()


def _require_leaf(mod, kinds, what):
    if type(mod) not in kinds:
        raise rt.B200Error(f"{what}: no CUDA path for leaf of type {type(mod).__module__}.{type(mod).__name__}")


def _act_name(m):
    if isinstance(m, nn.SiLU):
        return "silu"
    if isinstance(m, nn.ReLU):
        return "relu"
    if isinstance(m, nn.Identity):
        return None
    raise rt.B200Error(f"unsupported activation module {type(m).__name__}")


# ------------------------------------------------------------------------------------------ leaf operators
def _new_stats(n, oh, ow, c, device, phases=1):
    """Statistics buffer for an igemm writing an (n, oh, ow, c) bf16 tensor, or None when the staged epilogue that
    computes them does not apply (fewer than 33 channels / pitch not 16-byte aligned)."""
    if not FUSE_GN_STATS or c <= 32 or c % 8:
        return None
    p = rt.stats_slots(oh, ow) * phases
    return (torch.empty(n, p, c, 2, device=device, dtype=torch.float32), p, c)


def group_norm(a, mod, act=None):
    _require_leaf(mod, (nn.GroupNorm,), "GroupNorm")
    if not isinstance(a, Act):
        return _P().gn_split(a, mod, act=act)
    if training():
        segs = a.stats if (a.stats is not None and len(a.stats) <= 2) else [AG.tensor_stats(a.t, a.n, a.hw, a.c) + (a.c,)]
        if a.c % 8 or a.ld % 8 or a.c > 2048:
            raise rt.B200Error("GroupNorm backward needs channels % 8 == 0 and <= 2048")
        meta = (a.n, a.hw, a.c, mod.num_groups, mod.eps, act, [sg[1:] for sg in segs])
        out = AG.GroupNormFn.apply(a.t, mod.weight, mod.bias, segs[0][0], segs[1][0] if len(segs) > 1 else None, meta)
        return Act(out, a.n, a.h, a.w, a.c)
    out = Act.empty(a.n, a.h, a.w, a.c, a.t.device)
    if a.stats is not None and len(a.stats) <= 2 and a.c % 8 == 0 and a.ld % 8 == 0 and a.c <= 2048:
        rt.groupnorm_apply(a.t, a.n, a.hw, a.c, mod.num_groups, _f32(mod.weight), _f32(mod.bias), a.stats, mod.eps,
                           act=act, out=out.t, ld_x=a.ld, ld_y=a.c)
        return out
    rt.groupnorm(a.t, a.n, a.hw, a.c, mod.num_groups, _f32(mod.weight), _f32(mod.bias), mod.eps, act=act, out=out.t,
                 ld_x=a.ld, ld_y=a.c)
    return out


def conv2d(a, mod, rowbias=None, extra=None, act=None, out=None, c_off=0, out_dtype=torch.bfloat16, stats=True):
    """nn.Conv2d leaf on an Act. `extra` = (Act, 1x1 nn.Conv2d) fuses a second 1x1 convolution of another tensor
    into the same accumulator (ResNet shortcut)."""
    if not isinstance(a, Act):
        return _P().conv2d_any(a, mod, out_dtype)
    if hasattr(mod, "forward_act") and extra is None:
        y = mod.forward_act(a)
        if rowbias is not None:
            # a quantised leaf ends in its own output quantiser: the time-embedding row (blocks.py:118) is added afterwards,
            # as the reference does (a per-image broadcast add on a [n, hw, c] view; elementwise, autograd-visible)
            t = (y.t[:, :y.c].reshape(y.n, y.hw, y.c) + rowbias[:, None, :y.c].to(y.t.dtype)).reshape(y.n * y.hw, y.c)
            y = Act(t if t.shape[1] % 8 == 0 else torch.nn.functional.pad(t, (0, 8 - t.shape[1] % 8)), y.n, y.h, y.w, y.c)
        return y
    _require_leaf(mod, (nn.Conv2d,), "Conv2d")
    k, s, p = mod.kernel_size, mod.stride, mod.padding
    if mod.groups != 1 or mod.dilation != (1, 1):
        raise rt.B200Error("Conv2d: groups/dilation are not supported")
    cout = mod.out_channels
    if training() and out is None:
        if k in ((3, 3), (1, 1)) and s == (1, 1) and p == (k[0] // 2, k[0] // 2) and act is None:
            x2 = w2 = b2 = None
            cin2 = 0
            if extra is not None:
                xa, xm = extra
                _require_leaf(xm, (nn.Conv2d,), "Conv2d(shortcut)")
                x2, w2, b2, cin2 = xa.t, xm.weight, xm.bias, xa.c
            o, st = AG.ConvSameFn.apply(a.t, mod.weight, mod.bias, rowbias, x2, w2, b2, (a.n, a.h, a.w, a.c, cin2), stats,
                                        out_dtype == torch.float32)
            return Act(o, a.n, a.h, a.w, cout, _st(st, cout))
        if k == (4, 4) and s == (2, 2) and p == (1, 1) and extra is None and rowbias is None:
            o, st = AG.ConvDownFn.apply(a.t, mod.weight, mod.bias, (a.n, a.h, a.w, a.c), stats)
            return Act(o, a.n, a.h // 2, a.w // 2, cout, _st(st, cout))
        raise rt.B200Error(f"Conv2d: kernel {k} stride {s} padding {p} has no training path")
    if k in ((3, 3), (1, 1)) and s == (1, 1) and p == (k[0] // 2, k[0] // 2):
        if extra is None:
            wmat = CACHE.get(mod, "w", (mod.weight,), lambda: ops.pack_conv_weight(mod.weight))
            bias = _f32(mod.bias)
            ex = None
        else:
            xa, xm = extra
            _require_leaf(xm, (nn.Conv2d,), "Conv2d(shortcut)")
            if xm.kernel_size != (1, 1) or xm.stride != (1, 1):
                raise rt.B200Error("fused shortcut must be a 1x1 stride-1 convolution")
            wmat = CACHE.get(mod, "w+res", (mod.weight, xm.weight), lambda: torch.cat(
                [ops.pack_conv_weight(mod.weight), ops.pack_conv_weight(xm.weight)], dim=1).contiguous())
            bias = CACHE.get(mod, "b+res", (mod.bias, xm.bias), lambda: (
                (mod.bias if mod.bias is not None else 0) + (xm.bias if xm.bias is not None else 0)).float().contiguous())
            ex = (xa.t, xa.c, xa.ld)
        if out is None:
            # row pitch padded to 16 B so the next layer's TMA view of this tensor is legal (z_channels = 4)
            o = torch.empty(a.n * a.hw, (cout + 7) // 8 * 8, device=a.t.device, dtype=out_dtype)
        else:
            o = out.t
        st = _new_stats(a.n, a.h, a.w, cout, a.t.device) if (stats and out is None and o.dtype == torch.bfloat16) \
            else None
        ops.conv_same(a.t, a.n, a.h, a.w, a.c, wmat, cout, k[0], bias=bias, ld_x=a.ld, out=o, c_off=c_off,
                      rowbias=rowbias, act=act, extra=ex, stats=None if st is None else (st[0], st[1], 0, st[2]))
        return out if out is not None else Act(o, a.n, a.h, a.w, cout, None if st is None else [st + (cout,)])
    if k == (4, 4) and s == (2, 2) and p == (1, 1) and extra is None and rowbias is None:
        if a.h % 2 or a.w % 2:
            raise rt.B200Error("4x4 stride-2 convolution needs even spatial size")
        wmat = CACHE.get(mod, "w", (mod.weight,), lambda: ops.pack_conv_weight(mod.weight))
        o = Act.empty(a.n, a.h // 2, a.w // 2, cout, a.t.device) if out is None else out
        st = _new_stats(a.n, a.h // 2, a.w // 2, cout, a.t.device) if (stats and out is None) else None
        ops.conv_down4x4(a.t, a.n, a.h, a.w, a.c, wmat, cout, bias=_f32(mod.bias), ld_x=a.ld, out=o.t,
                         stats=None if st is None else (st[0], st[1], 0, st[2]))
        if st is not None:
            o.stats = [st + (cout,)]
        return o
    raise rt.B200Error(f"Conv2d: kernel {k} stride {s} padding {p} has no CUDA path")


def conv_transpose2d(a, mod, out=None, c_off=0):
    _require_leaf(mod, (nn.ConvTranspose2d,), "ConvTranspose2d")
    if mod.kernel_size != (4, 4) or mod.stride != (2, 2) or mod.padding != (1, 1) or mod.output_padding != (0, 0):
        raise rt.B200Error("ConvTranspose2d: only kernel 4 stride 2 padding 1 has a CUDA path")
    cout = mod.out_channels
    if training():
        if out is not None:
            raise rt.B200Error("conv_transpose2d: in-place concat target is an inference-only optimisation")
        o, st = AG.ConvUpFn.apply(a.t, mod.weight, mod.bias, (a.n, a.h, a.w, a.c), True)
        segs = _st(st, cout)
        act_o = Act(o, a.n, 2 * a.h, 2 * a.w, cout, segs)
        return act_o, (None if segs is None else segs[0])
    phases = CACHE.get(mod, "w", (mod.weight,), lambda: ops.pack_convT_weight(mod.weight))
    o = Act.empty(a.n, 2 * a.h, 2 * a.w, cout, a.t.device) if out is None else out
    st = _new_stats(a.n, a.h, a.w, cout, a.t.device, phases=4)
    ops.conv_up4x4(a.t, a.n, a.h, a.w, a.c, phases, cout, bias=_f32(mod.bias), ld_x=a.ld, out=o.t, c_off=c_off,
                   stats=st)
    if out is None and st is not None:
        o.stats = [st + (cout,)]
    return o, (None if st is None else st + (cout,))


def linear(x, mod, act=None, residual=None, out_dtype=torch.bfloat16, rows_per_img=None, rowbias=None, gate=None,
           stats=None):
    """nn.Linear leaf on a 2-D bf16 matrix."""
    if hasattr(mod, "forward_rows"):
        # quantised leaf (cim_layers drop-ins): its own CUDA path produces the layer output; what the float path fuses
        # into the GEMM epilogue is applied afterwards
        if rowbias is not None or stats is not None or out_dtype != torch.bfloat16:
            raise rt.B200Error("Linear: this epilogue combination is not available around a quantised leaf")
        y = mod.forward_rows(x, act=act)
        if gate is not None:
            if residual is None or rows_per_img is None:
                raise rt.B200Error("Linear: a gate needs its residual and rows_per_img")
            n = y.shape[0] // rows_per_img
            if training():
                return AG.GateResidualFn.apply(residual, y, gate, (n, rows_per_img, y.shape[1]))
            return rt.gate_residual(residual, y, gate, n, rows_per_img, y.shape[1])
        if residual is not None:
            if training():
                return residual[:, :y.shape[1]] + y
            out = torch.empty_like(y)
            rt.add_bf16(residual if residual.shape[1] == y.shape[1] else residual[:, :y.shape[1]].contiguous(), y, out)
            return out
        return y
    _require_leaf(mod, (nn.Linear, nn.modules.linear.NonDynamicallyQuantizableLinear), "Linear")
    if training() and act in (None, "relu") and gate is None and rowbias is None and out_dtype == torch.bfloat16:
        o, st = AG.LinearFn.apply(x, mod.weight, mod.bias, residual, rows_per_img, stats is not None, act)
        return (o, st) if stats is not None else o
    if training():
        raise rt.B200Error("Linear: this epilogue combination has no training path")
    wmat = CACHE.get(mod, "w", (mod.weight,), lambda: ops.pack_linear_weight(mod.weight))
    return rt.linear(x, wmat, bias=_f32(mod.bias), act=act, residual=residual, out_dtype=out_dtype,
                     rows_per_img=rows_per_img, rowbias=rowbias, gate=gate, stats=stats)


def _mha_params(mha):
    """(in_proj weight [3C, C], bias [3C] or None) as differentiable tensors (packed parameter of nn.MultiheadAttention, or
    the concatenation of CustomMultiheadAttention's q/k/v Linear parameters)."""
    if isinstance(mha, nn.MultiheadAttention):
        if mha.in_proj_weight is None:
            raise rt.B200Error("MultiheadAttention with kdim/vdim != embed_dim has no CUDA path")
        return mha.in_proj_weight, mha.in_proj_bias
    if all(hasattr(mha, k) for k in ("q_proj", "k_proj", "v_proj", "out_proj")):
        for k in ("q_proj", "k_proj", "v_proj"):
            _require_leaf(getattr(mha, k), (nn.Linear,), f"attention {k}")
        w = torch.cat([mha.q_proj.weight, mha.k_proj.weight, mha.v_proj.weight], 0)
        b = None if mha.q_proj.bias is None else torch.cat([mha.q_proj.bias, mha.k_proj.bias, mha.v_proj.bias], 0)
        return w, b
    raise rt.B200Error(f"attention module {type(mha).__name__} has no CUDA path")


def _mha_weights(mha):
    """(w_q, b_q, w_kv, b_kv, w_qkv, b_qkv, out_proj) for nn.MultiheadAttention (packed in_proj, blocks.py:83) or
    the reference's CustomMultiheadAttention (separate q/k/v Linear, models/multihead_attention.py:34-37)."""
    if isinstance(mha, nn.MultiheadAttention):
        if mha.in_proj_weight is None:
            raise rt.B200Error("MultiheadAttention with kdim/vdim != embed_dim has no CUDA path")
        w, b = mha.in_proj_weight, mha.in_proj_bias
        srcs = (w, b)
        get_w = lambda: w
        get_b = lambda: b
    elif all(hasattr(mha, k) for k in ("q_proj", "k_proj", "v_proj", "out_proj")):
        for k in ("q_proj", "k_proj", "v_proj"):
            _require_leaf(getattr(mha, k), (nn.Linear,), f"attention {k}")
        ws = [mha.q_proj.weight, mha.k_proj.weight, mha.v_proj.weight]
        bs = [mha.q_proj.bias, mha.k_proj.bias, mha.v_proj.bias]
        srcs = tuple(ws + bs)
        get_w = lambda: torch.cat(ws, 0)
        get_b = lambda: None if bs[0] is None else torch.cat(bs, 0)
    else:
        raise rt.B200Error(f"attention module {type(mha).__name__} has no CUDA path")

    def build():
        w, b = get_w(), get_b()
        C = w.shape[1]
        bf = None if b is None else b.float().contiguous()
        return {
            "qkv": ops.pack_linear_weight(w), "b_qkv": bf,
            "q": ops.pack_linear_weight(w[:C]), "b_q": None if bf is None else bf[:C].contiguous(),
            "kv": ops.pack_linear_weight(w[C:]), "b_kv": None if bf is None else bf[C:].contiguous(),
        }

    return CACHE.get(mha, "in_proj", srcs, build)


def _out_proj(o, mha, a):
    """out_proj GEMM with the residual add (and the next GroupNorm's statistics) in its epilogue."""
    C = a.c
    proj = mha.out_proj
    if hasattr(proj, "forward_rows"):
        return add(Act(linear(o, proj), a.n, a.h, a.w, C), a)
    if training():
        _require_leaf(proj, (nn.Linear, nn.modules.linear.NonDynamicallyQuantizableLinear), "Linear")
        out, stt = AG.LinearFn.apply(o, proj.weight, proj.bias, a.t, a.hw, True)
        return Act(out, a.n, a.h, a.w, C, _st(stt, C))
    st = _new_stats(a.n, 1, a.hw, C, a.t.device)
    out = linear(o, proj, residual=a.t, rows_per_img=a.hw, stats=None if st is None else (st[0], st[1], 0, st[2]))
    return Act(out, a.n, a.h, a.w, C, None if st is None else [st + (C,)])


def self_attention(a, norm, mha):
    """x + out_proj(MHA(GN(x))) — blocks.py:122-131. One QKV GEMM, flash attention, out-proj GEMM with the residual
    add in its epilogue."""
    if not isinstance(a, Act):
        return _P().self_attention(a, norm, mha)
    C, heads = a.c, mha.num_heads
    xn = group_norm(a, norm)
    if not _plain_projections(mha):
        # CustomMultiheadAttention whose q/k/v/out Linear leaves were swapped (the cim_qn_train surgery on the ReLU UNet,
        # models/multihead_attention.py:34-37): every projection runs through its own leaf
        q, k, v = linear(xn.t, mha.q_proj), linear(xn.t, mha.k_proj), linear(xn.t, mha.v_proj)
        if training():
            o = AG.AttentionFn.apply(q, torch.cat([k, v], dim=1), (a.n, heads, a.hw, a.hw, C // heads))
        else:
            o = rt.attention(q, k, v, a.n, heads, a.hw, a.hw, C // heads)
        return _out_proj(o, mha, a)
    if training():
        w, b = _mha_params(mha)
        qkv, _ = AG.LinearFn.apply(xn.t, w, b, None, None, False)
        o = AG.AttentionFn.apply(None, qkv, (a.n, heads, a.hw, a.hw, C // heads))
        return _out_proj(o, mha, a)
    pk = _mha_weights(mha)
    qkv = rt.linear(xn.t, pk["qkv"], bias=pk["b_qkv"])
    o = rt.attention(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], a.n, heads, a.hw, a.hw, C // heads)
    return _out_proj(o, mha, a)


class Context:
    """Text context of one forward: bf16 [n*lk, dim] rows plus, per cross-attention site, the projected K/V
    (context_proj -> in_proj[k,v]). The K/V of a site depend only on the context and that site's weights, not on
    x or t, so a Context that outlives one forward (the 1000-step sampling loop keeps `cond_input` fixed,
    tools/sample_ddpm_text_image_cond.py:64-82) computes them once instead of once per step."""
    __slots__ = ("t", "lk", "dim", "kv", "persistent", "sites")

    def __init__(self, t, lk, dim, persistent=False):
        self.t, self.lk, self.dim, self.kv, self.persistent = t, lk, dim, {}, persistent
        self.sites = {}     # id(mha) -> (context_proj module, mha): who to ask for a cached K/V again (rebind_cond)

    def __iter__(self):  # (ctx_t, lk, dim) unpacking used by the block classes
        return iter((self.t, self.lk, self.dim))


class CondState:
    """The (x, t)-independent part of a conditioned forward; built by Unet.prepare_cond."""
    __slots__ = ("inbuf", "cin", "ctx", "class_onehot", "raw")

    def __init__(self):
        self.inbuf = self.cin = self.ctx = self.class_onehot = self.raw = None


def rebind_cond(old, new):
    """Serve a NEW condition through the storage of an old CondState (same shapes), in place: the captured forward
    graphs that hold pointers into `old` (input buffer with the mask channels, text context rows, per-site K/V) stay
    valid, so a sampling job with fresh conditions costs the hoisted conditioning work once — not a re-capture of the
    whole forward (~65 ms per job, which a 20-step job cannot amortise). The cached K/V are recomputed from the new
    context with the sites' current weights. Returns False (nothing touched) when the two states do not match."""
    def same(a, b):
        return (a is None) == (b is None) and (a is None or (a.shape == b.shape and a.dtype == b.dtype
                                                              and a.device == b.device))
    if old.raw is not None or new.raw is not None or old.cin != new.cin:
        return False       # precise mode keeps the caller's own tensors: nothing to copy into
    if not same(old.inbuf, new.inbuf) or not same(old.class_onehot, new.class_onehot):
        return False
    if (old.ctx is None) != (new.ctx is None):
        return False
    if old.ctx is not None and (not same(old.ctx.t, new.ctx.t) or (old.ctx.lk, old.ctx.dim) != (new.ctx.lk, new.ctx.dim)
                                or set(old.ctx.kv) != set(old.ctx.sites)):
        return False
    with torch.no_grad():
        if old.inbuf is not None:
            old.inbuf.copy_(new.inbuf)
        if old.class_onehot is not None:
            old.class_onehot.copy_(new.class_onehot)
        if old.ctx is not None:
            old.ctx.t.copy_(new.ctx.t)
            for key, (proj, mha) in old.ctx.sites.items():
                pk = _mha_weights(mha)
                old.ctx.kv[key].copy_(rt.linear(linear(old.ctx.t, proj), pk["kv"], bias=pk["b_kv"]))
    return True


HOIST_CONDITIONING = True
BATCH_TEMB = True  # one GEMM for all per-block time-embedding projections  # tests flip this to compare hoisted and per-call conditioning


def cross_attention(a, norm, mha, ctx_proj_mod, ctx, lk, cache=None):
    """x + out_proj(MHA(q=GN(x), k=v=context_proj(ctx))) — blocks.py:133-145. ctx: bf16 [n*lk, ctx_dim].
    `cache`: a persistent Context whose per-site K/V are reused across forwards."""
    if not isinstance(a, Act):
        return _P().cross_attention(a, norm, mha, ctx_proj_mod, ctx, lk, cache=cache)
    C, heads = a.c, mha.num_heads
    xn = group_norm(a, norm)
    if not _plain_projections(mha) or type(ctx_proj_mod) is not nn.Linear:
        q = linear(xn.t, mha.q_proj)
        cp = linear(ctx, ctx_proj_mod)
        k, v = linear(cp, mha.k_proj), linear(cp, mha.v_proj)
        if training():
            o = AG.AttentionFn.apply(q, torch.cat([k, v], dim=1), (a.n, heads, a.hw, lk, C // heads))
        else:
            o = rt.attention(q, k, v, a.n, heads, a.hw, lk, C // heads)
        return _out_proj(o, mha, a)
    if training():
        w, b = _mha_params(mha)
        _require_leaf(ctx_proj_mod, (nn.Linear,), "Linear")
        cp, _ = AG.LinearFn.apply(ctx, ctx_proj_mod.weight, ctx_proj_mod.bias, None, None, False)
        q, kv = AG.CrossProjFn.apply(xn.t, cp, w, b, C)
        o = AG.AttentionFn.apply(q, kv, (a.n, heads, a.hw, lk, C // heads))
        return _out_proj(o, mha, a)
    pk = _mha_weights(mha)
    q = rt.linear(xn.t, pk["q"], bias=pk["b_q"])
    kv = cache.kv.get(id(mha)) if cache is not None else None
    if kv is None:
        cp = linear(ctx, ctx_proj_mod)
        kv = rt.linear(cp, pk["kv"], bias=pk["b_kv"])
        if cache is not None and cache.persistent:
            cache.kv[id(mha)] = kv
            cache.sites[id(mha)] = (ctx_proj_mod, mha)
    o = rt.attention(q, kv[:, :C], kv[:, C:], a.n, heads, a.hw, lk, C // heads)
    return _out_proj(o, mha, a)


class TembBank:
    """All per-block time-embedding projections `Linear(t_emb_dim -> C_out)` of one forward (22 in the full UNet,
    blocks.py:57-62) depend only on t, so they run as ONE GEMM against the row-concatenated weights at the top of the
    forward; each ResNet unit then picks its [n, C_out] slice as the conv epilogue's row bias."""

    def __init__(self, act, slices):
        self.act, self.slices = act, slices

    @staticmethod
    def build(owner, temb_act, linears):
        mods = [m for m in linears if type(m) is nn.Linear and m.out_features % 8 == 0]
        if len(mods) < 2:
            return TembBank(temb_act, {})
        if training():
            w = torch.cat([m.weight for m in mods], 0)
            b = torch.cat([m.bias if m.bias is not None else torch.zeros(m.out_features, device=m.weight.device)
                           for m in mods], 0)
            big, _ = AG.LinearFn.apply(temb_act, w, b, None, None, False)
            slices, off = {}, 0
            for m in mods:
                slices[id(m)] = big[:, off:off + m.out_features]
                off += m.out_features
            return TembBank(temb_act, slices)
        params = tuple(m.weight for m in mods) + tuple(m.bias for m in mods)

        def pack():
            w = ops.pack_linear_weight(torch.cat([m.weight for m in mods], 0))
            b = torch.cat([m.bias if m.bias is not None else torch.zeros(m.out_features, device=m.weight.device)
                           for m in mods], 0).float().contiguous()
            return w, b

        w, b = CACHE.get(owner, "temb_bank", params, pack)
        big = rt.linear(temb_act, w, bias=b)
        slices, off = {}, 0
        for m in mods:
            slices[id(m)] = big[:, off:off + m.out_features]
            off += m.out_features
        return TembBank(temb_act, slices)


def resnet(a, first, temb_layer, second, shortcut, temb_act):
    """GN-act-conv3x3 (+ time-embedding row bias) -> GN-act-conv3x3 + 1x1 shortcut of the block input, the shortcut
    fused into the second conv's accumulator — blocks.py:113-119. `first`/`second` are the reference's
    nn.Sequential(GroupNorm, activation, Conv2d); temb_layer is nn.Sequential(activation, Linear) or None."""
    if not isinstance(a, Act):
        return _P().resnet(a, first, temb_layer, second, shortcut, temb_act)
    act = _act_name(first[1])
    h = group_norm(a, first[0], act=act)
    rowbias = None
    if temb_layer is not None:
        if isinstance(temb_act, TembBank):
            rowbias = temb_act.slices.get(id(temb_layer[1]))
            if rowbias is None:
                rowbias = linear(temb_act.act, temb_layer[1])
        else:
            rowbias = linear(temb_act, temb_layer[1])
    h = conv2d(h, first[2], rowbias=rowbias)
    h = group_norm(h, second[0], act=_act_name(second[1]))
    if type(second[2]) is nn.Conv2d and type(shortcut) is nn.Conv2d:
        return conv2d(h, second[2], extra=(a, shortcut))
    # quantised leaves: run the two convolutions separately and add
    h = conv2d(h, second[2])
    s = conv2d(a, shortcut)
    return add(h, s)


def add(a, b):
    if training():
        return Act(a.t[:, :a.c] + b.t[:, :b.c], a.n, a.h, a.w, a.c)
    out = Act.empty(a.n, a.h, a.w, a.c, a.t.device)
    rt.add_bf16(a.t, b.t, out.t)
    return out


def activate_temb(t_emb, act_cls):
    """act(t_emb) as bf16 [B, D]: the shared first half of every block's t_emb_layers (blocks.py:57-62)."""
    if training() and t_emb.requires_grad:
        fn = {"silu": torch.nn.functional.silu, "relu": torch.relu, None: lambda v: v}[_act_name(act_cls())]
        return fn(t_emb.to(torch.bfloat16))  # (B, t_emb_dim): tiny
    src = t_emb if t_emb.dtype == torch.bfloat16 else rt.cast_bf16(t_emb.float())
    out = torch.empty_like(src)
    rt.copy_act(src, out, 0, act=_act_name(act_cls()))
    return out


def prepare_context(context):
    """(B, L, D) fp32 text context -> (bf16 [B*L, D], L, D): the autocast cast, done once per forward instead of
    once per cross-attention site."""
    b, l, d = context.shape
    if d % 8:
        raise rt.B200Error("context dimension must be a multiple of 8")
    t = context if context.dtype == torch.bfloat16 else rt.cast_bf16(context.float())
    return Context(t.reshape(b * l, d), l, d)


def upsample_concat(a, up_conv, skip):
    """x = up_sample_conv(x); x = cat([x, out_down], dim=1) — blocks.py:461-464. The transposed convolution writes
    straight into the left channel slice of the concat buffer; the skip tensor is copied into the right slice."""
    if not isinstance(a, Act):
        return _P().upsample_concat(a, up_conv, skip)
    if up_conv is None and skip is None:
        return a
    h, w = (2 * a.h, 2 * a.w) if up_conv is not None else (a.h, a.w)
    c_up = up_conv.out_channels if up_conv is not None else a.c
    c_tot = c_up + (skip.c if skip is not None else 0)
    if skip is not None and (skip.h, skip.w, skip.n) != (h, w, a.n):
        raise rt.B200Error("skip tensor shape does not match the upsampled tensor")
    if training():
        up = conv_transpose2d(a, up_conv)[0] if up_conv is not None else a
        if skip is None:
            return up
        out = Act(torch.cat([up.t[:, :up.c], skip.t[:, :skip.c]], dim=1), a.n, h, w, c_tot)
        segs = [x.stats[0] if x.stats is not None and len(x.stats) == 1 else None for x in (up, skip)]
        if all(sg is not None for sg in segs):
            out.stats = segs
        return out
    out = Act.empty(a.n, h, w, c_tot, a.t.device)
    if up_conv is not None:
        _, st_up = conv_transpose2d(a, up_conv, out=out, c_off=0)
    else:
        rt.copy_act(a.t, out.t, 0)
        st_up = a.stats[0] if a.stats is not None and len(a.stats) == 1 else None
    segs = [st_up]
    if skip is not None:
        rt.copy_act(skip.t[:, :skip.c] if skip.ld != skip.c else skip.t, out.t, c_up)
        segs.append(skip.stats[0] if skip.stats is not None and len(skip.stats) == 1 else None)
    if all(sg is not None for sg in segs):
        out.stats = segs  # channel-concat of tensors whose producers each left their own partial statistics
    return out


# ------------------------------------------------------------------------------------------ DiT token-row operators
# The DiT residual stream is a bf16 [n*L, hidden] row matrix (L tokens per image); per-image adaLN operands are
# [n, hidden] column slices of ONE batched projection of relu(t_emb) (TembBank over every layer's
# adaptive_norm_layer[1], transformer_layer.py:62-65 + transformer.py:126-129).
def layer_norm_mod(x, norm, n, L, shift=None, scale=None):
    """LayerNorm(elementwise_affine=False) * (1 + scale) + shift — transformer_layer.py:88-89,94,103-104."""
    _require_leaf(norm, (nn.LayerNorm,), "LayerNorm")
    if norm.elementwise_affine or len(norm.normalized_shape) != 1:
        raise rt.B200Error("LayerNorm: only the affine-free last-dimension form has a CUDA path")
    c = norm.normalized_shape[0]
    if training() and (x.requires_grad or (scale is not None and scale.requires_grad)):
        return AG.LayerNormModFn.apply(x, shift, scale, (n, L, c, norm.eps))
    return rt.layernorm_mod(x, n, L, c, shift=shift, scale=scale, eps=norm.eps)


def gated_linear(h_in, mod, x, gate, n, L):
    """x + gate[img] * Linear(h_in) — the adaLN-Zero residual (transformer_layer.py:90,105): one GEMM with the gate
    and the residual in its epilogue; in training the un-gated product is kept for the gate gradient."""
    if training():
        h = linear(h_in, mod)
        return AG.GateResidualFn.apply(x, h, gate, (n, L, x.shape[1]))
    return linear(h_in, mod, residual=x, gate=gate, rows_per_img=L)


def dit_self_attention(xn, attn, n, L):
    """models/attention.py:33-78 up to (not including) output_proj: packed QKV GEMM + flash attention."""
    A, heads, d = attn.att_dim, attn.n_heads, attn.head_dim
    qkv = linear(xn, attn.qkv_proj)
    if training():
        return AG.AttentionFn.apply(None, qkv, (n, heads, L, L, d))
    return rt.attention(qkv[:, :A], qkv[:, A:2 * A], qkv[:, 2 * A:], n, heads, L, L, d)


def _plain_projections(mha):
    return all(type(getattr(mha, k, None)) is nn.Linear for k in ("q_proj", "k_proj", "v_proj", "out_proj")) \
        or isinstance(mha, nn.MultiheadAttention)


def _cross_attention_leafwise(xn, x, mha, ctx_proj_mod, ctx, n, L):
    """Cross-attention whose projections are separate (possibly quantised) leaves — CustomMultiheadAttention exists for
    exactly this (models/multihead_attention.py:34-37): each projection runs through its own leaf."""
    C, heads = mha.embed_dim, mha.num_heads
    q = linear(xn, mha.q_proj)
    cp = linear(ctx.t, ctx_proj_mod)
    kv = torch.cat([linear(cp, mha.k_proj), linear(cp, mha.v_proj)], dim=1)
    if training():
        o = AG.AttentionFn.apply(q, kv, (n, heads, L, ctx.lk, C // heads))
    else:
        o = rt.attention(q, kv[:, :C], kv[:, C:], n, heads, L, ctx.lk, C // heads)
    return linear(o, mha.out_proj, residual=x)


def dit_cross_attention(x, norm, mha, ctx_proj_mod, ctx, n, L):
    """x + out_proj(MHA(q = LN(x), k = v = context_proj(context))) — transformer_layer.py:92-101 (un-gated). The K/V of
    a site depend only on the context and the weights: a persistent Context keeps them across sampling steps."""
    C, heads = mha.embed_dim, mha.num_heads
    xn = layer_norm_mod(x, norm, n, L)
    if not _plain_projections(mha) or type(ctx_proj_mod) is not nn.Linear:
        return _cross_attention_leafwise(xn, x, mha, ctx_proj_mod, ctx, n, L)
    if training():
        w, b = _mha_params(mha)
        q, _ = AG.LinearFn.apply(xn, w[:C], None if b is None else b[:C], None, None, False)
        _require_leaf(ctx_proj_mod, (nn.Linear,), "Linear")
        cp, _ = AG.LinearFn.apply(ctx.t, ctx_proj_mod.weight, ctx_proj_mod.bias, None, None, False)
        kv, _ = AG.LinearFn.apply(cp, w[C:], None if b is None else b[C:], None, None, False)
        o = AG.AttentionFn.apply(q, kv, (n, heads, L, ctx.lk, C // heads))
        out, _ = AG.LinearFn.apply(o, mha.out_proj.weight, mha.out_proj.bias, x, None, False)
        return out
    pk = _mha_weights(mha)
    q = rt.linear(xn, pk["q"], bias=pk["b_q"])
    kv = ctx.kv.get(id(mha))
    if kv is None:
        kv = rt.linear(linear(ctx.t, ctx_proj_mod), pk["kv"], bias=pk["b_kv"])
        if ctx.persistent:
            ctx.kv[id(mha)] = kv
            ctx.sites[id(mha)] = (ctx_proj_mod, mha)
    o = rt.attention(q, kv[:, :C], kv[:, C:], n, heads, L, ctx.lk, C // heads)
    return linear(o, mha.out_proj, residual=x)


def patch_embed(xbuf, n, H, W, cin, p, mod, pos):
    """PatchEmbedding.patch_embed[0] on the NHWC input buffer (+ position embedding) — patch_embed.py:84-95."""
    if hasattr(mod, "forward_rows"):
        # quantised leaf: its input quantiser needs the folded patches as rows (layout glue on an 8-channel tensor)
        gh, gw, ld = H // p, W // p, xbuf.stride(0)
        tok = xbuf.reshape(n, gh, p, gw, p, ld)[..., :cin].permute(0, 1, 3, 2, 4, 5).reshape(n * gh * gw, p * p * cin)
        kp = (p * p * cin + 7) // 8 * 8
        tok = torch.nn.functional.pad(tok, (0, kp - p * p * cin))
        y = mod.forward_rows(tok)
        pos_all = pos.repeat(n, 1)
        if training():
            return y + pos_all
        out = torch.empty_like(y)
        rt.add_bf16(y, pos_all, out)
        return out
    _require_leaf(mod, (nn.Linear,), "Linear(patch_embed)")
    if p * p > 4:
        raise rt.B200Error("patch embedding: patch sizes above 2 are not supported")
    if mod.in_features != p * p * cin:
        raise rt.B200Error("patch embedding: Linear.in_features does not match patch_size^2 * channels")
    if training():
        return AG.PatchEmbedFn.apply(xbuf, mod.weight, mod.bias, pos, (n, H, W, cin, p))
    wmat = CACHE.get(mod, "patch", (mod.weight,), lambda: ops.pack_patch_weight(mod.weight, p, cin))
    return ops.patch_embed(xbuf, n, H, W, cin, p, wmat, mod.out_features, bias=_f32(mod.bias), pos=pos)


def unpatchify(tokens, n, ch, gh, gw, p):
    if training() and tokens.requires_grad:
        return AG.UnpatchifyFn.apply(tokens, (n, ch, gh, gw, p))
    return rt.unpatchify(tokens, n, ch, gh, gw, p, p)

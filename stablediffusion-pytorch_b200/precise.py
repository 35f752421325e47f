"""Precise mode of the fused engine: the configuration that brings the whole-network noise prediction within north_star's
tolerance of the reference's fp32 output (profiles/r2_error_budget.md names where the fast mode's error comes from).

    import sdb200.engine as E;  E.set_parity_mode("precise")        # or B200SD_PARITY_MODE=precise

What changes against the fast (all-bf16) mode, family by family of the error budget:
  S  stored values   the residual stream and every GEMM result stay fp32 in HBM (`PAct`)
  W  weights         each weight matrix is SPLIT into hi = bf16(w) and lo = bf16(w - hi)
  A  activations     each tensor-core A operand is split the same way by its producer (csrc/precise.cu:
                     GroupNorm+activation of the fp32 stream, or a plain split of an fp32 tensor)
and a product is A_hi.[W_hi | W_lo] + A_lo.W_hi on the SAME tcgen05 kind::f16 implicit GEMM (fp32 accumulation in TMEM;
the second launch accumulates in place through the fp32 residual operand) — 3x the MMA work instead of a second
arithmetic type. Attention keeps the bf16 flash kernel (q, k, v, P, O in bf16: 2.2e-3 of the budget).
Inference only. The fast mode stays the default and the one bench.py measures; its measured error is in the bench line.
"""
import torch
import torch.nn as nn

from . import engine as E
from . import ops
from . import runtime as rt


class PAct:
    """fp32 NHWC activation [n*h*w, ld] (the precise-mode counterpart of engine.Act)."""
    __slots__ = ("f", "n", "h", "w", "c")

    def __init__(self, f, n, h, w, c):
        self.f, self.n, self.h, self.w, self.c = f, n, h, w, c

    @property
    def hw(self):
        return self.h * self.w

    @staticmethod
    def from_nchw(x):
        n, c, h, w = x.shape
        return PAct(x.float().permute(0, 2, 3, 1).reshape(n * h * w, c).contiguous(), n, h, w, c)

    def to_nchw(self):
        return rt.nhwc_to_nchw(self.f, self.n, self.c, self.h, self.w, ld=self.f.stride(0))


class Split:
    """A tensor-core A operand as bf16 hi (columns [0, c)) and, when `lo`, bf16 lo (columns [cp, cp + c))."""
    __slots__ = ("t", "c", "cp", "lo", "geom")

    def __init__(self, t, c, lo=True, geom=None):
        self.t, self.c, self.cp, self.lo, self.geom = t, c, rt.pad8(c), lo, geom

    def views(self):
        return [self.t] + ([self.t[:, self.cp:]] if self.lo else [])


def split(f32_2d, c=None, act=None):
    c = f32_2d.shape[1] if c is None else c
    return Split(rt.split_f32(f32_2d, c, act=act), c)


def gn_split(a, mod, act=None):
    E._require_leaf(mod, (nn.GroupNorm,), "GroupNorm")
    t = rt.groupnorm_split(a.f, a.n, a.hw, a.c, mod.num_groups, E._f32(mod.weight), E._f32(mod.bias), mod.eps, act=act)
    return Split(t, a.c, geom=(a.n, a.h, a.w))


def conv2d_any(a, mod, out_dtype=None):
    """engine.conv2d on a precise-mode tensor: a PAct (split here) or the Split a GroupNorm just produced."""
    if isinstance(a, PAct):
        return conv2d(split(a.f, a.c), (a.n, a.h, a.w), mod)
    if isinstance(a, Split) and a.geom is not None:
        return conv2d(a, a.geom, mod)
    raise rt.B200Error(f"conv2d: unexpected precise-mode operand {type(a).__name__}")


def _wsplit(mod, key, params, packers):
    """((hi, lo) packed bf16 matrices) of a leaf's weight(s): hi = pack(bf16(w)), lo = pack(bf16(w - hi))."""
    def build():
        out = []
        for w, pack in packers():
            w = w.detach().float()
            hi = w.to(torch.bfloat16)
            out.append((pack(hi), pack((w - hi.float()).to(torch.bfloat16))))
        return out
    return E.CACHE.get(mod, "precise:" + key, params, build)


def _cat(mod, key, params, mats):
    return E.CACHE.get(mod, "precise:cat:" + key, params, lambda: torch.cat(mats, dim=1).contiguous())


def _gemm(parts, wkey_mod, wkey, wparams, n_out, on, oh, ow, out, strides, bias=None, rowbias=None, residual=None,
          res_strides=None):
    """out (fp32) = sum over parts of A.W (+ bias + rowbias + residual), every product as the split-operand sum.
    parts: list of (Split, build_srcs(view) -> [Src], segs with part-local source indices, (w_hi, w_lo))."""
    assert out.dtype == torch.float32
    # launch 1: every part's hi view against [W_hi | W_lo]
    srcs, seg_hi, base = [], [], 0
    for sp, build, segs, _ in parts:
        ss = build(sp.t)
        seg_hi += [(s + base, dy, dx, c0, nch) for s, dy, dx, c0, nch in segs]
        srcs += ss
        base += len(ss)
    w1 = _cat(wkey_mod, wkey + ":1", wparams, [w[0] for *_, w in parts] + [w[1] for *_, w in parts])
    rt.igemm(srcs, seg_hi + seg_hi, w1, n_out, on, oh, ow, out, strides, bias=bias, rowbias=rowbias, residual=residual,
             res_strides=res_strides)
    # launch 2: the lo views against W_hi, accumulated in place
    lo_parts = [p for p in parts if p[0].lo]
    if lo_parts:
        srcs, seg_lo, base = [], [], 0
        for sp, build, segs, _ in lo_parts:
            ss = build(sp.t[:, sp.cp:])
            seg_lo += [(s + base, dy, dx, c0, nch) for s, dy, dx, c0, nch in segs]
            srcs += ss
            base += len(ss)
        tag = wkey + ":2:" + "".join("1" if p[0].lo else "0" for p in parts)
        w2 = _cat(wkey_mod, tag, wparams, [w[0] for *_, w in lo_parts])
        rt.igemm(srcs, seg_lo, w2, n_out, on, oh, ow, out, strides, residual=out, res_strides=strides)
    return out


def _pad64(c):
    return (c + 63) // 64 * 64


def linear(a, mod, residual=None):
    """nn.Linear on a Split [rows, K] -> fp32 [rows, N]."""
    E._require_leaf(mod, (nn.Linear, nn.modules.linear.NonDynamicallyQuantizableLinear), "Linear")
    (w,) = _wsplit(mod, "w", (mod.weight,), lambda: [(mod.weight, ops.pack_linear_weight)])
    return _linear_w(a, mod, "lin", (mod.weight,), w, mod.out_features, E._f32(mod.bias), residual)


def _linear_w(a, key_mod, key, params, w, n_out, bias, residual=None):
    rows = a.t.shape[0]
    out = torch.empty(rows, n_out, device=a.t.device, dtype=torch.float32)
    on, ow = 1, rows
    build = lambda v: [rt.make_src(v, on, 1, ow, a.c, ld=v.stride(0))]   # noqa: E731
    strides = (ow * n_out, 0, n_out)
    rs = None if residual is None else (ow * residual.stride(0), 0, residual.stride(0))
    return _gemm([(a, build, [(0, 0, 0, 0, _pad64(a.c) // 64)], w)], key_mod, key, params, n_out, on, 1, ow, out, strides,
                 bias=bias, residual=residual, res_strides=rs)


def conv2d(a_split, geom, mod, rowbias=None, extra=None, residual=None):
    """nn.Conv2d leaf on a Split of an (n, h, w, cin) tensor -> PAct (fp32). extra = (Split of the block input, 1x1 conv)."""
    E._require_leaf(mod, (nn.Conv2d,), "Conv2d")
    n, h, w = geom
    k, s, p = mod.kernel_size, mod.stride, mod.padding
    if mod.groups != 1 or mod.dilation != (1, 1):
        raise rt.B200Error("Conv2d: groups/dilation are not supported")
    cin, cout = mod.in_channels, mod.out_channels
    dev = a_split.t.device
    if k in ((3, 3), (1, 1)) and s == (1, 1) and p == (k[0] // 2, k[0] // 2):
        pad = k[0] // 2
        (wm,) = _wsplit(mod, "w", (mod.weight,), lambda: [(mod.weight, ops.pack_conv_weight)])
        segs = [(0, ky - pad, kx - pad, 0, _pad64(cin) // 64) for ky in range(k[0]) for kx in range(k[0])]
        parts = [(a_split, lambda v: [rt.make_src(v, n, h, w, cin, ld=v.stride(0))], segs, wm)]
        params, bias, key_mod = (mod.weight,), E._f32(mod.bias), mod
        if extra is not None:
            xs, xm = extra
            E._require_leaf(xm, (nn.Conv2d,), "Conv2d(shortcut)")
            if xm.kernel_size != (1, 1) or xm.stride != (1, 1):
                raise rt.B200Error("fused shortcut must be a 1x1 stride-1 convolution")
            (wx,) = _wsplit(xm, "w", (xm.weight,), lambda: [(xm.weight, ops.pack_conv_weight)])
            parts.append((xs, lambda v: [rt.make_src(v, n, h, w, xs.c, ld=v.stride(0))],
                          [(0, 0, 0, 0, _pad64(xs.c) // 64)], wx))
            params = (mod.weight, xm.weight)
            bias = E.CACHE.get(mod, "b+res", (mod.bias, xm.bias), lambda: (
                (mod.bias if mod.bias is not None else 0) + (xm.bias if xm.bias is not None else 0)).float().contiguous())
        ld = max(cout, 4)
        out = torch.empty(n * h * w, ld, device=dev, dtype=torch.float32) if ld == cout else \
            torch.zeros(n * h * w, ld, device=dev, dtype=torch.float32)
        _gemm(parts, key_mod, "conv", params, cout, n, h, w, out, (h * w * ld, w * ld, ld), bias=bias,
              rowbias=rowbias, residual=residual, res_strides=None if residual is None else
              (h * w * residual.stride(0), w * residual.stride(0), residual.stride(0)))
        return PAct(out, n, h, w, cout)
    if k == (4, 4) and s == (2, 2) and p == (1, 1) and extra is None and rowbias is None:
        if h % 2 or w % 2:
            raise rt.B200Error("4x4 stride-2 convolution needs even spatial size")
        oh, ow = h // 2, w // 2
        (wm,) = _wsplit(mod, "w", (mod.weight,), lambda: [(mod.weight, ops.pack_conv_weight)])
        par = {0: (1, -1), 1: (0, 0), 2: (1, 0), 3: (0, 1)}
        segs = []
        for ky in range(4):
            for kx in range(4):
                (py, dy), (px, dx) = par[ky], par[kx]
                segs.append((py * 2 + px, dy, dx, 0, _pad64(cin) // 64))
        build = lambda v: [rt.make_src(v, n, oh, ow, cin, ld=v.stride(0), sy=2, sx=2, oy=py, ox=px, hh=h, ww=w)  # noqa: E731
                           for py in (0, 1) for px in (0, 1)]
        out = torch.empty(n * oh * ow, cout, device=dev, dtype=torch.float32)
        _gemm([(a_split, build, segs, wm)], mod, "down", (mod.weight,), cout, n, oh, ow, out, (oh * ow * cout, ow * cout, cout),
              bias=E._f32(mod.bias))
        return PAct(out, n, oh, ow, cout)
    raise rt.B200Error(f"Conv2d: kernel {k} stride {s} padding {p} has no CUDA path")


def conv_transpose2d(a, mod):
    """nn.ConvTranspose2d(4, 2, 1) of a PAct as four sub-pixel phase GEMMs interleaved into an fp32 (2h, 2w) tensor."""
    E._require_leaf(mod, (nn.ConvTranspose2d,), "ConvTranspose2d")
    if mod.kernel_size != (4, 4) or mod.stride != (2, 2) or mod.padding != (1, 1) or mod.output_padding != (0, 0):
        raise rt.B200Error("ConvTranspose2d: only kernel 4 stride 2 padding 1 has a CUDA path")
    n, h, w, cin, cout = a.n, a.h, a.w, a.c, mod.out_channels
    sp = split(a.f, a.c)

    def build_phases():
        wf = mod.weight.detach().float()
        hi = wf.to(torch.bfloat16)
        ph_hi, ph_lo = ops.pack_convT_weight(hi), ops.pack_convT_weight((wf - hi.float()).to(torch.bfloat16))
        return [((mh, ml), segs) for (mh, segs), (ml, _) in zip(ph_hi, ph_lo)]

    phases = E.CACHE.get(mod, "precise:phases", (mod.weight,), build_phases)
    out = torch.empty(n * 4 * h * w, cout, device=a.f.device, dtype=torch.float32)
    flat = out.view(-1)
    build = lambda v: [rt.make_src(v, n, h, w, cin, ld=v.stride(0))]   # noqa: E731
    for idx, (wm, segs) in enumerate(phases):
        py, px = idx // 2, idx % 2
        o = flat[(py * 2 * w + px) * cout:]
        _gemm([(sp, build, list(segs), wm)], mod, f"phase{idx}", (mod.weight,), cout, n, h, w, o,
              (4 * h * w * cout, 4 * w * cout, 2 * cout), bias=E._f32(mod.bias))
    return PAct(out, n, 2 * h, 2 * w, cout)


# ------------------------------------------------------------------------------------------ block-level operators
class PTemb:
    """act(t_emb) as a split operand, shared by every block's time-embedding Linear (blocks.py:57-62)."""
    __slots__ = ("sp",)

    def __init__(self, sp):
        self.sp = sp


def resnet(a, first, temb_layer, second, shortcut, temb):
    act = E._act_name(first[1])
    rowbias = None
    if temb_layer is not None:
        rowbias = linear(temb.sp, temb_layer[1])          # fp32 [n, C_out], added per image in the conv epilogue
    h = conv2d(gn_split(a, first[0], act=act), (a.n, a.h, a.w), first[2], rowbias=rowbias)
    h2 = gn_split(h, second[0], act=E._act_name(second[1]))
    return conv2d(h2, (a.n, a.h, a.w), second[2], extra=(split(a.f, a.c), shortcut))


def _attend(q32, kv32_or_qkv, a, heads, lk, C, fused_qkv):
    d = C // heads
    if fused_qkv:
        qkv = rt.cast_bf16(kv32_or_qkv)
        return rt.attention(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], a.n, heads, a.hw, a.hw, d)
    q, kv = rt.cast_bf16(q32), kv32_or_qkv
    return rt.attention(q, kv[:, :C], kv[:, C:], a.n, heads, a.hw, lk, d)


def _in_proj(mha):
    """Split packed (qkv, q, kv) projection weights + fp32 biases of nn.MultiheadAttention / CustomMultiheadAttention."""
    w, b = E._mha_params(mha)
    C = w.shape[1]
    params = tuple(p for p in mha.parameters())

    def build():
        wf = w.detach().float()
        hi = wf.to(torch.bfloat16)
        lo = (wf - hi.float()).to(torch.bfloat16)
        pk = ops.pack_linear_weight
        bf = None if b is None else b.detach().float().contiguous()
        return {"qkv": (pk(hi), pk(lo)), "q": (pk(hi[:C]), pk(lo[:C])), "kv": (pk(hi[C:]), pk(lo[C:])),
                "b_qkv": bf, "b_q": None if bf is None else bf[:C].contiguous(),
                "b_kv": None if bf is None else bf[C:].contiguous()}

    return E.CACHE.get(mha, "precise:in_proj", params, build), params


def _out_proj(o_bf16, mha, a):
    sp = Split(o_bf16, a.c, lo=False)        # the attention kernel's output is a single bf16 operand
    return PAct(linear(sp, mha.out_proj, residual=a.f), a.n, a.h, a.w, a.c)


def self_attention(a, norm, mha):
    C, heads = a.c, mha.num_heads
    xn = gn_split(a, norm)
    pk, params = _in_proj(mha)
    qkv = _linear_w(xn, mha, "qkv", params, pk["qkv"], 3 * C, pk["b_qkv"])
    return _out_proj(_attend(None, qkv, a, heads, a.hw, C, True), mha, a)


def cross_attention(a, norm, mha, ctx_proj_mod, ctx, lk, cache=None):
    C, heads = a.c, mha.num_heads
    xn = gn_split(a, norm)
    pk, params = _in_proj(mha)
    q = _linear_w(xn, mha, "q", params, pk["q"], C, pk["b_q"])
    kv = cache.kv.get(id(mha)) if cache is not None else None
    if kv is None:
        cp = linear(ctx, ctx_proj_mod)                      # ctx: Split of the fp32 text context
        kv = rt.cast_bf16(_linear_w(split(cp), mha, "kv", params, pk["kv"], 2 * C, pk["b_kv"]))
        if cache is not None and cache.persistent:
            cache.kv[id(mha)] = kv
    return _out_proj(_attend(q, kv, a, heads, lk, C, False), mha, a)


def upsample_concat(a, up_conv, skip):
    if up_conv is None and skip is None:
        return a
    up = conv_transpose2d(a, up_conv) if up_conv is not None else a
    if skip is None:
        return up
    if (skip.h, skip.w, skip.n) != (up.h, up.w, up.n):
        raise rt.B200Error("skip tensor shape does not match the upsampled tensor")
    return PAct(torch.cat([up.f[:, :up.c], skip.f[:, :skip.c]], dim=1), up.n, up.h, up.w, up.c + skip.c)


def down_conv(a, mod):
    return conv2d(split(a.f, a.c), (a.n, a.h, a.w), mod)


class PContext:
    """Text context in precise mode: the fp32 rows as a split operand (+ per-site K/V cache, as engine.Context)."""
    __slots__ = ("t", "lk", "dim", "kv", "persistent")

    def __init__(self, sp, lk, dim):
        self.t, self.lk, self.dim, self.kv, self.persistent = sp, lk, dim, {}, False

    def __iter__(self):
        return iter((self.t, self.lk, self.dim))


def prepare_context(context):
    b, l, d = context.shape
    return PContext(split(context.float().reshape(b * l, d).contiguous()), l, d)


def unet_run(m, x, tt, st):
    """models/_unet.py UnetCore.run in precise mode: the same block walk on PAct tensors."""
    n, c, h, w = x.shape
    xin = PAct.from_nchw(x)
    if m.image_cond:
        # mask branch (unet_cond_base.py:134-138): nearest resize = a strided read of the one-hot mask, then the 1x1
        # conv 18 -> 3 as a split-operand GEMM, concatenated behind the latent channels
        mask = st.raw["image"].float()
        sy, sx = mask.shape[2] // h, mask.shape[3] // w
        if mask.shape[2] % h or mask.shape[3] % w:
            raise rt.B200Error("precise mode: mask size must be a multiple of the latent size")
        mk = mask[:, :, ::sy, ::sx].permute(0, 2, 3, 1).reshape(n * h * w, -1).contiguous()
        mc = conv2d(split(mk), (n, h, w), m.cond_conv_in)
        xin = PAct(torch.cat([xin.f, mc.f[:, :mc.c]], dim=1).contiguous(), n, h, w, c + mc.c)
        out = conv2d(split(xin.f, xin.c), (n, h, w), m.conv_in_concat)
    else:
        out = conv2d(split(xin.f, xin.c), (n, h, w), m.conv_in)

    emb = rt.time_embedding(tt, n, m.t_emb_dim, dtype=torch.float32)
    emb = linear(split(emb), m.t_proj[0])
    emb = linear(split(emb, act=E._act_name(m.t_proj[1])), m.t_proj[2])
    if m.class_cond:
        emb = emb + st.class_onehot @ m.class_emb.weight.detach()
    temb = PTemb(split(emb, act=E._act_name(m.ACT())))

    ctx = prepare_context(st.raw["text"]) if m.text_cond else None
    skips = []
    for down in m.downs:
        skips.append(out)
        out = down.run(out, temb, ctx) if m.CONDITIONED else down.run(out, temb)
    for mid in m.mids:
        out = mid.run(out, temb, ctx) if m.CONDITIONED else mid.run(out, temb)
    for up in m.ups:
        skip = skips.pop()
        out = up.run(out, skip, temb, ctx) if m.CONDITIONED else up.run(out, skip, temb)
    return conv2d(gn_split(out, m.norm_out, act=E._act_name(m.ACT())), (n, h, w), m.conv_out)

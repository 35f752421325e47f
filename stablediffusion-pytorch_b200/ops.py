"""Convolution / linear operators expressed on the implicit-GEMM core (csrc/igemm.cu).

Activations are NHWC bf16 2-D tensors [N*H*W, ld] (ld >= C, multiple of 8). Weights are repacked once into the
K-major bf16 matrices the kernel reads: [Cout, taps * Cin_pad] with Cin padded to 64 per tap.
"""
import torch

from . import runtime as rt


def _pad64(c):
    return (c + 63) // 64 * 64


def pack_conv_weight(w):
    """nn.Conv2d weight [Cout, Cin, kh, kw] -> bf16 [Cout, kh*kw*Cin_pad], tap-major then channel."""
    co, ci, kh, kw = w.shape
    cp = _pad64(ci)
    m = torch.zeros(co, kh * kw, cp, device=w.device, dtype=torch.bfloat16)
    m[:, :, :ci] = w.detach().permute(0, 2, 3, 1).reshape(co, kh * kw, ci).to(torch.bfloat16)
    return m.reshape(co, kh * kw * cp).contiguous()


def pack_linear_weight(w):
    """nn.Linear weight [N, K] -> bf16 [N, K_pad64]."""
    n, k = w.shape
    kp = _pad64(k)
    m = torch.empty(n, kp, device=w.device, dtype=torch.bfloat16)
    if kp != k:
        m[:, k:].zero_()
    m[:, :k].copy_(w.detach())      # one kernel: dtype conversion and any stride (transposed views) in the copy
    return m


def pack_convT_weight(w):
    """nn.ConvTranspose2d(4, 2, 1) weight [Cin, Cout, 4, 4] -> 4 sub-pixel phase matrices, each bf16
    [Cout, 4 * Cin_pad]. Phase (py, px) of the output reads a 2x2 input neighbourhood:
    oy = 2*iy - 1 + ky  =>  py = 0: (ky=1, dy=0), (ky=3, dy=-1);  py = 1: (ky=0, dy=+1), (ky=2, dy=0)."""
    ci, co = w.shape[0], w.shape[1]
    cp = _pad64(ci)
    taps = {0: ((1, 0), (3, -1)), 1: ((0, 1), (2, 0))}
    out = []
    for py in (0, 1):
        for px in (0, 1):
            m = torch.zeros(co, 4, cp, device=w.device, dtype=torch.bfloat16)
            segs = []
            i = 0
            for ky, dy in taps[py]:
                for kx, dx in taps[px]:
                    m[:, i, :ci] = w.detach()[:, :, ky, kx].t().to(torch.bfloat16)
                    segs.append((0, dy, dx, 0, cp // 64))
                    i += 1
            out.append((m.reshape(co, 4 * cp).contiguous(), segs))
    return out


def conv_same(x, n, h, w, cin, wmat, cout, k, bias=None, ld_x=None, out=None, ld_out=None, c_off=0, rowbias=None,
              residual=None, ld_res=None, act=None, extra=None, out_dtype=torch.bfloat16, stats=None):
    """k x k stride-1 'same' convolution (k in {1, 3}). `extra` = (x2, cin2, ld2): a second source read with a
    1x1 tap whose weights are appended along K (the fused residual_input_conv, blocks.py:104,119)."""
    pad = k // 2
    cp = _pad64(cin)
    ld_x = cin if ld_x is None else ld_x
    srcs = [rt.make_src(x, n, h, w, cin, ld=ld_x)]
    segs = [(0, ky - pad, kx - pad, 0, cp // 64) for ky in range(k) for kx in range(k)]
    if extra is not None:
        x2, cin2, ld2 = extra
        srcs.append(rt.make_src(x2, n, h, w, cin2, ld=ld2))
        segs.append((1, 0, 0, 0, _pad64(cin2) // 64))
    if out is None:
        out = torch.empty(n * h * w, cout, device=x.device, dtype=out_dtype)
        ld_out = cout
    ld_out = out.stride(0) if ld_out is None else ld_out
    o = out if c_off == 0 else out.view(-1)[c_off:]
    rs = None
    if residual is not None:
        ld_res = residual.stride(0) if ld_res is None else ld_res
        rs = (h * w * ld_res, w * ld_res, ld_res)
    rt.igemm(srcs, segs, wmat, cout, n, h, w, o, (h * w * ld_out, w * ld_out, ld_out), bias=bias, rowbias=rowbias,
             residual=residual, res_strides=rs, act=act, stats=stats)
    return out


def conv_down4x4(x, n, h, w, cin, wmat, cout, bias=None, ld_x=None, out=None, stats=None):
    """nn.Conv2d(C, C, 4, 2, 1) (blocks.py:108): iy = 2*oy + ky - 1. The input is read through four stride-2 parity
    views; tap ky uses row parity (ky+1)%2 at coarse row oy + {-1, 0, 0, +1}[ky]."""
    cp = _pad64(cin)
    ld_x = cin if ld_x is None else ld_x
    oh, ow = h // 2, w // 2
    srcs = [rt.make_src(x, n, oh, ow, cin, ld=ld_x, sy=2, sx=2, oy=py, ox=px, hh=h, ww=w) for py in (0, 1)
            for px in (0, 1)]
    par = {0: (1, -1), 1: (0, 0), 2: (1, 0), 3: (0, 1)}
    segs = []
    for ky in range(4):
        for kx in range(4):
            (py, dy), (px, dx) = par[ky], par[kx]
            segs.append((py * 2 + px, dy, dx, 0, cp // 64))
    if out is None:
        out = torch.empty(n * oh * ow, cout, device=x.device, dtype=torch.bfloat16)
    ld = out.stride(0)
    rt.igemm(srcs, segs, wmat, cout, n, oh, ow, out, (oh * ow * ld, ow * ld, ld), bias=bias, stats=stats)
    return out


def conv_up4x4(x, n, h, w, cin, phases, cout, bias=None, ld_x=None, out=None, c_off=0, stats=None):
    """nn.ConvTranspose2d(C, C, 4, 2, 1) (blocks.py:339,457) as four sub-pixel phase GEMMs writing the interleaved
    (2h, 2w) output directly (optionally into a channel slice of a wider concat buffer)."""
    ld_x = cin if ld_x is None else ld_x
    if out is None:
        out = torch.empty(n * 4 * h * w, cout, device=x.device, dtype=torch.bfloat16)
    ld = out.stride(0)
    src = rt.make_src(x, n, h, w, cin, ld=ld_x)
    flat = out.view(-1)
    for idx, (wmat, segs) in enumerate(phases):
        py, px = idx // 2, idx % 2
        o = flat[(py * 2 * w + px) * ld + c_off:]
        st = None
        if stats is not None:  # (tensor, P, ld): each sub-pixel phase owns a quarter of the slots
            st = (stats[0], stats[1], idx * (stats[1] // 4), stats[2])
        rt.igemm([src], segs, wmat, cout, n, h, w, o, (4 * h * w * ld, 4 * w * ld, 2 * ld), bias=bias, stats=st)
    return out


def pack_patch_weight(w, p, cin):
    """PatchEmbedding Linear weight [hidden, (py px c)] -> bf16 [hidden, p*p*64]: one 64-wide K chunk per tap."""
    hidden = w.shape[0]
    if cin > 64:
        raise rt.B200Error("patch embedding: more than 64 input channels per tap is not supported")
    m = torch.zeros(hidden, p * p, 64, device=w.device, dtype=torch.bfloat16)
    m[:, :, :cin] = w.detach().reshape(hidden, p * p, cin).to(torch.bfloat16)
    return m.reshape(hidden, p * p * 64)


def patch_embed(xbuf, n, H, W, cin, p, wmat, hidden, bias=None, pos=None):
    """tokens[(img, gy, gx)] = W . fold_{p x p}(x)[gy, gx] + bias + pos[(gy, gx)] (models/patch_embed.py:84-95) as an
    implicit GEMM over p*p strided views of the NHWC buffer xbuf [n*H*W, ld]."""
    if H % p or W % p:
        raise rt.B200Error("patch embedding: image size must be divisible by the patch size")
    gh, gw = H // p, W // p
    ld = xbuf.stride(0)
    srcs = [rt.make_src(xbuf, n, gh, gw, cin, ld=ld, sy=p, sx=p, oy=t // p, ox=t % p, hh=H, ww=W) for t in range(p * p)]
    segs = [(t, 0, 0, 0, 1) for t in range(p * p)]
    out = torch.empty(n * gh * gw, hidden, device=xbuf.device, dtype=torch.bfloat16)
    rs = None
    if pos is not None:
        rs = (0, gw * pos.stride(0), pos.stride(0))   # the same position rows for every image
    rt.igemm(srcs, segs, wmat, hidden, n, gh, gw, out, (gh * gw * hidden, gw * hidden, hidden), bias=bias, residual=pos,
             res_strides=rs)
    return out

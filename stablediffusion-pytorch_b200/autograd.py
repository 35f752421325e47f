"""torch.autograd.Function wrappers that give the sm_100a kernels a backward pass, so `loss.backward()` of the
reference's training loops (train_ddpm_cond_celebhq_multi_gpu.py:341-374, tools/train_ddpm_vqvae.py:95-107) flows
through the drop-in modules. Forward AND backward of every heavy op run on libb200sd.so:

    leaf                      forward                     backward
    Conv2d 3x3/1x1 (+1x1)     igemm                       dgrad = igemm (flipped / transposed weights), wgrad = b200sd_wgrad,
    Conv2d 4x4 s2             igemm (parity views)          bias / time-embedding grads = b200sd_colsum
    ConvTranspose2d 4x4 s2    4 phase igemms              dgrad = the stride-2 conv, wgrad per phase
    Linear                    igemm                       dgrad = igemm, wgrad = b200sd_wgrad
    GroupNorm + act           groupnorm_apply             b200sd_groupnorm_bwd
    attention                 attention_lse               attention_bwd (scores recomputed)

Activations and their gradients are NHWC bf16 2-D tensors; parameters and their gradients stay fp32 in the
reference's layouts (what DistributedDataParallel buckets and the optimizer see). What remains on torch in training
mode is glue on small tensors only: the channel concat of the skip connection, gradient accumulation where a tensor
has two consumers, SiLU on the (B, 512) time embedding, and un-packing weight gradients into parameter layout.
"""
import torch
from torch.autograd import Function

from . import bank as BK
from . import ops
from . import runtime as rt


_Side = rt.Side     # parameter gradients as a branch beside the input-gradient GEMM (runtime.Side)


def _bank(*ws):
    """The active weight bank if every given weight is parameter storage it knows (captured training step), else None."""
    b = BK.ACTIVE
    if b is None or not all(b.knows(w) for w in ws):
        return None
    return b


def _phase_taps(cp):
    """Sub-pixel phases of ConvTranspose2d(4, 2, 1): ([(ky, kx)] taps, igemm segments) per output parity (py, px)."""
    taps = {0: ((1, 0), (3, -1)), 1: ((0, 1), (2, 0))}
    out = []
    for py in (0, 1):
        for px in (0, 1):
            tl, segs = [], []
            for ky, dy in taps[py]:
                for kx, dx in taps[px]:
                    tl.append((ky, kx))
                    segs.append((0, dy, dx, 0, cp // 64))
            out.append((tl, segs))
    return out


def _pad8(c):
    return (c + 7) // 8 * 8


def _rows(t, c):
    """bf16 gradient as a 2-D row matrix with 16-byte aligned rows."""
    t = t if t.dtype == torch.bfloat16 else t.to(torch.bfloat16)
    if t.stride(-1) != 1 or t.stride(0) % 8 or t.data_ptr() % 16:
        if c % 8 == 0:
            return t.contiguous()
        buf = torch.zeros(t.shape[0], _pad8(c), device=t.device, dtype=torch.bfloat16)
        buf[:, :c] = t
        return buf[:, :c]
    return t


def _grad_buf(x, c):
    """Gradient buffer for an activation whose row matrix may carry padding columns beyond its c channels."""
    if x.shape[1] == c:
        return torch.empty(x.shape[0], _pad8(c), device=x.device, dtype=torch.bfloat16)
    return torch.zeros(x.shape[0], x.shape[1], device=x.device, dtype=torch.bfloat16)


def _stats_arg(st):
    return None if st is None else (st[0], st[1], 0, st[2])


def _new_stats(n, oh, ow, c, device, phases=1):
    if c <= 32 or c % 8:
        return None
    p = rt.stats_slots(oh, ow) * phases
    return (torch.empty(n, p, c, 2, device=device, dtype=torch.float32), p, c)


# ---------------------------------------------------------------------------------------------------- Linear
class LinearFn(Function):
    """y = x @ W^T + b (+ residual); x bf16 [M, K] rows, W fp32 [N, K]. Returns (y, stats or empty)."""

    @staticmethod
    def forward(ctx, x, weight, bias, residual, rows_per_img, want_stats, act=None):
        M, K = x.shape
        N = weight.shape[0]
        b = _bank(weight)
        wp = b.linear(weight) if b else ops.pack_linear_weight(weight)
        st = None
        if want_stats and rows_per_img:
            st = _new_stats(M // rows_per_img, 1, rows_per_img, N, x.device)
        out = torch.empty(M, _pad8(N), device=x.device, dtype=torch.bfloat16)[:, :N]
        if act not in (None, "relu") or (act and residual is not None):
            raise rt.B200Error("LinearFn: only a ReLU epilogue (without residual) has a backward")
        rt.linear(x, wp, bias=None if bias is None else bias.detach().float(), residual=residual, out=out, act=act,
                  rows_per_img=rows_per_img if st is not None else None, stats=_stats_arg(st))
        ctx.save_for_backward(x, weight, out if act else None)
        ctx.has_bias, ctx.has_res = bias is not None, residual is not None
        stats_t = st[0] if st is not None else torch.empty(0, device=x.device)
        ctx.mark_non_differentiable(stats_t)
        ctx.set_materialize_grads(False)   # no zero-filled gradient tensor for the statistics output
        return out, stats_t

    @staticmethod
    def backward(ctx, dy, _):
        x, weight, y_act = ctx.saved_tensors
        M, K = x.shape
        N = weight.shape[0]
        dy = _rows(dy, N)
        if y_act is not None:
            dy = rt.relu_bwd(dy, y_act)
        dx = dw = db = dres = gv = None
        want_b = ctx.has_bias and ctx.needs_input_grad[2]
        with _Side(ctx.needs_input_grad[0] and (ctx.needs_input_grad[1] or want_b), x.device) as side:
            if ctx.needs_input_grad[1]:
                kp = (K + 63) // 64 * 64
                src = rt.make_src(x, 1, 1, M, K, ld=x.stride(0))
                b = _bank(weight)
                gv = b.grad_view(weight) if b else None
                dwp = rt.wgrad([src], [(0, 0, 0, 0, kp // 64)], rt.make_src(dy, 1, 1, M, N, ld=dy.stride(0)), N, 1, 1, M,
                               scatter=None if gv is None else [(gv, 0, K, kp, 1, K, 1, [0])])
                dw = gv if gv is not None else dwp[:, :K]
            if want_b:
                db = rt.colsum(dy, 1, M, N) if N % 8 == 0 else dy.float().sum(0)
        if ctx.needs_input_grad[0]:
            b = _bank(weight)
            wt = b.linear_t(weight) if b else ops.pack_linear_weight(weight.detach().t())  # [K, N_pad64]
            dx = torch.empty(M, _pad8(K), device=x.device, dtype=torch.bfloat16)[:, :K]
            rt.linear(dy, wt, out=dx)
        side.join(dy, x, defer_ok=not ctx.needs_input_grad[1] or (dw is not None and gv is not None))
        if ctx.has_res and ctx.needs_input_grad[3]:
            dres = dy
        return dx, dw, db, dres, None, None, None


class CrossProjFn(Function):
    """The two input projections of a cross-attention site from ONE packed in_proj parameter (nn.MultiheadAttention,
    blocks.py:138-141): q = xq @ W[:C]^T + b[:C];  kv = xkv @ W[C:]^T + b[C:]. One autograd node for the parameter:
    its gradient is assembled in place (row ranges of one [3C, C] tensor) instead of two sliced gradients that
    autograd would zero-pad and add."""

    @staticmethod
    def forward(ctx, xq, xkv, weight, bias, C):
        bk = _bank(weight)
        wd = weight.detach()
        wq = bk.linear(weight, 0, C) if bk else ops.pack_linear_weight(wd[:C])
        wkv = bk.linear(weight, C, 2 * C) if bk else ops.pack_linear_weight(wd[C:])
        bf = None if bias is None else bias.detach().float()
        q = torch.empty(xq.shape[0], C, device=xq.device, dtype=torch.bfloat16)
        kv = torch.empty(xkv.shape[0], 2 * C, device=xq.device, dtype=torch.bfloat16)
        rt.linear(xq, wq, bias=None if bf is None else bf[:C], out=q)
        rt.linear(xkv, wkv, bias=None if bf is None else bf[C:], out=kv)
        ctx.save_for_backward(xq, xkv, weight)
        ctx.C, ctx.has_bias = C, bias is not None
        ctx.set_materialize_grads(False)
        return q, kv

    @staticmethod
    def backward(ctx, dq, dkv):
        xq, xkv, weight = ctx.saved_tensors
        C = ctx.C
        K = weight.shape[1]
        kp = (K + 63) // 64 * 64
        dev = xq.device
        bk = _bank(weight)
        wd = weight.detach()
        dxq = dxkv = dw = db = gv = None
        parts = ((dq, xq, 0, C, 0), (dkv, xkv, C, 2 * C, 1))
        if ctx.needs_input_grad[2]:
            gv = bk.grad_view(weight) if bk else None
            dw = gv if gv is not None else torch.zeros(3 * C, K, device=dev, dtype=torch.float32)
            if gv is not None and (dq is None or dkv is None):
                (dw[:C] if dq is None else dw[C:]).zero_()
        if ctx.has_bias and ctx.needs_input_grad[3]:
            db = torch.empty(3 * C, device=dev, dtype=torch.float32)
            if dq is None or dkv is None or C % 8:
                db.zero_()
        parts = [(None if dy is None else _rows(dy, rn), x, r0, rn, slot) for dy, x, r0, rn, slot in parts]
        with _Side((ctx.needs_input_grad[0] or ctx.needs_input_grad[1]) and (dw is not None or db is not None), dev) as side:
            for dy, x, r0, rn, slot in parts:       # parameter gradients: a branch of their own (see _Side)
                if dy is None:
                    continue
                M = x.shape[0]
                if dw is not None:
                    rt.wgrad([rt.make_src(x, 1, 1, M, K, ld=x.stride(0))], [(0, 0, 0, 0, kp // 64)],
                             rt.make_src(dy, 1, 1, M, rn, ld=dy.stride(0)), rn, 1, 1, M,
                             scatter=[(dw[r0:r0 + rn], 0, K, kp, 1, K, 1, [0])])
                if db is not None:
                    if rn % 8 == 0:
                        rt.colsum(dy, 1, M, rn, out=db[r0:r0 + rn])
                    else:
                        db[r0:r0 + rn] = dy.float().sum(0)
        for dy, x, r0, rn, slot in parts:
            if dy is None:
                continue
            M = x.shape[0]
            if ctx.needs_input_grad[slot]:
                wt = bk.linear_t(weight, r0, rn) if bk else ops.pack_linear_weight(wd[r0:r0 + rn].t())
                dx = torch.empty(M, _pad8(K), device=dev, dtype=torch.bfloat16)[:, :K]
                rt.linear(dy, wt, out=dx)
                if slot == 0:
                    dxq = dx
                else:
                    dxkv = dx
        side.join(xq, xkv, *[t for t, *_ in parts], defer_ok=dw is None or gv is not None)
        return dxq, dxkv, dw, db, None


# ---------------------------------------------------------------------------------------------------- Conv 3x3 / 1x1
class ConvSameFn(Function):
    """k x k stride-1 'same' conv (+ per-image row bias, + fused 1x1 conv of a second tensor into the accumulator).
    x: bf16 [n*h*w, >=cin]; weight fp32 [cout, cin, k, k]; x2/w2/b2: the ResNet shortcut (blocks.py:104,119)."""

    @staticmethod
    def forward(ctx, x, weight, bias, rowbias, x2, w2, b2, geom, want_stats, out_f32):
        n, h, w, cin, cin2 = geom
        cout, k = weight.shape[0], weight.shape[2]
        bk = _bank(weight) if x2 is None else _bank(weight, w2)
        wmat = bk.conv(weight, w2) if bk else ops.pack_conv_weight(weight)
        b = None if bias is None else bias.detach().float()
        extra = None
        if x2 is not None:
            if not bk:
                wmat = torch.cat([wmat, ops.pack_conv_weight(w2)], dim=1).contiguous()
            if b2 is not None:
                b = b2.detach().float() if b is None else b + b2.detach().float()
            extra = (x2, cin2, x2.stride(0))
        st = _new_stats(n, h, w, cout, x.device) if (want_stats and not out_f32) else None
        out = torch.empty(n * h * w, _pad8(cout), device=x.device,
                          dtype=torch.float32 if out_f32 else torch.bfloat16)[:, :cout]
        ops.conv_same(x, n, h, w, cin, wmat, cout, k, bias=b, ld_x=x.stride(0), out=out, rowbias=rowbias, extra=extra,
                      stats=_stats_arg(st))
        ctx.save_for_backward(x, weight, x2, w2)
        ctx.geom, ctx.flags = geom, (bias is not None, rowbias is not None, b2 is not None)
        stats_t = st[0] if st is not None else torch.empty(0, device=x.device)
        ctx.mark_non_differentiable(stats_t)
        ctx.set_materialize_grads(False)   # no zero-filled gradient tensor for the statistics output
        return out, stats_t

    @staticmethod
    def backward(ctx, dy, _):
        x, weight, x2, w2 = ctx.saved_tensors
        n, h, w, cin, cin2 = ctx.geom
        has_b, has_rb, has_b2 = ctx.flags
        cout, k = weight.shape[0], weight.shape[2]
        dy = _rows(dy, cout)
        P = n * h * w
        dx = dw = db = drb = dx2 = dw2 = db2 = sc = None
        want_bias = (has_b and ctx.needs_input_grad[2]) or (has_b2 and ctx.needs_input_grad[6]) or has_rb
        want_w = ctx.needs_input_grad[1] or (w2 is not None and ctx.needs_input_grad[5])

        def bias_sums():
            if cout % 8 == 0:
                per = rt.colsum(dy, n, h * w, cout, per_image=True)  # fp32 [n, cout]
            else:
                per = dy.float().reshape(n, h * w, cout).sum(1)
            tot = per.sum(0) if (has_b or has_b2) else None
            # (never hand one buffer to two parameters)
            return (tot if has_b else None), ((tot.clone() if has_b else tot) if has_b2 else None), \
                (per.to(torch.bfloat16) if has_rb else None)

        with _Side((ctx.needs_input_grad[0] or (x2 is not None and ctx.needs_input_grad[4])) and (want_w or want_bias),
                   x.device) as side:
            if ctx.needs_input_grad[1] or (w2 is not None and ctx.needs_input_grad[5]):
                pad, cp = k // 2, ops._pad64(cin)
                srcs = [rt.make_src(x, n, h, w, cin, ld=x.stride(0))]
                segs = [(0, ky - pad, kx - pad, 0, cp // 64) for ky in range(k) for kx in range(k)]
                if x2 is not None:
                    srcs.append(rt.make_src(x2, n, h, w, cin2, ld=x2.stride(0)))
                    segs.append((1, 0, 0, 0, ops._pad64(cin2) // 64))
                bk = _bank(weight) if x2 is None else _bank(weight, w2)
                gv = bk.grad_view(weight) if bk else None
                gv2 = bk.grad_view(w2) if (bk and x2 is not None) else None
                sc = None
                if gv is not None and (x2 is None or gv2 is not None):
                    kk = k * k
                    sc = [(gv, 0, cin, cp, kk, cin * kk, kk, list(range(kk)))]
                    if x2 is not None:
                        sc.append((gv2, kk * cp, cin2, ops._pad64(cin2), 1, cin2, 1, [0]))
                dwp = rt.wgrad(srcs, segs, rt.make_src(dy, n, h, w, cout, ld=dy.stride(0)), cout, n, h, w, scatter=sc)
                if sc is not None:
                    dw, dw2 = gv, gv2
                else:
                    dw = dwp[:, :k * k * cp].reshape(cout, k * k, cp)[:, :, :cin].permute(0, 2, 1).reshape(cout, cin, k, k)
                    if x2 is not None:
                        dw2 = dwp[:, k * k * cp:k * k * cp + cin2].reshape(cout, cin2, 1, 1)
            if want_bias and not has_rb:
                db, db2, drb = bias_sums()
        if want_bias and has_rb:
            # the row-bias gradient feeds the time-embedding branch of THIS backward pass: computed on the main stream,
            # so that the weight gradient beside it can still join late
            db, db2, drb = bias_sums()
        if ctx.needs_input_grad[0]:
            # dgrad of a stride-1 'same' conv = the same conv with taps flipped and in/out channels swapped
            bk = _bank(weight)
            wd = bk.conv_dgrad(weight) if bk else ops.pack_conv_weight(weight.detach().flip(2, 3).transpose(0, 1))
            full = _grad_buf(x, cin)
            dx = full[:, :cin]
            ops.conv_same(dy, n, h, w, cout, wd, cin, k, ld_x=dy.stride(0), out=dx)
            dx = full if full.shape[1] == x.shape[1] else dx
        if x2 is not None and ctx.needs_input_grad[4]:
            bk = _bank(w2)
            wt = bk.linear_t(w2) if bk else ops.pack_linear_weight(w2.detach().reshape(cout, cin2).t())
            full = _grad_buf(x2, cin2)
            dx2 = full[:, :cin2]
            rt.linear(dy, wt, out=dx2)
            dx2 = full if full.shape[1] == x2.shape[1] else dx2
        side.join(dy, x, x2, defer_ok=not want_w or sc is not None)
        return dx, dw, db, drb, dx2, dw2, db2, None, None, None


# ---------------------------------------------------------------------------------------------------- Conv 4x4 s2
def _down_segs(cp):
    par = {0: (1, -1), 1: (0, 0), 2: (1, 0), 3: (0, 1)}
    segs = []
    for ky in range(4):
        for kx in range(4):
            (py, dy), (px, dx) = par[ky], par[kx]
            segs.append((py * 2 + px, dy, dx, 0, cp // 64))
    return segs


class ConvDownFn(Function):
    """nn.Conv2d(C, C, 4, 2, 1) (blocks.py:108)."""

    @staticmethod
    def forward(ctx, x, weight, bias, geom, want_stats):
        n, h, w, cin = geom
        cout = weight.shape[0]
        st = _new_stats(n, h // 2, w // 2, cout, x.device) if want_stats else None
        out = torch.empty(n * (h // 2) * (w // 2), cout, device=x.device, dtype=torch.bfloat16)
        bk = _bank(weight)
        ops.conv_down4x4(x, n, h, w, cin, bk.conv(weight) if bk else ops.pack_conv_weight(weight), cout,
                         bias=None if bias is None else bias.detach().float(), ld_x=x.stride(0), out=out,
                         stats=_stats_arg(st))
        ctx.save_for_backward(x, weight)
        ctx.geom, ctx.has_bias = geom, bias is not None
        stats_t = st[0] if st is not None else torch.empty(0, device=x.device)
        ctx.mark_non_differentiable(stats_t)
        ctx.set_materialize_grads(False)   # no zero-filled gradient tensor for the statistics output
        return out, stats_t

    @staticmethod
    def backward(ctx, dy, _):
        x, weight = ctx.saved_tensors
        n, h, w, cin = ctx.geom
        cout = weight.shape[0]
        oh, ow = h // 2, w // 2
        dy = _rows(dy, cout)
        dx = dw = db = gv = None
        want_b = ctx.has_bias and ctx.needs_input_grad[2]
        with _Side(ctx.needs_input_grad[0] and (ctx.needs_input_grad[1] or want_b), x.device) as side:
            if ctx.needs_input_grad[1]:
                cp = ops._pad64(cin)
                ld = x.stride(0)
                srcs = [rt.make_src(x, n, oh, ow, cin, ld=ld, sy=2, sx=2, oy=py, ox=px, hh=h, ww=w) for py in (0, 1)
                        for px in (0, 1)]
                bk = _bank(weight)
                gv = bk.grad_view(weight) if bk else None
                dwp = rt.wgrad(srcs, _down_segs(cp), rt.make_src(dy, n, oh, ow, cout, ld=dy.stride(0)), cout, n, oh, ow,
                               scatter=None if gv is None else [(gv, 0, cin, cp, 16, cin * 16, 16, list(range(16)))])
                dw = gv if gv is not None else dwp.reshape(cout, 16, cp)[:, :, :cin].permute(0, 2, 1).reshape(cout, cin, 4, 4)
            if want_b:
                db = rt.colsum(dy, n, oh * ow, cout)
        if ctx.needs_input_grad[0]:
            # adjoint of Conv2d(4, 2, 1) with weight [co, ci] = ConvTranspose2d(4, 2, 1) with the same tensor as [in=co, out=ci]
            bk = _bank(weight)
            phases = bk.convT_phases(weight, _phase_taps(ops._pad64(cout))) if bk else ops.pack_convT_weight(weight.detach())
            dx = torch.empty(n * h * w, cin, device=x.device, dtype=torch.bfloat16)
            ops.conv_up4x4(dy, n, oh, ow, cout, phases, cin, ld_x=dy.stride(0), out=dx)
        side.join(dy, x, defer_ok=not ctx.needs_input_grad[1] or gv is not None)
        return dx, dw, db, None, None


# ---------------------------------------------------------------------------------------------------- ConvTranspose 4x4 s2
_T_TAPS = {0: ((1, 0), (3, -1)), 1: ((0, 1), (2, 0))}


class ConvUpFn(Function):
    """nn.ConvTranspose2d(C, C, 4, 2, 1) (blocks.py:339,457) as four sub-pixel phase GEMMs."""

    @staticmethod
    def forward(ctx, x, weight, bias, geom, want_stats):
        n, h, w, cin = geom
        cout = weight.shape[1]
        st = _new_stats(n, h, w, cout, x.device, phases=4) if want_stats else None
        out = torch.empty(n * 4 * h * w, cout, device=x.device, dtype=torch.bfloat16)
        bk = _bank(weight)
        phases = bk.convT_phases(weight, _phase_taps(ops._pad64(cin))) if bk else ops.pack_convT_weight(weight)
        ops.conv_up4x4(x, n, h, w, cin, phases, cout,
                       bias=None if bias is None else bias.detach().float(), ld_x=x.stride(0), out=out, stats=st)
        ctx.save_for_backward(x, weight)
        ctx.geom, ctx.has_bias = geom, bias is not None
        stats_t = st[0] if st is not None else torch.empty(0, device=x.device)
        ctx.mark_non_differentiable(stats_t)
        ctx.set_materialize_grads(False)   # no zero-filled gradient tensor for the statistics output
        return out, stats_t

    @staticmethod
    def backward(ctx, dy, _):
        x, weight = ctx.saved_tensors
        n, h, w, cin = ctx.geom
        cout = weight.shape[1]
        dy = _rows(dy, cout)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            # adjoint of ConvTranspose2d(4, 2, 1) with weight [ci, co] = Conv2d(4, 2, 1) with the same tensor as [out=ci, in=co]
            dx = torch.empty(n * h * w, cin, device=x.device, dtype=torch.bfloat16)
            bk = _bank(weight)
            ops.conv_down4x4(dy, n, 2 * h, 2 * w, cout, bk.conv(weight) if bk else ops.pack_conv_weight(weight.detach()),
                             cin, ld_x=dy.stride(0), out=dx)
        if ctx.needs_input_grad[1]:
            cp = ops._pad64(cin)
            src = rt.make_src(x, n, h, w, cin, ld=x.stride(0))
            bk = _bank(weight)
            gv = bk.grad_view(weight) if bk else None
            dw = gv if gv is not None else torch.empty_like(weight)
            ld = dy.stride(0)
            for py in (0, 1):
                for px in (0, 1):
                    segs, taps = [], []
                    for ky, ddy in _T_TAPS[py]:
                        for kx, ddx in _T_TAPS[px]:
                            segs.append((0, ddy, ddx, 0, cp // 64))
                            taps.append((ky, kx))
                    dyv = rt.make_src(dy, n, h, w, cout, ld=ld, sy=2, sx=2, oy=py, ox=px, hh=2 * h, ww=2 * w)
                    if gv is not None:   # dw[ci, co, ky, kx]: rows of the GEMM are co, channels ci
                        rt.wgrad([src], segs, dyv, cout, n, h, w,
                                 scatter=[(gv, 0, cin, cp, 4, 16, cout * 16, [ky * 4 + kx for ky, kx in taps])])
                        continue
                    dwp = rt.wgrad([src], segs, dyv, cout, n, h, w).reshape(cout, 4, cp)
                    for i, (ky, kx) in enumerate(taps):
                        dw[:, :, ky, kx] = dwp[:, i, :cin].t()
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = rt.colsum(dy, n, 4 * h * w, cout)
        return dx, dw, db, None, None


# ---------------------------------------------------------------------------------------------------- GroupNorm
class GroupNormFn(Function):
    """act(GroupNorm(x)) from the producer's partial statistics; meta = (n, hw, c, groups, eps, act, seg_meta) with
    seg_meta = [(P, ld, channels), ...] describing st0 / st1."""

    @staticmethod
    def forward(ctx, x, gamma, beta, st0, st1, meta):
        n, hw, c, groups, eps, act, seg_meta = meta
        segs = [(st0,) + tuple(seg_meta[0])]
        if st1 is not None:
            segs.append((st1,) + tuple(seg_meta[1]))
        out = torch.empty(n * hw, c, device=x.device, dtype=torch.bfloat16)
        rt.groupnorm_apply(x, n, hw, c, groups, gamma.detach().float(), beta.detach().float(), segs, eps, act=act,
                           out=out, ld_x=x.stride(0), ld_y=c)
        ctx.save_for_backward(x, gamma, beta, st0, st1)
        ctx.meta = meta
        return out

    @staticmethod
    def backward(ctx, dy):
        x, gamma, beta, st0, st1 = ctx.saved_tensors
        n, hw, c, groups, eps, act, seg_meta = ctx.meta
        segs = [(st0,) + tuple(seg_meta[0])]
        if st1 is not None:
            segs.append((st1,) + tuple(seg_meta[1]))
        dy = _rows(dy, c)
        dx, dgamma, dbeta = rt.groupnorm_bwd(x, dy, n, hw, c, groups, gamma.detach().float(), beta.detach().float(), segs,
                                             eps, act=act)
        return dx, dgamma, dbeta, None, None, None


def tensor_stats(x, n, hw, c):
    """Partial statistics of a tensor that did not come out of an igemm epilogue (a block called stand-alone on an
    NCHW input): one slot per image. Off the model's hot path; small reduction on torch."""
    xf = x[:, :c].float().reshape(n, hw, c)
    return (torch.stack([xf.sum(1), (xf * xf).sum(1)], dim=-1).reshape(n, 1, c, 2).contiguous(), 1, c)


# ---------------------------------------------------------------------------------------------------- attention
class AttentionFn(Function):
    """softmax(Q K^T / sqrt(d)) V. `packed` bf16 holds q | k | v side by side (self-attention, [n*l, 3C]) or k | v
    ([n*lk, 2C], cross-attention, with q given separately); gradients come back in the same packed buffers."""

    @staticmethod
    def forward(ctx, q, packed, meta):
        n, heads, lq, lk, d = meta
        C = heads * d
        if q is None:
            qv, kv, vv = packed[:, :C], packed[:, C:2 * C], packed[:, 2 * C:]
        else:
            qv, kv, vv = q, packed[:, :C], packed[:, C:]
        out, lse = rt.attention_lse(qv, kv, vv, n, heads, lq, lk, d)
        ctx.save_for_backward(q, packed, out, lse)
        ctx.meta = meta
        return out

    @staticmethod
    def backward(ctx, d_o):
        q, packed, out, lse = ctx.saved_tensors
        n, heads, lq, lk, d = ctx.meta
        C = heads * d
        d_o = _rows(d_o, C)
        dpk = torch.empty_like(packed)
        if q is None:
            rt.attention_bwd(packed[:, :C], packed[:, C:2 * C], packed[:, 2 * C:], out, d_o, lse, n, heads, lq, lk, d,
                             dpk[:, :C], dpk[:, C:2 * C], dpk[:, 2 * C:])
            return None, dpk, None
        dq = torch.empty_like(q)
        rt.attention_bwd(q, packed[:, :C], packed[:, C:], out, d_o, lse, n, heads, lq, lk, d, dq, dpk[:, :C], dpk[:, C:])
        return dq, dpk, None


# ---------------------------------------------------------------------------------------------------- boundary
class ToNchwFn(Function):
    """NHWC rows (bf16 or fp32) -> NCHW fp32 model output; the incoming NCHW fp32 gradient returns as NHWC bf16."""

    @staticmethod
    def forward(ctx, x, geom):
        n, c, h, w = geom
        ctx.geom, ctx.ld, ctx.cols, ctx.dt = geom, x.stride(0), x.shape[1], x.dtype
        return rt.nhwc_to_nchw(x, n, c, h, w, ld=x.stride(0))

    @staticmethod
    def backward(ctx, g):
        n, c, h, w = ctx.geom
        width = max(_pad8(c), ctx.cols)   # a quantised conv_out hands over rows padded to 8 columns
        buf = torch.zeros(n * h * w, width, device=g.device, dtype=torch.bfloat16)
        rt.nchw_to_nhwc(g.float().contiguous(), buf, 0)
        return buf[:, :ctx.cols].to(ctx.dt), None


class FromNchwFn(Function):
    """NCHW fp32 -> NHWC bf16 rows for an input that itself needs a gradient (blocks called stand-alone)."""

    @staticmethod
    def forward(ctx, x, ld):
        n, c, h, w = x.shape
        buf = torch.zeros(n * h * w, ld, device=x.device, dtype=torch.bfloat16)
        rt.nchw_to_nhwc(x.detach().float().contiguous(), buf, 0)
        ctx.geom = (n, c, h, w)
        return buf

    @staticmethod
    def backward(ctx, g):
        n, c, h, w = ctx.geom
        g = _rows(g, g.shape[1])
        return rt.nhwc_to_nchw(g, n, c, h, w, ld=g.stride(0)), None


class MaskCondFn(Function):
    """Mask branch (unet_cond_base.py:134-138): nearest resize + 1x1 conv (no bias) written into channels
    [c_off, c_off+cout) of the conv_in input buffer. The 18->3 weight gradient is a (3 x 18) contraction over the
    sampled pixels — tiny, done on torch."""

    @staticmethod
    def forward(ctx, weight, mask, buf, c_off, h, w):
        cout = weight.shape[0]
        rt.mask_cond(mask.float(), weight.detach().float().reshape(cout, -1), buf, c_off, h, w)
        ctx.save_for_backward(mask)
        ctx.meta = (c_off, h, w, cout)
        ctx.mark_dirty(buf)
        return buf

    @staticmethod
    def backward(ctx, g):
        (mask,) = ctx.saved_tensors
        c_off, h, w, cout = ctx.meta
        n, cin, mh, mw = mask.shape
        sy, sx = mh // h, mw // w
        ms = mask[:, :, ::sy, ::sx][:, :, :h, :w].float().permute(0, 2, 3, 1).reshape(n * h * w, cin)
        dw = g[:, c_off:c_off + cout].float().t() @ ms
        return dw.reshape(cout, cin, 1, 1), None, g, None, None, None


# ---------------------------------------------------------------------------------------------------- DiT rows
class LayerNormModFn(Function):
    """LayerNorm(no affine, eps) * (1 + scale[img]) + shift[img] on bf16 token rows (transformer_layer.py:88-89,
    103-104; transformer.py:196-197); shift = scale = None is the plain cross_attn_norm (:94)."""

    @staticmethod
    def forward(ctx, x, shift, scale, meta):
        n, rows_per_img, c, eps = meta
        y, stats = rt.layernorm_mod(x, n, rows_per_img, c, shift=shift, scale=scale, eps=eps, want_stats=True)
        ctx.save_for_backward(x, stats, scale)
        ctx.meta = meta
        return y

    @staticmethod
    def backward(ctx, dy):
        x, stats, scale = ctx.saved_tensors
        n, rows_per_img, c, _ = ctx.meta
        dy = _rows(dy, c)
        dx, dshift, dscale = rt.layernorm_mod_bwd(dy, x, stats, n, rows_per_img, c, scale=scale)
        if scale is None:
            return dx, None, None, None
        return dx, dshift.to(torch.bfloat16), dscale.to(torch.bfloat16), None


class GateResidualFn(Function):
    """y = x + gate[img] * h (adaLN-Zero gate, transformer_layer.py:90,105). Inference fuses this into the igemm
    epilogue; training keeps h, which the gate gradient needs."""

    @staticmethod
    def forward(ctx, x, h, gate, meta):
        n, rows_per_img, c = meta
        ctx.save_for_backward(h, gate)
        ctx.meta = meta
        return rt.gate_residual(x, h, gate, n, rows_per_img, c)

    @staticmethod
    def backward(ctx, dy):
        h, gate = ctx.saved_tensors
        n, rows_per_img, c = ctx.meta
        dy = _rows(dy, c)
        dh, dgate = rt.gate_residual_bwd(dy, h, gate, n, rows_per_img, c)
        return dy, dh, dgate.to(torch.bfloat16), None


class PatchEmbedFn(Function):
    """PatchEmbedding (patch_embed.py:75-96) on the NHWC input buffer: the (ph pw c) fold + Linear is a p x p stride-p
    convolution, run as one implicit GEMM over p*p strided views of the buffer, with the position embedding as the
    epilogue's (batch-broadcast) residual. xbuf: bf16 [n*H*W, ld]; weight fp32 [hidden, p*p*cin]; pos bf16 [gh*gw, hidden]."""

    @staticmethod
    def forward(ctx, xbuf, weight, bias, pos, geom):
        n, H, W, cin, p = geom
        out = ops.patch_embed(xbuf, n, H, W, cin, p, ops.pack_patch_weight(weight, p, cin), weight.shape[0],
                              bias=None if bias is None else bias.detach().float(), pos=pos)
        ctx.save_for_backward(xbuf, weight)
        ctx.geom, ctx.has_bias = geom, bias is not None
        return out

    @staticmethod
    def backward(ctx, dy):
        xbuf, weight = ctx.saved_tensors
        n, H, W, cin, p = ctx.geom
        hidden = weight.shape[0]
        gh, gw = H // p, W // p
        dy = _rows(dy, hidden)
        dx = dw = db = None
        ld = xbuf.stride(0)
        if ctx.needs_input_grad[0]:
            # adjoint: each (py, px) phase of the input grid is a GEMM against that tap's [cin, hidden] weights
            dx = torch.zeros_like(xbuf)
            wv = weight.detach().reshape(hidden, p * p, cin)
            flat = dx.view(-1)
            for t in range(p * p):
                py, px = t // p, t % p
                wt = ops.pack_linear_weight(wv[:, t, :].t())           # [cin, hidden_pad64]
                o = flat[(py * W + px) * ld:]
                src = rt.make_src(dy, n, gh, gw, hidden, ld=dy.stride(0))
                rt.igemm([src], [(0, 0, 0, 0, wt.shape[1] // 64)], wt, cin, n, gh, gw, o, (H * W * ld, p * W * ld, p * ld))
        if ctx.needs_input_grad[1]:
            srcs = [rt.make_src(xbuf, n, gh, gw, cin, ld=ld, sy=p, sx=p, oy=t // p, ox=t % p, hh=H, ww=W)
                    for t in range(p * p)]
            segs = [(t, 0, 0, 0, 1) for t in range(p * p)]
            dwp = rt.wgrad(srcs, segs, rt.make_src(dy, n, gh, gw, hidden, ld=dy.stride(0)), hidden, n, gh, gw)
            dw = dwp.reshape(hidden, p * p, 64)[:, :, :cin].reshape(hidden, p * p * cin)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = rt.colsum(dy, 1, n * gh * gw, hidden)
        return dx, dw, db, None, None


class UnpatchifyFn(Function):
    """Token rows -> fp32 NCHW model output (transformer.py:203-211); the NCHW gradient folds back into token rows."""

    @staticmethod
    def forward(ctx, tokens, geom):
        n, ch, gh, gw, p = geom
        ctx.geom, ctx.ld = geom, tokens.shape[1]
        return rt.unpatchify(tokens, n, ch, gh, gw, p, p)

    @staticmethod
    def backward(ctx, g):
        n, ch, gh, gw, p = ctx.geom
        tok = rt.patchify(g.float(), p, p, ld=(ctx.ld + 7) // 8 * 8)
        return tok[:, :ctx.ld], None

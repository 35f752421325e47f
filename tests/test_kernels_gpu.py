"""Per-kernel numerics through the C-ABI against plain PyTorch fp32 references of the same op (floating-point
kernels) on bf16-rounded inputs. Tolerances are written at each assert. Model-level parity against the oracle and
the golden fixtures lives in test_parity_gpu.py."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def relerr(a, b):
    """max |a-b| / max |b| — the 'max relative error' used throughout (per-element ratios blow up at zeros)."""
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-20)).item()


def bf(x):
    return x.to(torch.bfloat16)


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 256, 512), (1000, 384, 200), (64, 1536, 512),
                                   (4096, 192, 1152), (77 * 4, 96, 512), (128, 4, 128), (8192, 768, 768)])
def test_linear(rt, M, N, K):
    dev = _dev()
    from sdb200 import ops
    g = torch.Generator(device="cpu").manual_seed(M * 7 + N * 3 + K)
    Kld = (K + 7) // 8 * 8
    a = torch.zeros(M, Kld, dtype=torch.bfloat16)
    a[:, :K] = bf(torch.randn(M, K, generator=g))
    w = torch.randn(N, K, generator=g) / K ** 0.5
    b = torch.randn(N, generator=g)
    a_d = a.to(dev)[:, :K] if Kld != K else a.to(dev)
    wp = ops.pack_linear_weight(w.to(dev))
    out = rt.linear(a_d, wp, bias=b.to(dev))
    torch.cuda.synchronize()
    ref = a[:, :K].float() @ bf(w).float().t() + b
    # fp32 accumulate of bf16 products, one bf16 rounding at the output: 2^-8 relative
    assert relerr(out.cpu(), ref) < 6e-3


def test_linear_epilogue(rt):
    dev = _dev()
    from sdb200 import ops
    g = torch.Generator().manual_seed(5)
    B, L, K, N = 3, 256, 288, 288
    a = bf(torch.randn(B * L, K, generator=g))
    w = torch.randn(N, K, generator=g) / K ** 0.5
    b = torch.randn(N, generator=g)
    res = bf(torch.randn(B * L, N, generator=g))
    gate = bf(torch.randn(B, N, generator=g))
    rowb = bf(torch.randn(B, N, generator=g))
    wp = ops.pack_linear_weight(w.to(dev))
    out = rt.linear(a.to(dev), wp, bias=b.to(dev), act="silu", residual=res.to(dev), rows_per_img=L,
                    rowbias=rowb.to(dev), gate=gate.to(dev))
    torch.cuda.synchronize()
    v = a.float() @ bf(w).float().t() + b + rowb.float().repeat_interleave(L, 0)
    ref = F.silu(v) * gate.float().repeat_interleave(L, 0) + res.float()
    assert relerr(out.cpu(), ref) < 6e-3
    out32 = rt.linear(a.to(dev), wp, bias=b.to(dev), act="relu", out_dtype=torch.float32)
    ref32 = F.relu(a.float() @ bf(w).float().t() + b)
    assert relerr(out32.cpu(), ref32) < 1e-4  # fp32 output: only accumulation-order differences remain


CONV_CASES = [  # n, h, w, cin, cout, k
    (2, 32, 32, 64, 128, 3), (3, 16, 16, 128, 192, 3), (5, 8, 8, 192, 256, 3), (9, 4, 4, 256, 64, 3),
    (1, 64, 64, 64, 64, 3), (1, 128, 128, 64, 32, 3), (2, 32, 32, 7, 64, 3), (2, 32, 32, 128, 4, 3),
    (2, 16, 16, 320, 96, 1), (1, 256, 256, 64, 3, 3),
]


@pytest.mark.parametrize("n,h,w,cin,cout,k", CONV_CASES)
def test_conv_same(rt, n, h, w, cin, cout, k):
    dev = _dev()
    from sdb200 import ops
    g = torch.Generator().manual_seed(n * 100 + h + cin + cout)
    x = bf(torch.randn(n, cin, h, w, generator=g))
    wt = torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5
    b = torch.randn(cout, generator=g)
    ld = (cin + 7) // 8 * 8
    xh = torch.zeros(n * h * w, ld, dtype=torch.bfloat16)
    xh[:, :cin] = x.permute(0, 2, 3, 1).reshape(-1, cin)
    out = ops.conv_same(xh.to(dev), n, h, w, cin, ops.pack_conv_weight(wt.to(dev)), cout, k, bias=b.to(dev), ld_x=ld)
    torch.cuda.synchronize()
    ref = F.conv2d(x.float(), bf(wt).float(), b, padding=k // 2).permute(0, 2, 3, 1).reshape(-1, cout)
    assert relerr(out.cpu(), ref) < 6e-3


def test_resnet_second_half_fused(rt):
    """conv3x3(h) + temb row bias, then conv3x3 + fused 1x1 residual conv of the block input (blocks.py:115-119)."""
    dev = _dev()
    from sdb200 import ops
    g = torch.Generator().manual_seed(11)
    n, h, w, cin, cout = 3, 16, 16, 128, 192
    x = bf(torch.randn(n, cin, h, w, generator=g))
    hmid = bf(torch.randn(n, cout, h, w, generator=g))
    w2 = torch.randn(cout, cout, 3, 3, generator=g) / (cout * 9) ** 0.5
    wr = torch.randn(cout, cin, 1, 1, generator=g) / cin ** 0.5
    b2, br = torch.randn(cout, generator=g), torch.randn(cout, generator=g)
    temb = bf(torch.randn(n, cout, generator=g))
    to_nhwc = lambda t: t.permute(0, 2, 3, 1).reshape(-1, t.shape[1]).contiguous()
    wcat = torch.cat([ops.pack_conv_weight(w2.to(dev)), ops.pack_conv_weight(wr.to(dev))], dim=1).contiguous()
    out = ops.conv_same(to_nhwc(hmid).to(dev), n, h, w, cout, wcat, cout, 3, bias=(b2 + br).to(dev),
                        extra=(to_nhwc(x).to(dev), cin, cin), rowbias=temb.to(dev))
    torch.cuda.synchronize()
    ref = F.conv2d(hmid.float(), bf(w2).float(), b2, padding=1) + F.conv2d(x.float(), bf(wr).float(), br)
    ref = ref + temb.float()[:, :, None, None]
    assert relerr(out.cpu(), to_nhwc(ref)) < 6e-3


@pytest.mark.parametrize("n,h,w,c", [(2, 32, 32, 128), (3, 16, 16, 64), (5, 8, 8, 192)])
def test_conv_down_up(rt, n, h, w, c):
    dev = _dev()
    from sdb200 import ops
    g = torch.Generator().manual_seed(h + c)
    x = bf(torch.randn(n, c, h, w, generator=g))
    to_nhwc = lambda t: t.permute(0, 2, 3, 1).reshape(-1, t.shape[1]).contiguous()
    wd = torch.randn(c, c, 4, 4, generator=g) / (c * 16) ** 0.5
    b = torch.randn(c, generator=g)
    out = ops.conv_down4x4(to_nhwc(x).to(dev), n, h, w, c, ops.pack_conv_weight(wd.to(dev)), c, bias=b.to(dev))
    torch.cuda.synchronize()
    ref = F.conv2d(x.float(), bf(wd).float(), b, stride=2, padding=1)
    assert relerr(out.cpu(), to_nhwc(ref)) < 6e-3
    wu = torch.randn(c, c, 4, 4, generator=g) / (c * 4) ** 0.5
    # into a channel slice of a wider buffer, as the up blocks do before the skip concat
    buf = torch.zeros(n * 4 * h * w, 2 * c, dtype=torch.bfloat16, device=dev)
    ops.conv_up4x4(to_nhwc(x).to(dev), n, h, w, c, ops.pack_convT_weight(wu.to(dev)), c, bias=b.to(dev), out=buf)
    torch.cuda.synchronize()
    ref = F.conv_transpose2d(x.float(), bf(wu).float(), b, stride=2, padding=1)
    assert relerr(buf[:, :c].cpu(), to_nhwc(ref)) < 6e-3
    assert buf[:, c:].abs().max().item() == 0.0


@pytest.mark.parametrize("n,hw,c,groups,act", [(2, 1024, 384, 32, "silu"), (3, 256, 512, 32, None),
                                               (4, 16, 768, 32, "silu"), (2, 1024, 128, 32, "relu"),
                                               (1, 65536, 64, 32, "silu"), (2, 64, 48, 16, "silu"),
                                               (2, 1024, 7 * 8, 8, None)])
def test_groupnorm(rt, n, hw, c, groups, act):
    dev = _dev()
    g = torch.Generator().manual_seed(hw + c)
    x = bf(torch.randn(n, hw, c, generator=g) * 2 + 0.7)
    gamma, beta = torch.randn(c, generator=g), torch.randn(c, generator=g)
    out = rt.groupnorm(x.reshape(-1, c).to(dev), n, hw, c, groups, gamma.to(dev), beta.to(dev), act=act)
    torch.cuda.synchronize()
    ref = F.group_norm(x.float().permute(0, 2, 1), groups, gamma, beta, 1e-5).permute(0, 2, 1)
    ref = {"silu": F.silu, "relu": F.relu, None: lambda t: t}[act](ref)
    assert relerr(out.cpu().reshape(n, hw, c), ref) < 6e-3


@pytest.mark.parametrize("b,heads,lq,lk,d", [(2, 16, 1024, 1024, 24), (3, 16, 256, 256, 32), (2, 16, 64, 64, 48),
                                             (5, 16, 16, 16, 32), (2, 16, 1024, 77, 8), (2, 16, 256, 77, 16),
                                             (1, 4, 1024, 1024, 64), (2, 9, 256, 256, 32), (3, 16, 16, 77, 32),
                                             (2, 16, 1024, 1024, 8), (2, 8, 100, 90, 40), (2, 4, 130, 17, 56),
                                             (1, 2, 48, 200, 64), (2, 16, 64, 77, 24), (1, 16, 1000, 1000, 16)])
def test_attention(rt, b, heads, lq, lk, d):
    dev = _dev()
    g = torch.Generator().manual_seed(lq + lk + d)
    C = heads * d
    qkv = bf(torch.randn(b * lq, 3 * C, generator=g))
    if lk == lq:
        q, k, v = qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:]
        qd = qkv.to(dev)
        out = rt.attention(qd[:, :C], qd[:, C:2 * C], qd[:, 2 * C:], b, heads, lq, lk, d)
    else:
        q = qkv[:, :C]
        kv = bf(torch.randn(b * lk, 2 * C, generator=g))
        k, v = kv[:, :C], kv[:, C:]
        kvd = kv.to(dev)
        out = rt.attention(q.contiguous().to(dev), kvd[:, :C], kvd[:, C:], b, heads, lq, lk, d)
    torch.cuda.synchronize()
    sh = lambda t, L: t.float().reshape(b, L, heads, d).transpose(1, 2)
    ref = F.scaled_dot_product_attention(sh(q, lq), sh(k, lk), sh(v, lk)).transpose(1, 2).reshape(b * lq, C)
    # P is rounded to bf16 before the PV product and the output is bf16
    assert relerr(out.cpu(), ref) < 1e-2


@pytest.mark.parametrize("b,heads,lq,lk,d", [(2, 16, 1024, 1024, 24), (3, 16, 256, 256, 32), (2, 16, 64, 64, 48),
                                             (5, 16, 16, 16, 32), (2, 16, 1024, 77, 8), (2, 16, 256, 77, 16),
                                             (1, 4, 1024, 1024, 64), (2, 9, 256, 256, 32), (3, 16, 16, 77, 32),
                                             (2, 16, 1024, 1024, 8), (2, 8, 100, 90, 40), (2, 4, 130, 17, 56),
                                             (1, 2, 48, 200, 64), (2, 16, 64, 77, 24), (1, 16, 1000, 1000, 16),
                                             (1, 3, 129, 65, 8), (2, 2, 128, 64, 64)])
def test_attention_tcgen05_kernel(rt, b, heads, lq, lk, d):
    """The tcgen05 / TMEM kernel on every shape family of the path (ragged tiles, partial key tiles, every head dim),
    output and log-sum-exp against fp32 torch on the same bf16 inputs."""
    dev = _dev()
    g = torch.Generator().manual_seed(7 * lq + lk + d)
    C = heads * d
    q = bf(torch.randn(b * lq, C, generator=g) * 1.5)
    kv = bf(torch.randn(b * lk, 2 * C, generator=g) * 1.5)
    kvd = kv.to(dev)
    out, lse = rt.attention_tc(q.to(dev), kvd[:, :C], kvd[:, C:], b, heads, lq, lk, d, lse=True)
    torch.cuda.synchronize()
    sh = lambda t, L: t.float().reshape(b, L, heads, d).transpose(1, 2)
    qr, kr, vr = sh(q, lq), sh(kv[:, :C], lk), sh(kv[:, C:], lk)
    ref = F.scaled_dot_product_attention(qr, kr, vr).transpose(1, 2).reshape(b * lq, C)
    assert relerr(out.cpu(), ref) < 1e-2
    ref_lse = torch.logsumexp((qr @ kr.transpose(-1, -2)) * d ** -0.5, -1) * 1.4426950408889634
    assert torch.allclose(lse.cpu(), ref_lse, atol=2e-3, rtol=1e-4)


def test_attention_unsupported_head_dim(rt):
    dev = _dev()
    x = torch.zeros(16, 3 * 16 * 3, dtype=torch.bfloat16, device=dev)
    with pytest.raises(rt.B200Error):
        rt.attention(x, x, x, 1, 16, 16, 16, 3)


def test_layout_and_mask(rt):
    dev = _dev()
    g = torch.Generator().manual_seed(3)
    n, h, w = 3, 32, 32
    x = torch.randn(n, 4, h, w, generator=g)
    mask = (torch.rand(n, 18, 512, 512, generator=g) > 0.5).float()
    wc = torch.randn(3, 18, generator=g)
    buf = torch.zeros(n * h * w, 8, dtype=torch.bfloat16, device=dev)
    rt.nchw_to_nhwc(x.to(dev), buf, 0)
    rt.mask_cond(mask.to(dev), wc.to(dev), buf, 4, h, w)
    torch.cuda.synchronize()
    ref_m = F.conv2d(F.interpolate(mask, size=(h, w)), bf(wc).float()[:, :, None, None])
    ref = torch.cat([x, ref_m], 1).permute(0, 2, 3, 1).reshape(-1, 7)
    assert relerr(buf[:, :7].cpu(), ref) < 5e-3
    assert buf[:, 7].abs().max().item() == 0
    back = rt.nhwc_to_nchw(buf, n, 7, h, w, ld=8)
    assert torch.equal(back.cpu(), buf[:, :7].float().cpu().reshape(n, h, w, 7).permute(0, 3, 1, 2))


def test_time_embedding(rt):
    dev = _dev()
    t = torch.tensor([0, 1, 17, 500, 999], dtype=torch.int64)
    out = rt.time_embedding(t.to(dev), 5, 512, dtype=torch.float32)
    half = 256
    factor = 10000 ** (torch.arange(0, half, dtype=torch.float32) / half)
    e = t[:, None].repeat(1, half) / factor
    ref = torch.cat([torch.sin(e), torch.cos(e)], -1)
    assert (out.cpu() - ref).abs().max().item() < 2e-4  # fp32 sin/cos of arguments up to 999 (ulp of arg ~6e-5)
    out1 = rt.time_embedding(t[2:3].to(dev), 4, 512, dtype=torch.float32)  # (1,) t broadcast over the batch
    assert torch.equal(out1.cpu(), out.cpu()[2:3].expand(4, -1))


# ------------------------------------------------------------------------- fused GroupNorm statistics (igemm v2)
STAT_CASES = [  # n, h, w, cin, cout, k, groups
    (3, 32, 32, 64, 128, 3, 32), (5, 8, 8, 128, 192, 3, 16), (9, 4, 4, 64, 256, 3, 32), (2, 16, 16, 96, 64, 1, 16),
    (2, 24, 20, 64, 96, 3, 8), (70, 4, 4, 64, 512, 1, 32),
    # CTA-pair kernels (>= 16 K-chunks): 192-column tile; an odd number of row tiles (the pair's second tile lies
    # outside the tensor); more work items than clusters; several images per tile with an odd tile count
    (3, 16, 16, 192, 384, 3, 32), (5, 8, 8, 192, 384, 3, 32), (24, 32, 32, 128, 256, 3, 32),
    (17, 4, 4, 256, 512, 3, 32), (5, 16, 16, 128, 128, 3, 32),
]


@pytest.mark.parametrize("n,h,w,cin,cout,k,groups", STAT_CASES)
def test_conv_stats_then_groupnorm_apply(rt, n, h, w, cin, cout, k, groups):
    """The igemm epilogue's partial statistics + groupnorm_apply reproduce F.group_norm of the conv output."""
    dev = _dev()
    from sdb200 import ops
    g = torch.Generator().manual_seed(n + h + cin + cout)
    x = bf(torch.randn(n, cin, h, w, generator=g))
    wt = torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5
    b = torch.randn(cout, generator=g)
    gamma, beta = torch.randn(cout, generator=g), torch.randn(cout, generator=g)
    xd = x.permute(0, 2, 3, 1).reshape(n * h * w, cin).contiguous().to(dev)
    P = rt.stats_slots(h, w)
    stats = torch.full((n, P, cout, 2), float("nan"), device=dev)  # every element must be overwritten
    y = ops.conv_same(xd, n, h, w, cin, ops.pack_conv_weight(wt.to(dev)), cout, k, bias=b.to(dev),
                      stats=(stats, P, 0, cout))
    out = rt.groupnorm_apply(y, n, h * w, cout, groups, gamma.to(dev), beta.to(dev), [(stats, P, cout, cout)],
                             act="silu")
    torch.cuda.synchronize()
    assert torch.isfinite(stats).all()
    yref = y.float().cpu().reshape(n, h, w, cout).permute(0, 3, 1, 2)  # GroupNorm of the bf16 tensor we produced
    ref = F.silu(F.group_norm(yref, groups, gamma, beta, 1e-5))
    got = out.float().cpu().reshape(n, h, w, cout).permute(0, 3, 1, 2)
    assert relerr(got, ref) < 6e-3  # one bf16 rounding at the output
    # statistics themselves: sum over slots equals the per-(image, channel) sums of the bf16 output
    s = stats.sum(1).cpu()
    assert torch.allclose(s[..., 0], yref.sum((2, 3)), rtol=1e-3, atol=1e-2)
    assert torch.allclose(s[..., 1], (yref ** 2).sum((2, 3)), rtol=1e-3, atol=1e-2)
    # bit-reproducible: no atomics anywhere
    stats2 = torch.empty_like(stats)
    y2 = ops.conv_same(xd, n, h, w, cin, ops.pack_conv_weight(wt.to(dev)), cout, k, bias=b.to(dev),
                       stats=(stats2, P, 0, cout))
    assert torch.equal(stats, stats2) and torch.equal(y, y2)


def test_groupnorm_apply_two_segments(rt):
    """Channel concat of two producers (up-block skip connection): statistics come from two buffers."""
    dev = _dev()
    g = torch.Generator().manual_seed(4)
    n, hw, c0, c1, groups = 3, 256, 64, 128, 16
    x = bf(torch.randn(n, hw, c0 + c1, generator=g) * 2 + 0.5)
    xf = x.float()

    def mk(part, P):
        st = torch.zeros(n, P, part.shape[-1], 2)
        chunks = part.chunk(P, dim=1)
        for i, ch in enumerate(chunks):
            st[:, i, :, 0], st[:, i, :, 1] = ch.sum(1), (ch ** 2).sum(1)
        return st.to(dev)

    s0, s1 = mk(xf[..., :c0], 2), mk(xf[..., c0:], 4)
    gamma, beta = torch.randn(c0 + c1, generator=g), torch.randn(c0 + c1, generator=g)
    out = rt.groupnorm_apply(x.reshape(n * hw, -1).to(dev), n, hw, c0 + c1, groups, gamma.to(dev), beta.to(dev),
                             [(s0, 2, c0, c0), (s1, 4, c1, c1)], act=None)
    ref = F.group_norm(xf.permute(0, 2, 1), groups, gamma, beta, 1e-5).permute(0, 2, 1)
    assert relerr(out.float().cpu().reshape(n, hw, -1), ref) < 6e-3


def test_linear_residual_tma_epilogue_large(rt):
    """The attention out-projection shape (K = N = 384, residual) at many tiles per CTA: exercises the persistent
    loop, both TMEM accumulators, the residual prefetch and the staged TMA store."""
    dev = _dev()
    from sdb200 import ops
    g = torch.Generator().manual_seed(8)
    M, K, N = 40000, 384, 384
    a, res = bf(torch.randn(M, K, generator=g)), bf(torch.randn(M, N, generator=g))
    w, b = torch.randn(N, K, generator=g) / K ** 0.5, torch.randn(N, generator=g)
    out = rt.linear(a.to(dev), ops.pack_linear_weight(w.to(dev)), bias=b.to(dev), residual=res.to(dev))
    ref = a.float() @ bf(w).float().t() + b + res.float()
    assert relerr(out.cpu(), ref) < 6e-3


# ------------------------------------------------------------------------------------------- weight gradient
@pytest.mark.parametrize("n,h,w,cin,cout,k", [(2, 32, 32, 64, 128, 3), (4, 16, 16, 128, 192, 3), (5, 8, 8, 192, 256, 3),
                                              (9, 4, 4, 256, 64, 3), (2, 32, 32, 7, 64, 3), (3, 16, 16, 320, 96, 1),
                                              (2, 24, 20, 72, 40, 3), (16, 32, 32, 128, 128, 3)])
def test_wgrad_conv(rt, n, h, w, cin, cout, k):
    """MN-major tcgen05 wgrad vs torch autograd of F.conv2d on the same bf16-rounded operands."""
    dev = _dev()
    from sdb200 import ops
    g = torch.Generator().manual_seed(n + h + cin + cout)
    x = bf(torch.randn(n, cin, h, w, generator=g))
    dy = bf(torch.randn(n, cout, h, w, generator=g))
    wt = torch.zeros(cout, cin, k, k, requires_grad=True)
    F.conv2d(x.float(), wt, padding=k // 2).backward(dy.float())
    ldx, ldy = (cin + 7) // 8 * 8, (cout + 7) // 8 * 8
    xd = torch.zeros(n * h * w, ldx, dtype=torch.bfloat16)
    xd[:, :cin] = x.permute(0, 2, 3, 1).reshape(-1, cin)
    dyd = torch.zeros(n * h * w, ldy, dtype=torch.bfloat16)
    dyd[:, :cout] = dy.permute(0, 2, 3, 1).reshape(-1, cout)
    xd, dyd = xd.to(dev), dyd.to(dev)
    pad = k // 2
    cp = (cin + 63) // 64 * 64
    srcs = [rt.make_src(xd, n, h, w, cin, ld=ldx)]
    segs = [(0, ky - pad, kx - pad, 0, cp // 64) for ky in range(k) for kx in range(k)]
    dwp = rt.wgrad(srcs, segs, rt.make_src(dyd, n, h, w, cout, ld=ldy), cout, n, h, w)
    torch.cuda.synchronize()
    got = dwp.cpu().reshape(cout, k * k, cp)[:, :, :cin].permute(0, 2, 1).reshape(cout, cin, k, k)
    assert torch.all(dwp.cpu().reshape(cout, k * k, cp)[:, :, cin:] == 0)  # padded columns read zero-filled channels
    assert relerr(got, wt.grad) < 2e-3  # bf16 products, fp32 accumulation over n*h*w pixels, fp32 output
    dwp2 = rt.wgrad(srcs, segs, rt.make_src(dyd, n, h, w, cout, ld=ldy), cout, n, h, w)
    assert torch.equal(dwp, dwp2)  # split-K partials are reduced in fixed order


# ------------------------------------------------------------------------------------- GroupNorm backward, colsum
@pytest.mark.parametrize("two_pass", [False, True])
@pytest.mark.parametrize("n,hw,c,groups,act", [(3, 1024, 128, 32, "silu"), (5, 64, 384, 32, "silu"),
                                               (4, 16, 512, 32, "relu"), (2, 256, 96, 8, None), (16, 1024, 256, 32, "silu"),
                                               (2, 1024, 320, 32, "silu"),     # 10 channels per group: 8 groups per cluster
                                               (2, 1024, 960, 32, "silu"),     # 30 per group: 120-channel slabs, 8 CTAs
                                               (3, 100, 80, 8, "relu"),        # rows not a multiple of the row step
                                               (2, 64, 48, 8, "silu"),         # 6 per group: all 8 groups in one cluster
                                               (1, 64, 24, 4, None),           # whole layer narrower than 64 channels
                                               (2, 4096, 1024, 32, "silu"),    # slab with the extra gradient too large: two-pass
                                               (1, 16384, 256, 32, "silu")])   # slab too large: two-pass either way
def test_groupnorm_backward(rt, n, hw, c, groups, act, two_pass, monkeypatch):
    """One-pass cluster kernel (slab of x / dy in shared memory) and the two-pass route (B200SD_GN_BWD_FUSED=0, also what
    shapes whose slab does not fit use) against autograd through F.group_norm."""
    if two_pass:
        monkeypatch.setenv("B200SD_GN_BWD_FUSED", "0")
    dev = _dev()
    g = torch.Generator().manual_seed(n + hw + c)
    x = bf(torch.randn(n, hw, c, generator=g) * 1.5 + 0.3)
    dy = bf(torch.randn(n, hw, c, generator=g))
    extra = bf(torch.randn(n, hw, c, generator=g))
    gamma = (torch.randn(c, generator=g) * 0.5 + 1).requires_grad_()
    beta = torch.randn(c, generator=g).requires_grad_()
    xr = x.float().requires_grad_()
    y = F.group_norm(xr.permute(0, 2, 1), groups, gamma, beta, 1e-5).permute(0, 2, 1)
    y = {"silu": F.silu, "relu": F.relu, None: lambda t: t}[act](y)
    y.backward(dy.float())
    # forward statistics in the layout the igemm epilogue writes: 2 partial slots per image
    st = torch.zeros(n, 2, c, 2)
    for i, part in enumerate(x.float().chunk(2, dim=1)):
        st[:, i, :, 0], st[:, i, :, 1] = part.sum(1), (part ** 2).sum(1)
    dx, dgamma, dbeta = rt.groupnorm_bwd(x.reshape(n * hw, c).to(dev), dy.reshape(n * hw, c).to(dev), n, hw, c, groups,
                                         gamma.detach().to(dev), beta.detach().to(dev), [(st.to(dev), 2, c, c)], act=act,
                                         add=extra.reshape(n * hw, c).to(dev))
    torch.cuda.synchronize()
    ref_dx = xr.grad + extra.float()
    assert relerr(dx.float().cpu().reshape(n, hw, c), ref_dx) < 6e-3   # bf16 output
    assert relerr(dgamma.cpu(), gamma.grad) < 2e-3
    assert relerr(dbeta.cpu(), beta.grad) < 2e-3


def test_colsum(rt):
    dev = _dev()
    g = torch.Generator().manual_seed(2)
    n, hw, c = 7, 300, 192
    x = bf(torch.randn(n, hw, c, generator=g))
    xd = x.reshape(n * hw, c).to(dev)
    per = rt.colsum(xd, n, hw, c, per_image=True)
    tot = rt.colsum(xd, n, hw, c)
    assert torch.allclose(per.cpu(), x.float().sum(1), rtol=1e-4, atol=1e-3)
    assert torch.allclose(tot.cpu(), x.float().sum((0, 1)), rtol=1e-4, atol=1e-3)


# --------------------------------------------------------------------------------------------- attention backward
@pytest.mark.parametrize("b,heads,lq,lk,d", [(2, 16, 1024, 1024, 24), (2, 16, 1024, 1024, 8), (3, 16, 256, 256, 32),
                                             (2, 16, 64, 64, 48), (5, 16, 16, 16, 32), (2, 16, 1024, 77, 8),
                                             (2, 16, 256, 77, 16), (2, 9, 256, 256, 32), (2, 4, 100, 90, 40)])
def test_attention_backward(rt, b, heads, lq, lk, d):
    dev = _dev()
    g = torch.Generator().manual_seed(lq + lk + d)
    C = heads * d
    q = bf(torch.randn(b * lq, C, generator=g))
    kv = bf(torch.randn(b * lk, 2 * C, generator=g))
    d_o = bf(torch.randn(b * lq, C, generator=g))
    sh = lambda t, L: t.float().reshape(b, L, heads, d).transpose(1, 2)
    qr, kr, vr = (t.clone().requires_grad_() for t in (sh(q, lq), sh(kv[:, :C], lk), sh(kv[:, C:], lk)))
    ref = F.scaled_dot_product_attention(qr, kr, vr)
    ref.backward(sh(d_o, lq))
    un = lambda t, L: t.transpose(1, 2).reshape(b * L, C)
    qd, kvd, dod = q.to(dev), kv.to(dev), d_o.to(dev)
    out, lse = rt.attention_lse(qd, kvd[:, :C], kvd[:, C:], b, heads, lq, lk, d)
    assert relerr(out.cpu(), un(ref.detach(), lq)) < 1e-2
    ref_lse = torch.logsumexp((qr.detach() @ kr.detach().transpose(-1, -2)) * d ** -0.5, -1) * 1.4426950408889634
    assert torch.allclose(lse.cpu(), ref_lse, atol=2e-3, rtol=1e-4)
    dq = torch.empty_like(qd)
    dkv = torch.empty_like(kvd)
    rt.attention_bwd(qd, kvd[:, :C], kvd[:, C:], out, dod, lse, b, heads, lq, lk, d, dq, dkv[:, :C], dkv[:, C:])
    torch.cuda.synchronize()
    # P and dS are rounded to bf16 before the second GEMMs; outputs are bf16
    assert relerr(dq.cpu(), un(qr.grad, lq)) < 1.5e-2
    assert relerr(dkv[:, :C].cpu(), un(kr.grad, lk)) < 1.5e-2
    assert relerr(dkv[:, C:].cpu(), un(vr.grad, lk)) < 1.5e-2


@pytest.mark.parametrize("n,h,w,cin,cout,k", [(16, 4, 4, 768, 768, 3), (16, 8, 8, 512, 384, 3), (4, 8, 8, 384, 200, 3),
                                              (2, 4, 4, 2048, 96, 1)])
def test_split_k_matches_the_single_pass_gemm(rt, n, h, w, cin, cout, k):
    """Split-K (training passes only: the K loop of a tile cut into ranges run by different CTAs, fp32 partials parked in a
    workspace, the last CTA to arrive sums them in split order and runs the full epilogue — csrc/igemm.cu) against the
    same launch without it: same epilogue (bias, time-embedding row bias, SiLU, residual, fused GroupNorm statistics),
    results within fp32 summation-order noise of each other and of the fp32 reference; run twice in a row (the tile
    counters are returned to zero) and bit-identical between the two runs (fixed summation order)."""
    dev = _dev()
    from sdb200 import ops
    g = torch.Generator().manual_seed(n + h + cin + cout)
    x = bf(torch.randn(n, cin, h, w, generator=g))
    wt = torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5
    b = torch.randn(cout, generator=g)
    temb = bf(torch.randn(n, cout, generator=g))
    res = bf(torch.randn(n * h * w, cout, generator=g))
    xh = x.permute(0, 2, 3, 1).reshape(-1, cin).contiguous().to(dev)
    wp = ops.pack_conv_weight(wt.to(dev))
    want_stats = cout % 8 == 0 and cout > 32

    def run():
        st = None
        if want_stats:
            P = rt.stats_slots(h, w)
            st = (torch.full((n, P, cout, 2), float("nan"), device=dev), P, 0, cout)
        out = ops.conv_same(xh, n, h, w, cin, wp, cout, k, bias=b.to(dev), rowbias=temb.to(dev), residual=res.to(dev),
                            act="silu", stats=st)
        return out, (st[0] if st else None), int(rt.lib().b200sd_igemm_last_splits())

    with torch.no_grad():
        plain, pst, s_plain = run()
    keep = rt.SPLITK_MAX
    rt.SPLITK_MAX = 8          # opt-in (off by default: it pays for a handful of launches only, profiles/r2i_splitk.md)
    try:
        with torch.enable_grad():
            split, sst, s_split = run()
            split2, sst2, _ = run()
    finally:
        rt.SPLITK_MAX = keep
    torch.cuda.synchronize()
    assert s_plain == 1 and s_split >= 2, (s_plain, s_split)
    ref = F.silu(F.conv2d(x.float(), bf(wt).float(), b, padding=k // 2) + temb.float()[:, :, None, None])
    ref = ref.permute(0, 2, 3, 1).reshape(-1, cout) + res.float()
    assert relerr(split.cpu(), ref) < 6e-3 and relerr(plain.cpu(), ref) < 6e-3
    assert relerr(split, plain) < 4e-3                     # one bf16 ulp where the fp32 sums round differently
    assert torch.equal(split, split2)
    if want_stats:
        assert torch.isfinite(sst).all() and torch.equal(sst, sst2)
        assert relerr(sst.sum(1), pst.sum(1)) < 2e-3

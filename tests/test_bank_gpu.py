"""Weight bank of the captured training step: the table-driven pack kernel must produce exactly the matrices the
per-leaf torch packers produce, and the scatter form of the weight-gradient reduction must leave exactly the values the
packed reduction + permute leaves (bit-exact: same bf16 rounding, same split order)."""
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu

rt = importlib.import_module("stablediffusion-pytorch_b200.runtime")
ops = importlib.import_module("stablediffusion-pytorch_b200.ops")
BK = importlib.import_module("stablediffusion-pytorch_b200.bank")
AG = importlib.import_module("stablediffusion-pytorch_b200.autograd")


def _bank(params):
    slots, off = [], 0
    for p in params:
        slots.append((off, p.numel(), tuple(p.shape)))
        off += (p.numel() + 63) // 64 * 64
    flat = torch.full((off,), float("nan"), device="cuda")
    return BK.WeightBank(params, flat, slots)


def test_pack_forms_match_torch_packers():
    g = torch.Generator(device="cuda").manual_seed(0)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    lin, conv3, conv1, sc, down, up = r(200, 72), r(96, 40, 3, 3), r(130, 70, 1, 1), r(96, 24, 1, 1), r(64, 48, 4, 4), \
        r(48, 80, 4, 4)
    params = [lin, conv3, conv1, sc, down, up]
    b = _bank(params)
    got = {
        "lin": b.linear(lin), "linT": b.linear_t(lin), "conv3": b.conv(conv3), "conv3+sc": b.conv(conv3, sc),
        "conv1": b.conv(conv1), "conv3D": b.conv_dgrad(conv3), "conv1T": b.linear_t(conv1), "down": b.conv(down),
        "downT": b.convT_phases(down, AG._phase_taps(64)), "up": b.convT_phases(up, AG._phase_taps(64)),
        "upD": b.conv(up),
    }

    def check():
        want = {
            "lin": ops.pack_linear_weight(lin), "linT": ops.pack_linear_weight(lin.t()),
            "conv3": ops.pack_conv_weight(conv3),
            "conv3+sc": torch.cat([ops.pack_conv_weight(conv3), ops.pack_conv_weight(sc)], 1),
            "conv1": ops.pack_conv_weight(conv1), "conv3D": ops.pack_conv_weight(conv3.flip(2, 3).transpose(0, 1)),
            "conv1T": ops.pack_linear_weight(conv1.reshape(130, 70).t()), "down": ops.pack_conv_weight(down),
            "downT": ops.pack_convT_weight(down), "up": ops.pack_convT_weight(up), "upD": ops.pack_conv_weight(up),
        }
        for k, w in want.items():
            if isinstance(w, list):
                for (gm, gs), (wm, ws) in zip(got[k], w):
                    assert list(gs) == list(ws), k
                    assert torch.equal(gm, wm), k
            else:
                assert got[k].shape == w.shape and torch.equal(got[k], w), k

    check()
    # one launch refreshes every form after the parameters changed
    b.finalize()
    assert b.n_jobs == 1 + 1 + 1 + 2 + 1 + 1 + 1 + 1 + 4 + 4 + 1
    for p in params:
        p.mul_(-1.7).add_(0.01)
    for v in got.values():
        for m in ([x[0] for x in v] if isinstance(v, list) else [v]):
            m.fill_(7.0)
    before = rt.launch_count()
    b.refresh()
    assert rt.launch_count() - before <= 1
    check()


def test_bank_refuses_new_forms_during_capture():
    w = torch.randn(64, 64, device="cuda")
    b = _bank([w])
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with pytest.raises(rt.B200Error):
        with torch.cuda.stream(s):
            with torch.cuda.graph(g):
                b.linear(w)
    torch.cuda.synchronize()


@pytest.mark.parametrize("leaf", ["linear", "conv3", "conv3_shortcut", "down", "up"])
def test_gradients_written_in_place_match_packed_path(leaf):
    torch.manual_seed(1)
    dev = "cuda"
    n, h, w_ = 2, 16, 16
    bf = lambda *s: torch.randn(*s, device=dev).to(torch.bfloat16)

    def run(use_bank):
        torch.manual_seed(2)
        if leaf == "linear":
            ws = [torch.randn(150, 72, device=dev, requires_grad=True)]
            x = bf(n * h * w_, 72).requires_grad_()
            f = lambda: AG.LinearFn.apply(x, ws[0], None, None, 0, False)
        elif leaf in ("conv3", "conv3_shortcut"):
            ws = [torch.randn(96, 40, 3, 3, device=dev, requires_grad=True)]
            x = bf(n * h * w_, 40).requires_grad_()
            if leaf == "conv3":
                f = lambda: AG.ConvSameFn.apply(x, ws[0], None, None, None, None, None, (n, h, w_, 40, 0), False, False)
            else:
                ws.append(torch.randn(96, 24, 1, 1, device=dev, requires_grad=True))
                x2 = bf(n * h * w_, 24).requires_grad_()
                f = lambda: AG.ConvSameFn.apply(x, ws[0], None, None, x2, ws[1], None, (n, h, w_, 40, 24), False, False)
        elif leaf == "down":
            ws = [torch.randn(64, 48, 4, 4, device=dev, requires_grad=True)]
            x = bf(n * h * w_, 48).requires_grad_()
            f = lambda: AG.ConvDownFn.apply(x, ws[0], None, (n, h, w_, 48), False)
        else:
            ws = [torch.randn(48, 80, 4, 4, device=dev, requires_grad=True)]
            x = bf(n * h * w_, 48).requires_grad_()
            f = lambda: AG.ConvUpFn.apply(x, ws[0], None, (n, h, w_, 48), False)
        b = _bank([p.detach() for p in ws]) if use_bank else None
        BK.ACTIVE = b
        try:
            y = f()
            y = y[0] if isinstance(y, tuple) else y
            gy = torch.randn(y.shape, device=dev, generator=torch.Generator(device=dev).manual_seed(3)).to(y.dtype)
            grads = torch.autograd.grad((y,), [x] + ws, (gy,))
        finally:
            BK.ACTIVE = None
        if use_bank:
            for p, g in zip(ws, grads[1:]):
                assert g.data_ptr() == b.grad_view(p.detach()).data_ptr()
        return [y.detach().float()] + [g.detach().float().clone() for g in grads]

    a, c = run(False), run(True)
    for u, v in zip(a, c):
        assert torch.equal(u, v)


@pytest.mark.parametrize("use_bank", [False, True])
def test_cross_projection_matches_sliced_linears(use_bank):
    dev, C, Mq, Mk = "cuda", 128, 512, 154
    torch.manual_seed(4)
    w = torch.randn(3 * C, C, device=dev, requires_grad=True)
    b = torch.randn(3 * C, device=dev, requires_grad=True)
    xq = torch.randn(Mq, C, device=dev).to(torch.bfloat16).requires_grad_()
    xkv = torch.randn(Mk, C, device=dev).to(torch.bfloat16).requires_grad_()
    gq = torch.randn(Mq, C, device=dev).to(torch.bfloat16)
    gkv = torch.randn(Mk, 2 * C, device=dev).to(torch.bfloat16)

    q0, _ = AG.LinearFn.apply(xq, w[:C], b[:C], None, None, False)
    kv0, _ = AG.LinearFn.apply(xkv, w[C:], b[C:], None, None, False)
    want = torch.autograd.grad((q0, kv0), [xq, xkv, w, b], (gq, gkv))

    BK.ACTIVE = _bank([w.detach(), b.detach()]) if use_bank else None
    try:
        q1, kv1 = AG.CrossProjFn.apply(xq, xkv, w, b, C)
        got = torch.autograd.grad((q1, kv1), [xq, xkv, w, b], (gq, gkv))
        if use_bank:
            assert got[2].data_ptr() == BK.ACTIVE.grad_view(w.detach()).data_ptr()
    finally:
        BK.ACTIVE = None
    assert torch.equal(q0, q1) and torch.equal(kv0, kv1)
    for u, v in zip(want, got):
        assert torch.equal(u.float(), v.float())

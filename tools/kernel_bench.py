"""Per-kernel microbenchmarks at the shapes the conditioned UNet hits at batch 64 (SURVEY §2b shape census).

    python tools/kernel_bench.py [--only igemm|groupnorm|groupnorm_bwd|attention|sched] [--iters N] [--json out.json]

Each case: 3 warm-up launches, then N timed launches bracketed by CUDA events on the launching stream, with a 256 MB
write between launches to flush the 126 MB L2 (the flush is outside the event pair). Prints achieved TFLOP/s (tensor-
bound kernels) or GB/s (HBM-bound kernels) and the fraction of MEASURED_PEAKS.json (burst figures: the kernel is
timed alone). This is the script the `ncu --set full` captures under profiles/ are taken on.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import sdb200  # noqa: E402,F401
from sdb200 import ops  # noqa: E402
from sdb200 import runtime as rt  # noqa: E402


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"]
    return 6650.0, 1590.0


class Timer:
    def __init__(self, iters):
        self.iters = iters
        self.flush = torch.empty(64 * 1024 * 1024, device="cuda", dtype=torch.float32)

    def run(self, fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        tot = 0.0
        for _ in range(self.iters):
            self.flush.fill_(1.0)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            b.synchronize()
            tot += a.elapsed_time(b)
        return tot / self.iters * 1e3  # us


def igemm_cases():
    B = 64
    cases = []
    # (name, n, h, w, cin, cout, k, residual)
    for name, h, cin, cout, k, res in [
        ("conv3x3 384->384 @32", 32, 384, 384, 3, False),
        ("conv3x3 512->512 @16", 16, 512, 512, 3, False),
        ("conv3x3 768->768 @8", 8, 768, 768, 3, False),
        ("conv3x3 512->512 @4", 4, 512, 512, 3, False),
        ("conv3x3 128->128 @32", 32, 128, 128, 3, False),
        ("conv3x3 256->384 @32", 32, 256, 384, 3, False),
        ("linear 384->1152 @32 (qkv)", 32, 384, 1152, 1, False),
        ("linear 384->384 @32 +res (out_proj)", 32, 384, 384, 1, True),
        ("linear 512->512 @16 +res", 16, 512, 512, 1, True),
        ("linear 128->128 @32 +res", 32, 128, 128, 1, True),
        ("linear 768->2304 @8 (qkv)", 8, 768, 2304, 1, False),
    ]:
        cases.append((name, B, h, h, cin, cout, k, res))
    return cases


def bench_igemm(tm, out):
    hbm, tf = peaks()
    for name, n, h, w, cin, cout, k, res in igemm_cases():
        x = torch.randn(n * h * w, cin, device="cuda").to(torch.bfloat16)
        wt = torch.randn(cout, cin, k, k, device="cuda") * 0.05
        wmat = ops.pack_conv_weight(wt)
        bias = torch.randn(cout, device="cuda")
        r = torch.randn(n * h * w, cout, device="cuda").to(torch.bfloat16) if res else None
        y = torch.empty(n * h * w, cout, device="cuda", dtype=torch.bfloat16)
        fn = lambda: ops.conv_same(x, n, h, w, cin, wmat, cout, k, bias=bias, out=y, residual=r)
        us = tm.run(fn)
        flops = 2.0 * n * h * w * cout * cin * k * k
        byts = 2.0 * (x.numel() + y.numel() + wmat.numel() + (r.numel() if res else 0))
        t = flops / us / 1e6
        print(f"igemm     {name:38s} {us:9.1f} us  {t:7.1f} TF/s ({t / tf:5.1%} of burst)  {byts / us / 1e3:7.0f} GB/s")
        out.append({"kernel": "igemm", "case": name, "us": us, "tflops": t, "frac": t / tf, "gbs": byts / us / 1e3})


def bench_groupnorm(tm, out):
    """gn_apply (the GroupNorm the models run: statistics come from the producer's epilogue) at the batch-64 UNet
    shapes, plus two batch-256 cases whose 400-800 MB of traffic cannot sit in the 126 MB L2."""
    hbm, tf = peaks()
    cases = [(64, 1024, 384), (64, 1024, 128), (64, 256, 512), (64, 256, 256), (64, 64, 768), (64, 64, 384),
             (64, 16, 512), (64, 1024, 512), (256, 1024, 384), (256, 1024, 512)]
    for n, hw, c in cases:
        x = torch.randn(n * hw, c, device="cuda").to(torch.bfloat16)
        g, b = torch.randn(c, device="cuda"), torch.randn(c, device="cuda")
        xf = x.float().view(n, hw, c)
        stats = torch.stack([xf.sum(1), (xf * xf).sum(1)], dim=-1).view(n, 1, c, 2).contiguous()
        del xf
        y = torch.empty_like(x)
        us = tm.run(lambda: rt.groupnorm_apply(x, n, hw, c, 32, g, b, [(stats, 1, c, c)], act="silu", out=y))
        gbs = 4.0 * n * hw * c / us / 1e3
        print(f"gn_apply  n={n:4d} hw={hw:5d} c={c:4d} {us:9.1f} us  {gbs:7.0f} GB/s ({gbs / hbm:5.1%} of measured copy)")
        out.append({"kernel": "gn_apply", "case": f"{n}x{hw}x{c}", "us": us, "gbs": gbs, "frac": gbs / hbm})


def bench_groupnorm_bwd(tm, out):
    """GroupNorm (+ SiLU) backward at the batch-16 training shapes of the conditioned UNet: the one-pass cluster kernel (the
    slab of x and dy stays in shared memory between the reduction and the apply phase) against the two-pass route
    (B200SD_GN_BWD_FUSED=0). Eight calls on eight different input sets inside one CUDA graph (how the training step
    runs them; L2 flushed before each replay, so every call reads cold inputs). Algorithmic bytes: x, dy, the extra
    gradient in; dx out = 8 bytes per element."""
    hbm, tf = peaks()
    cases = [(16, 1024, 128), (16, 1024, 256), (16, 1024, 384), (16, 1024, 512), (16, 1024, 640), (16, 256, 256),
             (16, 256, 512), (16, 256, 768), (16, 256, 1024), (16, 64, 768), (16, 64, 1536), (16, 16, 768), (64, 1024, 384)]
    reps = 8
    for n, hw, c in cases:
        sets = []
        for _ in range(reps):
            x = torch.randn(n * hw, c, device="cuda").to(torch.bfloat16)
            dy = torch.randn(n * hw, c, device="cuda").to(torch.bfloat16)
            extra = torch.randn(n * hw, c, device="cuda").to(torch.bfloat16)
            xf = x.float().view(n, hw, c)
            stats = torch.stack([xf.sum(1), (xf * xf).sum(1)], dim=-1).view(n, 1, c, 2).contiguous()
            del xf
            sets.append((x, dy, extra, stats))
        g, b = torch.randn(c, device="cuda"), torch.randn(c, device="cuda")
        res = {}
        for mode in ("1", "0"):
            os.environ["B200SD_GN_BWD_FUSED"] = mode

            def calls():
                for x, dy, extra, stats in sets:
                    rt.groupnorm_bwd(x, dy, n, hw, c, 32, g, b, [(stats, 1, c, c)], act="silu", add=extra)
            calls()
            torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                calls()
            res[mode] = tm.run(gr.replay) / reps
            del gr
        os.environ.pop("B200SD_GN_BWD_FUSED", None)
        gbs = 8.0 * n * hw * c / res["1"] / 1e3
        print(f"gn_bwd    n={n:4d} hw={hw:5d} c={c:4d} one-pass {res['1']:8.1f} us  two-pass {res['0']:8.1f} us  "
              f"{gbs:7.0f} GB/s ({gbs / hbm:5.1%} of measured copy)", flush=True)
        out.append({"kernel": "gn_bwd", "case": f"{n}x{hw}x{c}", "us": res["1"], "us_two_pass": res["0"], "gbs": gbs,
                    "frac": gbs / hbm})


def bench_attention(tm, out):
    hbm, tf = peaks()
    for lq, lk, c, heads in [(1024, 1024, 384, 16), (1024, 1024, 128, 16), (256, 256, 512, 16), (256, 256, 256, 16),
                             (64, 64, 768, 16), (1024, 77, 384, 16), (1024, 77, 128, 16), (256, 77, 512, 16),
                             (256, 77, 256, 16), (16, 16, 512, 16), (1024, 1024, 256, 4), (256, 256, 288, 9),
                             (256, 77, 288, 9)]:
        n = 64
        d = c // heads
        q = torch.randn(n * lq, c, device="cuda").to(torch.bfloat16)
        kv = torch.randn(n * lk, 2 * c, device="cuda").to(torch.bfloat16)
        o = torch.empty_like(q)
        us = tm.run(lambda: rt.attention(q, kv[:, :c], kv[:, c:], n, heads, lq, lk, d, out=o))
        t = 4.0 * n * heads * lq * lk * d / us / 1e6
        print(f"attention lq={lq:5d} lk={lk:5d} d={d:3d} {us:9.1f} us  {t:7.1f} TF/s  "
              f"{n * heads * lq * lk / us / 1e3:7.1f} Gscore/s")
        out.append({"kernel": "attention", "case": f"{lq}x{lk}x{d}", "us": us, "tflops": t})


def bench_sched(tm, out):
    hbm, tf = peaks()
    from scheduler.linear_noise_scheduler import LinearNoiseScheduler
    s = LinearNoiseScheduler(1000, 0.00085, 0.012)
    for n in (64, 65536):
        x = torch.randn(n, 4, 32, 32, device="cuda")
        e = torch.randn_like(x)
        t = torch.randint(0, 1000, (n,), device="cuda")
        us = tm.run(lambda: s.add_noise(x, e, t))
        gbs = 3 * 4.0 * x.numel() / us / 1e3
        print(f"add_noise B={n:6d} {us:9.1f} us {gbs:7.0f} GB/s ({gbs / hbm:5.1%})")
        out.append({"kernel": "add_noise", "case": f"B={n}", "us": us, "gbs": gbs, "frac": gbs / hbm})
        xp, x0 = torch.empty_like(x), torch.empty_like(x)
        tb = s._tables(x.device)[:4]
        us = tm.run(lambda: rt.sched_prev_step(x, e, 500, tb, x_prev=xp, x0=x0))
        gbs = 4 * 4.0 * x.numel() / us / 1e3
        print(f"prev_step B={n:6d} {us:9.1f} us {gbs:7.0f} GB/s ({gbs / hbm:5.1%})  [z drawn in-kernel]")
        out.append({"kernel": "prev_step", "case": f"B={n}", "us": us, "gbs": gbs, "frac": gbs / hbm})


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    tm = Timer(a.iters)
    out = []
    for name, fn in [("igemm", bench_igemm), ("groupnorm", bench_groupnorm), ("groupnorm_bwd", bench_groupnorm_bwd),
                     ("attention", bench_attention),
                     ("sched", bench_sched)]:
        if not a.only or a.only == name:
            fn(tm, out)
    if a.json:
        with open(a.json, "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()

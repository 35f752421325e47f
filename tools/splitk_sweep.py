"""Ground truth for the split-K / column-tile choice of small training GEMMs: times one convolution shape per
(B200SD_FORCE_BN, B200SD_FORCE_SPLITS) setting, each in its own process (the switches are read at launch time, the
kernel attributes are latched per process).

    python tools/splitk_sweep.py            # driver: runs every case x BN x S as a subprocess, prints a table
    python tools/splitk_sweep.py n h w cin cout k   # worker: prints the time of that shape under the current env
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = [(16, 4, 4, 512, 512, 3), (16, 4, 4, 768, 768, 3), (16, 8, 8, 768, 768, 3), (16, 8, 8, 384, 384, 3),
         (16, 16, 16, 768, 256, 3), (16, 16, 16, 256, 256, 3), (16, 8, 8, 1280, 512, 3)]


def worker(n, h, w, cin, cout, k):
    import torch
    import sdb200  # noqa: F401
    from sdb200 import ops, runtime as rt
    dev = torch.device("cuda:0")
    x = torch.randn(n * h * w, cin, device=dev).to(torch.bfloat16)
    wt = ops.pack_conv_weight(torch.randn(cout, cin, k, k, device=dev) / (cin * k * k) ** 0.5)
    out = torch.empty(n * h * w, cout, device=dev, dtype=torch.bfloat16)
    flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)
    with torch.enable_grad():
        for _ in range(3):
            ops.conv_same(x, n, h, w, cin, wt, cout, k, out=out)
        ts = []
        for _ in range(10):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops.conv_same(x, n, h, w, cin, wt, cout, k, out=out)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    print(f"{ts[len(ts) // 2]:.1f} {int(rt.lib().b200sd_igemm_last_splits())}")


def main():
    if len(sys.argv) > 1:
        return worker(*map(int, sys.argv[1:7]))
    for case in CASES:
        n, h, w, cin, cout, k = case
        print(f"M={n * h * w} N={cout} chunks={k * k * ((cin + 63) // 64)}")
        for bn in (64, 128, 256):
            row = []
            for s in (1, 2, 3, 4, 6, 8):
                env = dict(os.environ, B200SD_FORCE_BN=str(bn), B200SD_FORCE_SPLITS=str(s))
                r = subprocess.run([sys.executable, __file__] + [str(v) for v in case], env=env, capture_output=True,
                                   text=True)
                row.append(r.stdout.strip().replace(" ", "/S") if r.returncode == 0 else "err")
            print(f"   BN={bn:3d}: " + "  ".join(f"S{s}:{v}" for s, v in zip((1, 2, 3, 4, 6, 8), row)))


if __name__ == "__main__":
    main()

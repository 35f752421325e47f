"""Sweep the igemm column tile (B200SD_FORCE_BN) over plain GEMM shapes, timing 20 back-to-back launches inside one
CUDA graph replay (no host launch overhead, inputs rotate over 4 buffers so consecutive launches do not hit the same
L2 lines):
    python tools/gemm_sweep.py [--shapes dit|unet]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import sdb200  # noqa: E402,F401
from sdb200 import ops  # noqa: E402
from sdb200 import runtime as rt  # noqa: E402

SHAPES = {
    "dit": [(16384, 1152, 288, "relu"), (16384, 864, 288, None), (16384, 288, 1152, "res"), (16384, 288, 288, "res")],
    "unet": [(65536, 1152, 384, None), (65536, 384, 384, "res"), (16384, 1536, 512, None), (16384, 512, 512, "res"),
             (4096, 2304, 768, None), (4928, 384, 512, None)],
    # the 1x1 / projection GEMMs of one batch-64 UNet step whose K is short (profiles/r1k: 1.25 ms of the step)
    "short": [(65536, 384, 384, "res"), (65536, 384, 384, None), (16384, 512, 512, "res"), (65536, 128, 128, "res"),
              (16384, 256, 256, "res"), (4096, 768, 768, "res"), (4096, 384, 384, "res"), (1024, 512, 512, "res"),
              (65536, 384, 128, None), (16384, 768, 256, None), (65536, 1152, 384, None), (16384, 256, 1024, None)],
    # out-projections with the fused partial GroupNorm statistics ("res+stats:<rows per image>")
    "stats": [(65536, 384, 384, "res"), (65536, 384, 384, "res+stats:1024"), (65536, 128, 128, "res"),
              (65536, 128, 128, "res+stats:1024"), (16384, 512, 512, "res"), (16384, 512, 512, "res+stats:256"),
              (4096, 768, 768, "res"), (4096, 768, 768, "res+stats:64"), (1024, 512, 512, "res+stats:16"),
              (65536, 1152, 384, None), (65536, 1152, 384, "stats:1024")],
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="dit")
    ap.add_argument("--dual-maxchunks", type=int, default=None,
                    help="B200SD_DUAL_MAXCHUNKS for this process (0 = never two CTAs per SM)")
    ap.add_argument("--bns", default="auto,64,128,256")
    a = ap.parse_args()
    if a.dual_maxchunks is not None:
        os.environ["B200SD_DUAL_MAXCHUNKS"] = str(a.dual_maxchunks)
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(0)
    for M, N, K, epi in SHAPES[a.shapes]:
        xs = [torch.randn(M, K, generator=g).to(dev, torch.bfloat16) for _ in range(4)]
        w = ops.pack_linear_weight(torch.randn(N, K, generator=g).to(dev) * 0.05)
        bias = torch.randn(N, generator=g).to(dev)
        res = torch.randn(M, N, generator=g).to(dev, torch.bfloat16) if epi and epi.startswith("res") else None
        rpi, stats = None, None
        if epi and "stats:" in epi:
            rpi = int(epi.split(":")[1])
            slots = rt.stats_slots(1, rpi)
            stats = (torch.zeros(M // rpi, slots, N, 2, device=dev), slots, 0, N)
        outs = [torch.empty(M, N, device=dev, dtype=torch.bfloat16) for _ in range(4)]
        line = f"M={M:6d} N={N:5d} K={K:5d} {str(epi):15s}"
        for bn in a.bns.split(","):
            if bn == "auto":
                os.environ.pop("B200SD_FORCE_BN", None)
            else:
                os.environ["B200SD_FORCE_BN"] = bn

            def launch(i):
                rt.linear(xs[i % 4], w, bias=bias, act="relu" if epi == "relu" else None, residual=res, out=outs[i % 4],
                          rows_per_img=rpi, stats=stats)

            for i in range(4):
                launch(i)
            torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                for i in range(20):
                    launch(i)
            gr.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                gr.replay()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / 100
            line += f" | BN={bn:4s} {us:6.1f} us {2.0 * M * N * K / us / 1e6:6.0f} TF/s"
        print(line)
    os.environ.pop("B200SD_FORCE_BN", None)


if __name__ == "__main__":
    main()

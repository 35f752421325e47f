"""Per-launch time of b200sd_lsq_quant_bwd inside a replayed CUDA graph (how the training legs run it) against the number
of CTAs per SM it is launched with (B200SD_QBWD_CAP, read once per process -> one subprocess per value).

    python tools/quant_bwd_sweep.py            # sweep
    python tools/quant_bwd_sweep.py --one      # one measurement with the current environment
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SHAPES = [(4096, 384), (4096, 1152), (4096, 1536), (1232, 384), (16384, 384)]


def one():
    import torch
    import sdb200  # noqa: F401
    from sdb200 import runtime as rt
    dev = torch.device("cuda:0")
    out = {}
    for rows, c in SHAPES:
        n = 24
        xs = [torch.randn(rows, c, device=dev).bfloat16() for _ in range(n)]
        dys = [torch.randn(rows, c, device=dev).bfloat16() for _ in range(n)]
        step = torch.tensor(0.05, device=dev)
        for i in range(2):
            rt.lsq_quant_bwd(dys[i], xs[i], step, 127.0)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(n):
                rt.lsq_quant_bwd(dys[i], xs[i], step, 127.0)
        for _ in range(3):
            g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        out[f"{rows}x{c}"] = round(e0.elapsed_time(e1) / (20 * n) * 1000, 2)
    print(json.dumps({"cap": os.environ.get("B200SD_QBWD_CAP", "default"), "us_per_launch": out}), flush=True)


if __name__ == "__main__":
    if "--one" in sys.argv:
        one()
    else:
        for cap in ("1", "2", "4", "8", "16"):
            env = dict(os.environ, B200SD_QBWD_CAP=cap)
            subprocess.run([sys.executable, os.path.abspath(__file__), "--one"], env=env, check=False)

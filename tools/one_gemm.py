"""Launch one plain GEMM shape a few times (target for `ncu --set full`): python tools/one_gemm.py M N K [relu|res|res+stats:ROWS_PER_IMG|stats:ROWS_PER_IMG]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import sdb200  # noqa: E402,F401
from sdb200 import ops  # noqa: E402
from sdb200 import runtime as rt  # noqa: E402

M, N, K = (int(v) for v in sys.argv[1:4])
epi = sys.argv[4] if len(sys.argv) > 4 else None
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
x = torch.randn(M, K, generator=g).to(dev, torch.bfloat16)
w = ops.pack_linear_weight(torch.randn(N, K, generator=g).to(dev) * 0.05)
bias = torch.randn(N, generator=g).to(dev)
res = torch.randn(M, N, generator=g).to(dev, torch.bfloat16) if epi and epi.startswith("res") else None
rpi, stats = None, None
if epi and "stats:" in epi:
    rpi = int(epi.split(":")[1])
    slots = rt.stats_slots(1, rpi)
    stats = (torch.zeros(M // rpi, slots, N, 2, device=dev), slots, 0, N)
out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
for _ in range(4):
    rt.linear(x, w, bias=bias, act="relu" if epi == "relu" else None, residual=res, out=out, rows_per_img=rpi,
              stats=stats)
torch.cuda.synchronize()
print("ok", out.float().abs().mean().item())

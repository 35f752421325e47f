"""One GroupNorm-backward shape, a few calls (for ncu): python tools/one_gn_bwd.py n hw c [calls]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import sdb200  # noqa: E402,F401
from sdb200 import runtime as rt  # noqa: E402

n, hw, c = (int(v) for v in sys.argv[1:4])
calls = int(sys.argv[4]) if len(sys.argv) > 4 else 4
x = torch.randn(n * hw, c, device="cuda").to(torch.bfloat16)
dy = torch.randn(n * hw, c, device="cuda").to(torch.bfloat16)
extra = torch.randn(n * hw, c, device="cuda").to(torch.bfloat16)
xf = x.float().view(n, hw, c)
stats = torch.stack([xf.sum(1), (xf * xf).sum(1)], dim=-1).view(n, 1, c, 2).contiguous()
g, b = torch.randn(c, device="cuda"), torch.randn(c, device="cuda")
for _ in range(calls):
    rt.groupnorm_bwd(x, dy, n, hw, c, 32, g, b, [(stats, 1, c, c)], act="silu", add=extra)
torch.cuda.synchronize()
print("ok")

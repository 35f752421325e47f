import sys, torch, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from util import cond_of, l2err, load_golden, relerr
from sdb200 import engine as E
from sdb200 import runtime as rt
dev = torch.device("cuda:0")
# unit checks of the new kernels first
g = torch.Generator().manual_seed(0)
x = torch.randn(300, 37, generator=g).to(dev)
sp = rt.split_f32(x, 37)
hi, lo = sp[:, :37].float(), sp[:, 40:77].float()
print("split err", ((hi + lo) - x).abs().max().item(), "hi-only err", (hi - x).abs().max().item())
n, hw, c, G = 3, 64, 96, 16
xf = torch.randn(n * hw, c, generator=g).to(dev) * 2 + 0.5
gam, bet = torch.randn(c, generator=g).to(dev), torch.randn(c, generator=g).to(dev)
y = rt.groupnorm_split(xf, n, hw, c, G, gam, bet, 1e-5, act="silu")
ref = torch.nn.functional.silu(torch.nn.functional.group_norm(xf.view(n, hw, c).permute(0, 2, 1), G, gam, bet, 1e-5)).permute(0, 2, 1).reshape(n * hw, c)
print("gn_split err", ((y[:, :c].float() + y[:, c:2 * c].float()) - ref).abs().max().item())
for name in sys.argv[1:]:
    fx = load_golden(name)
    if "uncond" in name:
        from models.unet_base import Unet
    elif fx.get("relu"):
        from models.unet_cond_base_relu import Unet
    else:
        from models.unet_cond_base import Unet
    torch.manual_seed(fx["seed"])
    model = Unet(4, fx["config"]).eval().to(dev)
    inp, cfg = fx["inputs"], fx["config"]
    cond = cond_of(inp, cfg, dev)
    args = (inp["x"].to(dev), inp["t"].to(dev)) + ((cond,) if cond is not None else ())
    for mode in ("fast", "precise"):
        E.set_parity_mode(mode)
        with torch.no_grad():
            out = model(*args)
        torch.cuda.synchronize()
        print(name, mode, f"max-norm {relerr(out.cpu(), fx['output']):.3e} L2 {l2err(out.cpu(), fx['output']):.3e}", flush=True)

"""Where does the whole-network bf16 error come from?  

Runs the fp32 CPU oracle of the UNet (oracle/unet_torch.py — test infrastructure; this tool is a test tool, not product
code) on the committed golden fixtures with the PRODUCT pipeline's rounding points emulated family by family, and prints
max-norm / relative-L2 error against the reference's fp32 output for: everything on (= what the sm_100a pipeline does),
each family alone, and everything on except one family. No GPU needed.

Families (the places where the sm_100a pipeline holds a value in bf16):
  W     tensor-core B operand: conv / linear weights rounded to bf16
  A     tensor-core A operand: the activation a conv / linear reads (GroupNorm+SiLU output, attention output, t-emb)
  S     stored GEMM results: residual stream, conv outputs, time-embedding rows, q/k/v (what the next GroupNorm /
        attention / residual add reads back from HBM)
  P     attention internals: softmax probabilities rounded to bf16 before P.V
  Q     attention operands: q, k, v as the attention kernel reads them and its output O as out_proj reads it (part of S / A
        in the fast pipeline; separate here because the precise mode keeps the bf16 attention kernel)
`split` columns emulate an operand kept as TWO bf16 values (hi + lo, 16 mantissa bits): the split-operand GEMM of the
precise mode.

    python tools/error_budget.py [--cases unet_full_cond unet_full_uncond] [--out profiles/r2_error_budget.md]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
import torch.nn.functional as TF  # noqa: E402

from oracle import unet_torch as O  # noqa: E402
from util import cond_of, l2err, load_golden, relerr  # noqa: E402


def bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


def split2(x):
    hi = bf(x)
    return hi + bf(x - hi)


class Rounding:
    def __init__(self, W="f32", A="f32", S="f32", P="f32", Q=None):
        self.W, self.A, self.S, self.P = W, A, S, P
        self.Q = Q            # None: q/k/v follow S and O follows A
        self.mha_calls = -1   # >= 0 while inside _mha: index of the next F.linear call (q, k, v, out_proj)

    @staticmethod
    def _r(mode, x):
        if x is None or mode == "f32":
            return x
        return bf(x) if mode == "bf16" else split2(x)

    def w(self, x):
        return self._r(self.W, x)

    def a(self, x):
        return self._r(self.A, x)

    def s(self, x):
        return self._r(self.S, x)

    def p(self, x):
        return self._r(self.P, x)


class _F:
    """Stand-in for torch.nn.functional inside the oracle: GEMM-shaped calls round their operands / results."""

    def __init__(self, r):
        self.r = r

    def __getattr__(self, name):
        return getattr(TF, name)

    def conv2d(self, x, w, b=None, **kw):
        return self.r.s(TF.conv2d(self.r.a(x), self.r.w(w), b, **kw))

    def conv_transpose2d(self, x, w, b=None, **kw):
        return self.r.s(TF.conv_transpose2d(self.r.a(x), self.r.w(w), b, **kw))

    def linear(self, x, w, b=None):
        r = self.r
        if r.Q is not None and r.mha_calls >= 0:
            i, r.mha_calls = r.mha_calls, r.mha_calls + 1
            if i < 3:      # q, k, v projections: output read by the attention kernel
                return r._r(r.Q, TF.linear(r.a(x), r.w(w), b))
            return r.s(TF.linear(r._r(r.Q, x), r.w(w), b))   # out_proj reads the attention kernel's output
        return r.s(TF.linear(r.a(x), r.w(w), b))


def run(case, r):
    fx = load_golden(case)
    relu = bool(fx.get("relu"))
    if "uncond" in case:
        from models.unet_base import Unet
    elif relu:
        from models.unet_cond_base_relu import Unet
    else:
        from models.unet_cond_base import Unet
    torch.manual_seed(fx["seed"])
    sd = Unet(4, fx["config"]).state_dict()
    inp, cfg = fx["inputs"], fx["config"]
    saved_F, saved_softmax = O.F, torch.softmax
    O.F = _F(r)

    def softmax(x, dim=-1):
        return r.p(saved_softmax(x, dim=dim))

    O.torch = type("T", (), {"softmax": staticmethod(softmax), "__getattr__": lambda self, n: getattr(torch, n)})()
    saved_mha = O._mha

    def mha(*a, **k):
        r.mha_calls = 0
        try:
            return saved_mha(*a, **k)
        finally:
            r.mha_calls = -1

    O._mha = mha
    saved_blocks = {n: getattr(O, n) for n in ("_resnet", "_self_attn", "_cross_attn")}
    for n, fn in saved_blocks.items():   # the residual add happens in the GEMM epilogue; its sum is what gets stored
        setattr(O, n, (lambda fn: lambda *a, **k: r.s(fn(*a, **k)))(fn))
    try:
        with torch.no_grad():
            out = O.unet_forward(sd, cfg, inp["x"], inp["t"], cond_of(inp, cfg), relu=relu)
    finally:
        O.F, O.torch, O._mha = saved_F, torch, saved_mha
        for n, fn in saved_blocks.items():
            setattr(O, n, fn)
    return relerr(out, fx["output"]), l2err(out, fx["output"]), fx


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", nargs="*", default=["unet_full_cond", "unet_full_uncond"])
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    fams = ["W", "A", "S", "P"]
    rows = [("fp32 everywhere (oracle vs golden)", {})]
    rows.append(("all bf16 = the fast sm_100a pipeline's rounding points", {f: "bf16" for f in fams}))
    for f in fams:
        rows.append((f"only {f} bf16", {f: "bf16"}))
    for f in fams:
        rows.append((f"all bf16 except {f} (fp32)", {g: "bf16" for g in fams if g != f}))
    rows.append(("operands W, A bf16; stored values fp32", {"W": "bf16", "A": "bf16"}))
    rows.append(("W split (hi+lo), rest bf16", {"W": "split", "A": "bf16", "S": "bf16", "P": "bf16"}))
    rows.append(("W, A split; S, P bf16", {"W": "split", "A": "split", "S": "bf16", "P": "bf16"}))
    rows.append(("W, A split; S fp32; P bf16   <- precise mode", {"W": "split", "A": "split", "P": "bf16"}))
    rows.append(("W, A, S split; P bf16", {"W": "split", "A": "split", "S": "split", "P": "bf16"}))
    rows.append(("W, A split; S fp32; P, q/k/v/O bf16   <- precise mode with the bf16 attention kernel",
                 {"W": "split", "A": "split", "P": "bf16", "Q": "bf16"}))
    rows.append(("W split; A bf16; S fp32; P, q/k/v/O bf16", {"W": "split", "A": "bf16", "P": "bf16", "Q": "bf16"}))
    lines = ["| rounding points | " + " | ".join(f"{c}: max-norm / rel-L2" for c in args.cases) + " |",
             "|---|" + "---|" * len(args.cases)]
    ref_line = []
    for name, kw in rows:
        cells = []
        for c in args.cases:
            e, l2, fx = run(c, Rounding(**kw))
            cells.append(f"{e:.2e} / {l2:.2e}")
            if name.startswith("fp32"):
                ref_line.append(f"{c}: reference's own bf16-autocast run vs its fp32 run "
                                f"{fx['ref_bf16_autocast_relerr']:.2e} / {fx['ref_bf16_autocast_l2err']:.2e}")
        lines.append(f"| {name} | " + " | ".join(cells) + " |")
        print(lines[-1], flush=True)
    text = "\n".join(["# Error budget of the whole-network noise prediction (tools/error_budget.py, CPU emulation)", "",
                      "max|a-b|/max|b| and ||a-b||/||b|| against the reference's fp32 output (tests/golden).", ""]
                     + ref_line + [""] + lines) + "\n"
    if args.out:
        with open(args.out, "w") as f:
            f.write(text)
    print(text)


if __name__ == "__main__":
    main()

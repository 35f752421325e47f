"""Per-function SASS census of the shipped library: which kernels use the Blackwell tensor-core / TMA instructions.

    python tools/sass_census.py [--out profiles/sass_census.txt]

Counts, per kernel of stablediffusion-pytorch_b200/lib/libb200sd.so (cuobjdump -sass): UTCHMMA / UTCQMMA (tcgen05.mma),
LDTM / STTM (tcgen05.ld / st), UTMALDG / UTMASTG (TMA tensor load / store), UTCBAR (tcgen05.commit), HMMA (warp-level
mma.sync), LDGSTS (cp.async), MUFU. Runs without a GPU."""
import argparse
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "stablediffusion-pytorch_b200", "lib", "libb200sd.so")
OPS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "HMMA", "LDGSTS", "MUFU"]


def demangle(names):
    try:
        out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout
        return out.splitlines()
    except (OSError, subprocess.CalledProcessError):
        return names


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts, order, cur = collections.defaultdict(collections.Counter), [], None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            order.append(cur)
            continue
        if cur is None:
            continue
        m = re.search(r"/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(1).split(".")[0]
            counts[cur]["_total"] += 1
            if op in OPS:
                counts[cur][op] += 1
    names = demangle(order)
    rows = []
    for mangled, name in zip(order, names):
        short = re.sub(r"\(.*", "", name).replace("b200::", "")
        rows.append((short, counts[mangled]))
    rows.sort(key=lambda r: r[0])
    w = max(len(r[0]) for r in rows)
    lines = [f"# SASS census of {os.path.relpath(LIB, ROOT)} (tools/sass_census.py; cuobjdump -sass, sm_100a)",
             f"{'kernel':<{w}}  {'instrs':>7}  " + "  ".join(f"{o:>7}" for o in OPS)]
    for short, c in rows:
        lines.append(f"{short:<{w}}  {c['_total']:>7}  " + "  ".join(f"{c[o]:>7}" for o in OPS))
    text = "\n".join(lines) + "\n"
    if args.out:
        with open(args.out, "w") as f:
            f.write(text)
    print(text)


if __name__ == "__main__":
    main()

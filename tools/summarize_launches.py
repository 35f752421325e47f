"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel totals and shares.

    python tools/summarize_launches.py gpurun_out/launches.csv [--last N] > profiles/launches_rXX.md

--step N keeps the launches of the N-th sampling step (between advance_t_kernel launches); --last N keeps only the final N launches (e.g. the launches of the last timed step). Times under ncu are cold-cache
and serialised: compare SHARES with bench.py's CUDA-event shares, not absolutes.
"""
import csv
import re
import sys
from collections import OrderedDict


def short(name):
    name = re.sub(r"^void ", "", name)
    m = re.match(r"(?:b200::)?([A-Za-z0-9_]+)(<[^>]*>)?", name)
    return (m.group(1) + (m.group(2) or "")) if m else name[:60]


def main():
    path = sys.argv[1]
    last = int(sys.argv[sys.argv.index("--last") + 1]) if "--last" in sys.argv else 0
    rows = []
    with open(path, newline="") as f:
        started = False
        for r in csv.reader(f):
            if not started:
                started = bool(r) and r[0] == "ID"
                continue
            if len(r) >= 15 and r[-3] == "gpu__time_duration.sum":
                ns = float(r[-1]) * {"ns": 1, "us": 1e3, "ms": 1e6}.get(r[-2], 1)
                rows.append((short(r[4]), ns))
    if last:
        rows = rows[-last:]
    if "--step" in sys.argv:  # one sampling step = the launches between two consecutive advance_t_kernel launches
        idx = [i for i, (k, _) in enumerate(rows) if k.startswith("advance_t_kernel")]
        n = int(sys.argv[sys.argv.index("--step") + 1])
        rows = rows[idx[n - 1] + 1: idx[n] + 1]
    agg = OrderedDict()
    for k, ns in rows:
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += ns
    tot = sum(v[1] for v in agg.values())
    print(f"launches: {len(rows)}  total device time {tot / 1e6:.3f} ms (ncu: cold-cache, serialised)\n")
    print("| kernel | launches | total us | avg us | share |")
    print("|---|---:|---:|---:|---:|")
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {n} | {ns / 1e3:.1f} | {ns / 1e3 / n:.2f} | {100 * ns / tot:.1f}% |")


if __name__ == "__main__":
    main()

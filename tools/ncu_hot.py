"""Top stall locations of one kernel from an .ncu-rep source page (SASS view): python tools/ncu_hot.py file.ncu-rep [N]"""
import csv
import subprocess
import sys


def main():
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    h = rows[hi]
    ci = {k: i for i, k in enumerate(h)}
    stall_cols = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
    data = []
    for r in rows[hi + 1:]:
        if len(r) < len(h):
            continue
        try:
            n = int(r[ci["# Samples"]] or 0)
        except ValueError:
            continue
        data.append((n, r))
    tot = sum(n for n, _ in data) or 1
    N = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    print(f"total samples {tot}")
    agg = {k: 0 for k in stall_cols}
    for n, r in data:
        for k in stall_cols:
            try:
                agg[k] += int(r[ci[k]] or 0)
            except ValueError:
                pass
    print("stall mix:", ", ".join(f"{k[6:]} {100 * v / tot:.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
    for n, r in sorted(data, key=lambda x: -x[0])[:N]:
        top = sorted(((int(r[ci[k]] or 0), k[6:]) for k in stall_cols), reverse=True)[:2]
        print(f"{100 * n / tot:5.1f}%  {r[ci['Source']][:90]:90s} {top[0][1]}:{top[0][0]} {top[1][1]}:{top[1][0]}")


if __name__ == "__main__":
    main()

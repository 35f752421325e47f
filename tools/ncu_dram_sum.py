"""Average DRAM traffic per launch of one kernel family from an ncu metrics CSV.

    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
        -k regex:igemm --csv --log-file gpurun_out/igemm_dram.csv python bench.py --steps 2 --warmup 3 ...
    python tools/ncu_dram_sum.py gpurun_out/igemm_dram.csv --last 106 --out profiles/igemm_traffic.json

--last N keeps the final N launches (one sampling step has 106 igemm launches). bench.py copies
`dram_bytes_per_launch` from the JSON into roofline.traffic.
"""
import csv
import json
import sys

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}


def main():
    path = sys.argv[1]
    last = int(sys.argv[sys.argv.index("--last") + 1]) if "--last" in sys.argv else 0
    out = sys.argv[sys.argv.index("--out") + 1] if "--out" in sys.argv else ""
    per = {}
    order = []
    with open(path, newline="") as f:
        started = False
        for r in csv.reader(f):
            if not started:
                started = bool(r) and r[0] == "ID"
                continue
            if len(r) < 15:
                continue
            kid, metric, unit, val = r[0], r[-3], r[-2], float(r[-1].replace(",", ""))
            if kid not in per:
                per[kid] = {"name": r[4]}
                order.append(kid)
            per[kid][metric] = val * UNIT.get(unit, 1)
    ids = order[-last:] if last else order
    rd = sum(per[k].get("dram__bytes_read.sum", 0.0) for k in ids)
    wr = sum(per[k].get("dram__bytes_write.sum", 0.0) for k in ids)
    us = sum(per[k].get("gpu__time_duration.sum", 0.0) for k in ids)
    res = {"launches": len(ids), "dram_read_bytes": rd, "dram_write_bytes": wr,
           "dram_bytes_per_launch": (rd + wr) / max(len(ids), 1), "sum_duration_us": us,
           "source": "ncu dram__bytes_read.sum + dram__bytes_write.sum over the igemm launches of one sampling step"}
    print(json.dumps(res, indent=1))
    if out:
        with open(out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()

"""Where does the gradient exchange of the captured data-parallel training step run?

One variant per invocation (environment switches such as NCCL_MAX_CTAS / B200SD_SM_LIMIT must be set before the first
CUDA / NCCL call, so the caller sets them on the command line):

    [NCCL_MAX_CTAS=8 B200SD_SM_LIMIT=140] python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 \
        --master-addr 127.0.0.1 --master-port 29541 tools/overlap_probe.py --label ctas8_sm140 \
        [--mode buckets|after|none] [--bucket-mb 64] [--timeline gpurun_out/timeline.json]

  --mode buckets  the product path: bucketed NCCL all-reduce captured inside the backward graph (graphs._BucketReducer)
  --mode after    one collective after the backward graph (nothing can overlap: the upper bound of the exposure)
  --mode none     no exchange at all (each rank trains alone: the lower bound)

Prints one JSON line: ms per forward+backward step (max over ranks) and, from a torch.profiler (CUPTI) trace of one
replayed step on rank 0, where the ncclDevKernel_* launches sit relative to the compute kernels: their total busy time,
how much of it is covered by a compute kernel running at the same time, and how long the step runs on after the last
compute kernel has finished (the exposed tail). --timeline writes the per-kernel (start, duration, stream, name) list.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
import sdb200  # noqa: E402,F401
from sdb200.parallel import DataParallel  # noqa: E402
from models.unet_cond_base import Unet  # noqa: E402


def union_len(iv):
    iv = sorted(iv)
    tot, end = 0.0, None
    for a, b in iv:
        if end is None or a > end:
            tot += b - a
            end = b
        elif b > end:
            tot += b - end
            end = b
    return tot


def covered(iv, by):
    """Length of the part of the intervals `iv` that lies inside the union of `by`."""
    by = sorted(by)
    merged = []
    for a, b in by:
        if merged and a <= merged[-1][1]:
            merged[-1][1] = max(merged[-1][1], b)
        else:
            merged.append([a, b])
    tot = 0.0
    for a, b in iv:
        for c, d in merged:
            lo, hi = max(a, c), min(b, d)
            if hi > lo:
                tot += hi - lo
    return tot


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--label", default="default")
    ap.add_argument("--mode", default="buckets", choices=["buckets", "after", "none"])
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--bucket-mb", type=float, default=None)
    ap.add_argument("--timeline", default=None)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    ucfg, _ = bench.load_cfg()
    torch.manual_seed(1111)
    model = Unet(4, ucfg).train().to(dev)
    g = torch.Generator().manual_seed(100 + rank)
    x = torch.randn(a.batch, 4, 32, 32, generator=g).to(dev)
    text = torch.randn(a.batch, 77, 512, generator=g).to(dev)
    mask = (torch.rand(a.batch, 18, 64, 64, generator=g) > 0.5).float().to(dev)
    t = torch.randint(0, 1000, (a.batch,), generator=g).to(dev)
    noise = torch.randn(a.batch, 4, 32, 32, generator=g).to(dev)
    net = model
    if world > 1 and a.mode != "none":
        net = DataParallel(model, device_ids=[dev.index], overlap=a.mode == "buckets",
                           bucket_cap_mb=a.bucket_mb if a.mode == "buckets" else None)

    def step():
        model.zero_grad(set_to_none=True)
        pred = net(x, t, {"text": text, "image": mask})
        torch.nn.functional.mse_loss(pred, noise).backward()

    for _ in range(4):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / a.steps], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    out = {"label": a.label, "mode": a.mode, "world": world, "batch_per_gpu": a.batch, "ms_per_step": round(ms.item(), 3),
           "env": {k: os.environ.get(k) for k in ("NCCL_MAX_CTAS", "NCCL_MIN_CTAS", "NCCL_NTHREADS", "B200SD_SM_LIMIT", "NCCL_ALGO",
                                                   "NCCL_CGA_CLUSTER_SIZE", "NCCL_NVLS_ENABLE") if os.environ.get(k)}}
    tg = model.__dict__.get("_train_graph")
    if tg is not None:
        out["reduced_in_graph"], out["buckets"] = bool(tg.reduced_in_graph), tg.n_buckets

    # ---- timeline of one replayed step (all ranks run it: the collectives need every rank)
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        step()
        torch.cuda.synchronize()
    if rank == 0:
        ev = []
        for e in prof.events():
            if e.device_type is not None and "cuda" in str(e.device_type).lower():
                dur = e.device_time if hasattr(e, "device_time") else e.cuda_time
                ev.append((e.time_range.start, dur, e.name))
        ev.sort()
        if ev:
            t0 = ev[0][0]
            nccl = [(s - t0, s - t0 + d) for s, d, n in ev if "nccl" in n.lower()]
            comp = [(s - t0, s - t0 + d) for s, d, n in ev if "nccl" not in n.lower() and not n.startswith("Mem")]
            heavy = [(s - t0, s - t0 + d) for s, d, n in ev if "igemm_kernel" in n or "wgrad_kernel" in n]
            end_all = max(b for _, b in nccl + comp)
            end_comp = max(b for _, b in comp)
            out["timeline"] = {
                "kernels": len(ev), "span_us": round(end_all, 1), "compute_busy_us": round(union_len(comp), 1),
                "nccl_launches": len(nccl), "nccl_busy_us": round(union_len(nccl), 1),
                "nccl_under_any_compute_us": round(covered(nccl, comp), 1),
                "nccl_under_igemm_wgrad_us": round(covered(nccl, heavy), 1),
                "tail_after_last_compute_us": round(end_all - end_comp, 1),
                "nccl_intervals_us": [[round(p, 1), round(q, 1)] for p, q in nccl],
            }
            if a.timeline:
                os.makedirs(os.path.dirname(os.path.abspath(a.timeline)), exist_ok=True)
                with open(a.timeline, "w") as f:
                    json.dump([[round(s - t0, 2), round(d, 2), n[:60]] for s, d, n in ev], f)
        print(json.dumps(out), flush=True)
    if world > 1:
        bench.shutdown(dist)


if __name__ == "__main__":
    main()

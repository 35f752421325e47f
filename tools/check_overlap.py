"""Two-rank check of the in-graph bucketed gradient all-reduce (graphs._BucketReducer) against the one-collective
path: same seeds, same per-rank batches, gradients must be bit-identical; then times both.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tools/check_overlap.py [--batch 16] [--steps 10]
"""
import argparse
import copy
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--bucket-mb", type=float, default=None)
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    import faulthandler
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    logf = open(os.path.join(ROOT, "gpurun_out", f"check_overlap_rank{rank}.log"), "w")
    faulthandler.dump_traceback_later(90, exit=True, file=logf)

    def say(*a_):
        print(*a_, file=logf, flush=True)
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    ucfg, _ = bench.load_cfg()
    torch.manual_seed(1111)
    base = Unet(4, ucfg).train().to(dev)
    g = torch.Generator().manual_seed(100 + rank)   # every rank its own batch
    x = torch.randn(a.batch, 4, 32, 32, generator=g).to(dev)
    text = torch.randn(a.batch, 77, 512, generator=g).to(dev)
    mask = (torch.rand(a.batch, 18, 64, 64, generator=g) > 0.5).float().to(dev)
    t = torch.randint(0, 1000, (a.batch,), generator=g).to(dev)
    noise = torch.randn(a.batch, 4, 32, 32, generator=g).to(dev)
    res = {}
    for name, ov in (("one_collective", False), ("in_graph_buckets", True)):
        say("variant", name)
        model = copy.deepcopy(base)
        net = DataParallel(model, device_ids=[dev.index], overlap=ov, bucket_cap_mb=a.bucket_mb if ov else None)

        def step():
            model.zero_grad(set_to_none=True)
            pred = net(x, t, {"text": text, "image": mask})
            torch.nn.functional.mse_loss(pred, noise).backward()

        for k in range(3):
            step()
            torch.cuda.synchronize()
            say(" warm step", k, "done")
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / a.steps], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        tg = model.__dict__["_train_graph"]
        res[name] = ([p.grad.clone() for p in model.parameters() if p.grad is not None], ms.item(),
                     tg.reduced_in_graph, tg.n_buckets)
        say(" timed", ms.item())
        del net, model
    ga, gb = res["one_collective"][0], res["in_graph_buckets"][0]
    assert len(ga) == len(gb)
    same = all(torch.equal(u, v) for u, v in zip(ga, gb))
    # the mean really is a mean over ranks: rank 0 and rank 1 hold the same gradients
    chk = torch.stack([v.double().sum() for v in gb]).sum().reshape(1)
    both = [torch.zeros_like(chk) for _ in range(world)]
    dist.all_gather(both, chk)
    if rank == 0:
        print(f"one collective after the graph : {res['one_collective'][1]:.3f} ms/step")
        print(f"buckets inside the graph       : {res['in_graph_buckets'][1]:.3f} ms/step "
              f"(reduced_in_graph={res['in_graph_buckets'][2]}, buckets={res['in_graph_buckets'][3]})")
        print("gradients identical:", same, " ranks agree:", all(torch.equal(b, both[0]) for b in both))
    assert same and res["in_graph_buckets"][2]
    logf.flush()
    bench.shutdown(dist)


if __name__ == "__main__":
    main()

"""Host-side logic of the N>1 training path on CPU: two gloo ranks (SURVEY.md §8 row e / a15). The device path is
the same code over NCCL."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sdb200.parallel import DataParallel, _mean_
        torch.manual_seed(100 + rank)                       # ranks start from different weights
        net = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.ReLU(), torch.nn.Linear(7, 3))
        dp = DataParallel(net, bucket_cap_mb=None)
        w0 = torch.cat([p.detach().flatten() for p in net.parameters()])
        gathered = [torch.empty_like(w0) for _ in range(world)]
        dist.all_gather(gathered, w0)
        same_init = all(torch.equal(g, gathered[0]) for g in gathered)

        # eager path: hooks reduce once the last gradient has landed; two backward passes to check the latch resets
        ok_grad = True
        for it in range(2):
            g = torch.Generator().manual_seed(7 * it + rank)
            x = torch.randn(4, 5, generator=g)
            net.zero_grad()
            dp(x).square().sum().backward()
            mine = torch.cat([p.grad.flatten() for p in net.parameters()])
            # the same thing by hand: local gradients of every rank's batch, averaged
            ref = torch.zeros_like(mine)
            for r in range(world):
                xr = torch.randn(4, 5, generator=torch.Generator().manual_seed(7 * it + r))
                net2 = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.ReLU(), torch.nn.Linear(7, 3))
                net2.load_state_dict(net.state_dict())
                net2(xr).square().sum().backward()
                ref += torch.cat([p.grad.flatten() for p in net2.parameters()]) / world
            ok_grad &= torch.allclose(mine, ref, rtol=1e-6, atol=1e-7)

        # a trainable parameter that never receives a gradient (unused branch) must not leave a short count behind:
        # three backward passes, each must end with every used gradient averaged over ranks
        class Branchy(torch.nn.Module):
            def __init__(self):
                super().__init__()
                self.used, self.unused = torch.nn.Linear(5, 3), torch.nn.Linear(5, 3)

            def forward(self, x):
                return self.used(x)

        torch.manual_seed(5)
        bn = Branchy()
        dpb = DataParallel(bn)
        ok_unused = True
        for it in range(3):
            xr = [torch.randn(4, 5, generator=torch.Generator().manual_seed(31 * it + r)) for r in range(world)]
            bn.zero_grad()
            dpb(xr[rank]).square().sum().backward()
            ref = torch.zeros_like(bn.used.weight)
            for r in range(world):
                b2 = Branchy()
                b2.load_state_dict(bn.state_dict())
                b2(xr[r]).square().sum().backward()
                ref += b2.used.weight.grad / world
            ok_unused &= torch.allclose(bn.used.weight.grad, ref, rtol=1e-6, atol=1e-7) and bn.unused.weight.grad is None
            ok_unused &= dpb._pending == 0
        ok_grad &= ok_unused

        # flat-buffer path (what a captured step calls), bucketed and unbucketed give the same mean
        flat = torch.arange(1000, dtype=torch.float32) * (rank + 1)
        a = _mean_(flat.clone(), None, None)
        b = _mean_(flat.clone(), None, 96)
        want = torch.arange(1000, dtype=torch.float32) * (sum(range(1, world + 1)) / world)
        ok_flat = torch.equal(a, b) and torch.allclose(a, want)
        dp._sync_flat(flat.clone())
        latch = dp._graph_synced
        q.put((rank, same_init, bool(ok_grad), bool(ok_flat), latch))
    finally:
        dist.destroy_process_group()


def test_data_parallel_two_ranks_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for rank, same_init, ok_grad, ok_flat, latch in res:
        assert same_init, f"rank {rank}: parameters not broadcast from rank 0"
        assert ok_grad, f"rank {rank}: eager-path gradients are not the mean over ranks"
        assert ok_flat, f"rank {rank}: flat-buffer mean wrong"
        assert latch


def test_bench_batch_sharding_is_disjoint():
    """Sampling shards by batch with no collective: every rank draws its own latents / conditions."""
    sys.path.insert(0, ROOT)
    import bench
    cfg = bench.load_cfg()[0]
    a = bench.synth_inputs(2, cfg, rank=0, pin=False)
    b = bench.synth_inputs(2, cfg, rank=1, pin=False)
    assert a[0].shape == b[0].shape and not torch.equal(a[0], b[0])


def test_gradient_bucket_plan_covers_the_flat_buffer_once():
    """Host logic of the in-graph gradient exchange (graphs.arrival_slots / plan_buckets / bucket_range): the flat
    buffer is laid out in gradient-arrival order (trainable parameters that never fired follow at the end, frozen ones
    get no slot); buckets walk the used slots from the first to arrive to the last, each owns one contiguous slice of
    the flat buffer, the slices are disjoint, ordered, and together cover every used slot."""
    import importlib
    G = importlib.import_module("sdb200.graphs")
    sizes = [300, 7, 4096, 64, 1, 9000, 128, 33, 70000, 5]
    trainable = [i != 4 for i in range(len(sizes))]                  # 4: requires_grad = False
    fired = [9, 8, 7, 5, 3, 0, 2]       # arrival order; 0 (used at the start of the forward) lands before 2; 6, 1 never fire
    slots, total = G.arrival_slots(sizes, [(n,) for n in sizes], trainable, fired)
    assert slots[4] is None and all(slots[i] is not None for i in range(len(sizes)) if i != 4)
    by_off = sorted((sl[0], i) for i, sl in enumerate(slots) if sl is not None)
    assert [i for _, i in by_off] == fired + [1, 6]                  # arrival order, then the rest in parameter order
    assert all(o % 64 == 0 for o, _ in by_off) and total == sum((sizes[i] + 63) // 64 * 64 for i in range(10) if i != 4)
    for cap in (1, 4000, 10000, 10 ** 9):
        buckets = G.plan_buckets(slots, fired + [2], cap)            # (2 reported twice)
        flat = [i for b in buckets for i in b]
        assert flat == fired
        ranges = [G.bucket_range(slots, b) for b in buckets]
        for (lo, hi), b in zip(ranges, buckets):
            assert lo % 64 == 0 and hi % 64 == 0 and lo < hi <= total
            assert all(lo <= slots[i][0] and slots[i][0] + slots[i][1] <= hi for i in b)
        for (lo_a, hi_a), (lo_b, hi_b) in zip(ranges, ranges[1:]):
            assert hi_a <= lo_b          # later buckets lie above earlier ones, no overlap
        if cap == 10 ** 9:
            assert len(buckets) == 1
        if cap == 1:
            assert len(buckets) == 7
        assert all(sum(slots[i][1] for i in b) >= cap for b in buckets[:-1])
    # without an arrival order (no overlap requested) the layout is plain parameter order
    plain, _ = G.arrival_slots(sizes, [(n,) for n in sizes], trainable, [])
    offs = [sl[0] for sl in plain if sl is not None]
    assert offs == sorted(offs)

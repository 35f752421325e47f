"""Data-parallel training of the denoiser: the one collective on the path is the gradient mean over ranks
(reference: DistributedDataParallel(model, device_ids=[rank], broadcast_buffers=False),
train_ddpm_cond_celebhq_multi_gpu.py:257-263; the reducer fires from loss.backward(), :362).

`DataParallel` is the one-line replacement for that wrap. It keeps DDP's contract — parameters broadcast from rank 0
at wrap time, `.module` is the wrapped model, after `backward()` every rank's `.grad` holds the mean over ranks — but
owns the schedule: a captured training step (graphs.TrainGraph) leaves all parameter gradients in ONE contiguous fp32
buffer, which is reduced in place by NCCL over NVLink/NVSwitch in `bucket_cap_mb` slices (default: one collective for
the whole model — NVSwitch cost is latency + bytes, not per-link, so fewer, larger messages win) before the `.grad`
views are handed to autograd. No per-bucket copy-in/copy-out as in the stock reducer.

A model that runs eagerly (ENABLE_TRAIN_GRAPHS off, class-conditioned variants) is covered by post-accumulate hooks
that flatten -> reduce -> scatter the gradients once the last one has landed.

torch's own DistributedDataParallel also works around these models (its hooks fire when the backward graph has
replayed); it then reduces 25 MB buckets after, not during, the backward.
"""
import torch
import torch.distributed as dist
from torch import nn


def _mean_(flat, group, bucket_elems=None):
    """In-place mean over the ranks of `group`, optionally in slices of `bucket_elems`."""
    world = dist.get_world_size(group)
    if world == 1:
        return flat
    avg = dist.get_backend(group) == "nccl"
    step = flat.numel() if not bucket_elems else int(bucket_elems)
    for lo in range(0, flat.numel(), step):
        part = flat[lo:lo + step]
        if avg:
            dist.all_reduce(part, op=dist.ReduceOp.AVG, group=group)
        else:  # gloo (CPU tests) has no AVG
            dist.all_reduce(part, op=dist.ReduceOp.SUM, group=group)
            part.div_(world)
    return flat


class DataParallel(nn.Module):
    def __init__(self, module, device_ids=None, process_group=None, broadcast_buffers=False, bucket_cap_mb=None,
                 overlap=True, **unused):
        super().__init__()
        if not dist.is_initialized():
            raise RuntimeError("DataParallel needs an initialised torch.distributed process group")
        self.module = module
        self.group = process_group
        self.bucket_elems = None if not bucket_cap_mb else int(bucket_cap_mb * (1 << 20) // 4)
        self.broadcast_buffers = broadcast_buffers
        self._params = [p for p in module.parameters() if p.requires_grad]
        self._pending = 0
        self._graph_synced = False
        self._broadcast([p.data for p in module.parameters()] + [b.data for b in module.buffers()])
        # captured steps: graphs.TrainGraph calls this with the flat gradient buffer
        module.__dict__["_grad_sync"] = self._sync_flat
        # NCCL: let the captured backward graph reduce bucket by bucket while it is still producing gradients
        # (graphs._BucketReducer); 32 MB buckets unless bucket_cap_mb says otherwise
        module.__dict__.pop("_grad_overlap", None)
        if overlap and dist.get_backend(self.group) == "nccl" and dist.get_world_size(self.group) > 1:
            module.__dict__["_grad_overlap"] = (self.group, self.bucket_elems or (32 << 20) // 4)
        for p in self._params:
            p.register_post_accumulate_grad_hook(self._landed)

    def _broadcast(self, tensors):
        by_dtype = {}
        for t in tensors:
            by_dtype.setdefault(t.dtype, []).append(t)
        for ts in by_dtype.values():
            flat = torch._utils._flatten_dense_tensors(ts)
            dist.broadcast(flat, src=dist.get_global_rank(self.group, 0) if self.group is not None else 0,
                           group=self.group)
            for t, s in zip(ts, torch._utils._unflatten_dense_tensors(flat, ts)):
                t.copy_(s)

    def _sync_flat(self, flat, reduced=False):
        if not reduced:   # `reduced`: the backward graph already holds the collective
            _mean_(flat, self.group, self.bucket_elems)
        self._graph_synced = True

    def _landed(self, p):
        """Post-accumulate hook of every trainable parameter. The FIRST gradient of a backward pass queues one
        end-of-backward callback; the exchange happens there, so a parameter that receives no gradient in some step
        (an unused branch, a frozen-by-usage conditioning path) cannot leave a short count behind that would fire the
        collective in the middle of the next backward."""
        self._pending += 1
        if self._pending == 1:
            torch.autograd.Variable._execution_engine.queue_callback(self._backward_done)

    def _backward_done(self):
        self._pending = 0
        if self._graph_synced:  # this backward already reduced its flat buffer
            self._graph_synced = False
            return
        # every rank reduces the SAME layout: one slot per trainable parameter, zeros where this rank has no gradient
        ref = next((q.grad for q in self._params if q.grad is not None), None)
        if ref is None:
            return
        sizes = [q.numel() for q in self._params]
        flat = torch.zeros(sum(sizes), device=ref.device, dtype=torch.float32)
        views = list(flat.split(sizes))
        have = [(v, q.grad) for v, q in zip(views, self._params) if q.grad is not None]
        torch._foreach_copy_([v.view(g.shape) for v, g in have], [g for _, g in have])
        _mean_(flat, self.group, self.bucket_elems)
        for v, q in zip(views, self._params):
            if q.grad is not None:
                q.grad.copy_(v.view(q.grad.shape))

    def forward(self, *args, **kwargs):
        if self.broadcast_buffers and any(True for _ in self.module.buffers()):
            self._broadcast([b.data for b in self.module.buffers()])
        return self.module(*args, **kwargs)

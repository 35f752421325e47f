"""Host -> device input pipeline of the training loop.

The reference loop uploads each batch on the compute stream right before it is used
(train_ddpm_cond_celebhq_multi_gpu.py:308-339: `.to(device, non_blocking=True)` on latents, text embeddings and the
(B, 18, 512, 512) fp32 one-hot mask — 305 MB per step at batch 16), so the copy and the step run back to back.
DevicePrefetcher uploads batch i+1 on its own stream while step i computes:

    for im, cond_input in DevicePrefetcher(data_loader, device):      # instead of `for im, cond_input in data_loader`
        ...
"""
import torch


def _map(obj, fn):
    if isinstance(obj, torch.Tensor):
        return fn(obj)
    if isinstance(obj, dict):
        return {k: _map(v, fn) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_map(v, fn) for v in obj)
    return obj


class DevicePrefetcher:
    """Iterates `loader` (any iterable of tensors / tuples / lists / dicts of tensors; pin the host memory for the copy
    to be asynchronous), yielding the same structure with every tensor on `device`. One batch is in flight."""

    def __init__(self, loader, device):
        self.loader = loader
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device)

    def __len__(self):
        return len(self.loader)

    def _upload(self, batch):
        with torch.cuda.stream(self.stream):
            return _map(batch, lambda t: t.to(self.device, non_blocking=True))

    def __iter__(self):
        it = iter(self.loader)
        try:
            nxt = self._upload(next(it))
        except StopIteration:
            return
        while nxt is not None:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_stream(self.stream)
            batch = nxt
            _map(batch, lambda t: t.record_stream(cur) if t.is_cuda else None)
            try:
                nxt = self._upload(next(it))
            except StopIteration:
                nxt = None
            yield batch

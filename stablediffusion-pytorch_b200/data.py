"""Host -> device input pipeline of the training loop.

The reference loop uploads each batch on the compute stream right before it is used
(train_ddpm_cond_celebhq_multi_gpu.py:308-339: `.to(device, non_blocking=True)` on latents, text embeddings and the
(B, 18, 512, 512) fp32 one-hot mask — 305 MB per step at batch 16), so the copy and the step run back to back.
DevicePrefetcher uploads batch i+1 on its own stream while step i computes:

    for im, cond_input in DevicePrefetcher(data_loader, device):      # instead of `for im, cond_input in data_loader`
        ...

`shrink={"image": (32, 32)}` additionally stops uploading what the model never reads: the mask branch resizes the
(B, 18, 512, 512) one-hot mask to the latent size with F.interpolate(mode="nearest") (models/unet_cond_base.py:134-135),
which for an integer ratio selects exactly the pixels [::16, ::16]. The prefetcher gathers those pixels on the HOST into
a pinned (B, 18, 32, 32) buffer and uploads 1/256 of the bytes (295 KB instead of 75.5 MB per batch of 16 masks); the
model's own resize is then the identity and the result is bit-identical (tests/test_data_gpu.py).
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

    def __init__(self, loader, device, shrink=None):
        self.loader = loader
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device)
        self.shrink = dict(shrink or {})
        self._stage = {}      # (key, shape) -> [pinned buffers, events, next index]

    def __len__(self):
        return len(self.loader)

    def _gather(self, key, t):
        """Nearest-neighbour pre-selection of a (B, C, H, W) host tensor to `shrink[key]` into a pinned staging buffer
        (two per shape, guarded by the event of the copy that last read them)."""
        h, w = self.shrink[key]
        if t.is_cuda or t.dim() != 4 or t.shape[2] % h or t.shape[3] % w or (t.shape[2] == h and t.shape[3] == w):
            return t
        shape = (t.shape[0], t.shape[1], h, w)
        ent = self._stage.setdefault((key, shape, t.dtype), [[torch.empty(shape, dtype=t.dtype).pin_memory() for _ in range(2)],
                                                              [None, None], 0])
        i = ent[2]
        ent[2] = (i + 1) % 2
        if ent[1][i] is not None:
            ent[1][i].synchronize()
        ent[0][i].copy_(t[:, :, ::t.shape[2] // h, ::t.shape[3] // w])
        return ent[0][i], ent, i

    def _upload(self, batch):
        def up(obj, key=None):
            if isinstance(obj, torch.Tensor):
                if key in self.shrink:
                    g = self._gather(key, obj)
                    if isinstance(g, tuple):
                        buf, ent, i = g
                        dev = buf.to(self.device, non_blocking=True)
                        ent[1][i] = torch.cuda.Event()
                        ent[1][i].record(self.stream)
                        return dev
                return obj.to(self.device, non_blocking=True)
            if isinstance(obj, dict):
                return {k: up(v, k) for k, v in obj.items()}
            if isinstance(obj, (list, tuple)):
                return type(obj)(up(v, key) for v in obj)
            return obj
        with torch.cuda.stream(self.stream):
            return up(batch)

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

"""CUDA-graph plumbing: the per-timestep work of DDPM sampling is ~300 kernel launches of a few microseconds
each, so issuing them from Python costs more than running them. Both users of this file capture the launch
sequence once and replay it:

  GraphedForward     — inside `Unet.forward` under torch.no_grad(): the drop-in call the reference's sampling loops
                       make (tools/sample_ddpm_text_image_cond.py:82) replays a graph when shapes, condition state
                       and parameters are unchanged since capture
  SamplingPipeline   — the whole `for i in reversed(range(T))` loop body (model [+ unconditional model] ->
                       sample_prev_timestep -> i -= 1) as one graph with the loop counter on the device
"""
import torch

from . import engine as E
from . import bank as BK
from . import runtime as rt

ENABLE_FORWARD_GRAPHS = True
ENABLE_TRAIN_GRAPHS = True   # forward AND backward of a training step replay captured graphs (see TrainGraph)


import contextlib
import gc


@contextlib.contextmanager
def quiesced_capture():
    """Region in which a stream capture is safe from THIS process's own garbage. Root cause of the "capture invalidated"
    failures seen in long sequences (round 1: the batch-32 sampling capture inside a full test run; round 2: the DiT leg of
    bench.py after the training legs): captured graphs of earlier models sit in reference cycles (TrainGraph <-> autograd
    Function <-> model.__dict__), so they are freed by Python's CYCLIC collector — at whatever allocation happens to
    trigger it, including one in the middle of a later capture. Destroying a CUDAGraph releases its private memory pool
    (cudaFree), and a cudaFree while a stream is capturing invalidates the capture. So: collect before the capture, let
    the device settle, and keep the cyclic collector off until the capture has ended."""
    gc.collect()
    torch.cuda.synchronize()
    was = gc.isenabled()
    gc.disable()
    try:
        yield
    finally:
        if was:
            gc.enable()


def _capture_inference(body):
    """Capture `body()` (no autograd) into a fresh CUDA graph -> (graph, body's result, library launches captured).
    `thread_local` error mode: a CUDA call another thread of the process makes (bench.py's clock sampler thread, a
    DataLoader's pin-memory thread) must not invalidate this thread's capture."""
    graph = torch.cuda.CUDAGraph()
    before = rt.launch_count()
    with quiesced_capture():
        with torch.cuda.graph(graph, capture_error_mode="thread_local"), torch.no_grad():
            out = body()
    return graph, out, rt.launch_count() - before


def as_nchw(out):
    """A model's `run` returns either an NHWC Act (UNet) or the fp32 NCHW tensor itself (DiT un-patchifies)."""
    return out.to_nchw() if hasattr(out, "to_nchw") else out


def _param_stamp(params):
    return tuple((p.data_ptr(), p._version) for p in params)


def cond_key(cond_input):
    """Identity stamp of a condition dict: (name, weak reference, _version, shape) per tensor. Identity, not address:
    the caching allocator reuses the address of a freed same-shape tensor at _version 0, so (data_ptr, _version) alone
    would match the NEXT batch's conditions against the previous batch's hoisted state."""
    import weakref
    if cond_input is None:
        return None
    return tuple((k, weakref.ref(v), v._version, tuple(v.shape)) for k, v in sorted(cond_input.items())
                 if isinstance(v, torch.Tensor))


def cond_key_matches(key, cond_input):
    if key is None or cond_input is None:
        return key is None and cond_input is None
    cur = [(k, v) for k, v in sorted(cond_input.items()) if isinstance(v, torch.Tensor)]
    return len(cur) == len(key) and all(k == kk and ref() is v and v._version == ver and tuple(v.shape) == shp
                                        for (k, v), (kk, ref, ver, shp) in zip(cur, key))


def cond_key_alive(key):
    return key is None or all(ref() is not None for _, ref, _, _ in key)


class GraphedForward:
    """One captured forward of `model.run` for a fixed (shape, t-length, CondState)."""

    def __init__(self, model, x, tt, st):
        self.params = list(model.parameters())
        self.stamp = _param_stamp(self.params)
        self.st = st
        self.x = x.detach().float().clone()
        self.t = tt.clone()
        n, c, h, w = x.shape
        side = torch.cuda.Stream(device=x.device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(2):  # warm-up: packs weights, fills the K/V cache, sets kernel attributes
                model.run(self.x, self.t, st)
        torch.cuda.current_stream().wait_stream(side)
        self.graph, self.out, self.launches = _capture_inference(lambda: as_nchw(model.run(self.x, self.t, st)))

    def valid(self, st):
        return st is self.st and _param_stamp(self.params) == self.stamp

    def __call__(self, x, tt):
        self.x.copy_(x)
        self.t.copy_(tt)
        self.graph.replay()
        rt.count_external(self.launches)
        return self.out.clone()


class SamplingPipeline:
    """Device-resident ancestral sampling loop (reference tools/sample_ddpm_text_image_cond.py:79-98 and
    tools/sample_ddpm_vqvae.py:28-40): x_T -> x_0 -> VQVAE.decode, every timestep one graph replay, no host
    synchronisation and no per-step host<->device traffic. z is drawn in-kernel (Philox) unless `z_fn(i)`
    supplies it (parity tests)."""

    def __init__(self, model, scheduler, vae=None, use_graph=True):
        self.model, self.scheduler, self.vae, self.use_graph = model, scheduler, vae, use_graph
        self._plan = None

    def _build(self, xt, cond_input, uncond_input, cfg_scale, inject_z):
        m, dev = self.model, xt.device
        n = xt.shape[0]
        P = type("Plan", (), {})()
        P.key = self._key(xt, cf_scale_key=float(cfg_scale), inject_z=inject_z)
        P.cond_key, P.uncond_key = cond_key(cond_input), cond_key(uncond_input)
        P.xt = torch.empty_like(xt, dtype=torch.float32)
        P.x0 = torch.empty_like(P.xt)
        P.t = torch.zeros(n, device=dev, dtype=torch.int64)
        P.z = torch.empty_like(P.xt) if inject_z else None
        P.st = m.prepare_cond(cond_input, xt.shape, dev) if m.cond else E.CondState()
        P.st_u = None
        if m.cond and uncond_input is not None and cfg_scale > 1:
            P.st_u = m.prepare_cond(uncond_input, xt.shape, dev)
        for s in (P.st, P.st_u):
            if s is not None and s.ctx is not None:
                s.ctx.persistent = True
        tables = self.scheduler._tables(dev)[:4]
        # Philox key of the in-kernel noise, read from device memory by the captured launch: {seed, call index}.
        # sample() refreshes it before every run, so successive batches do not share their noise trajectories and a
        # changed scheduler.seed takes effect without re-capturing.
        P.rng = torch.zeros(2, device=dev, dtype=torch.int64)

        def body():
            eps = as_nchw(m.run(P.xt, P.t, P.st))
            eps_u = as_nchw(m.run(P.xt, P.t, P.st_u)) if P.st_u is not None else None
            rt.sched_prev_step(P.xt, eps, P.t[:1], tables, z=P.z, eps_uncond=eps_u, cfg_scale=cfg_scale,
                               seed=0, offset=0, x_prev=P.xt, x0=P.x0, rng_state=P.rng)
            rt.advance_t(P.t, -1)

        P.body = body
        P.graph = None
        if self.use_graph:
            P.xt.copy_(xt)
            P.t.fill_(self.scheduler.num_timesteps - 1)
            if inject_z:
                P.z.zero_()
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side), torch.no_grad():
                body()
                body()
            torch.cuda.current_stream().wait_stream(side)
            P.graph, _, P.launches = _capture_inference(body)
        return P

    def _key(self, xt, cf_scale_key, inject_z):
        return (tuple(xt.shape), cf_scale_key, inject_z, E.PARITY_MODE, _param_stamp(list(self.model.parameters())))

    def plan(self, xt, cond_input=None, uncond_input=None, cf_guidance_scale=1.0, inject_z=False):
        """The captured loop body for these shapes / conditions / weights / seed. The condition dicts are matched by the
        identity and _version of the tensors inside them (graphs.cond_key), never by id() of the dict or by address."""
        key = self._key(xt, float(cf_guidance_scale), inject_z)
        P = self._plan
        if P is None or P.key != key or not cond_key_matches(P.cond_key, cond_input) \
                or not cond_key_matches(P.uncond_key, uncond_input):
            self._plan = None   # release the old plan's graph and buffers before building the next
            self._plan = self._build(xt, cond_input, uncond_input, cf_guidance_scale, inject_z)
        return self._plan

    def step(self, P):
        """One timestep on the plan's resident state (xt, t)."""
        if P.graph is not None:
            P.graph.replay()
            rt.count_external(P.launches)
        else:
            with torch.no_grad():
                P.body()

    @torch.no_grad()
    def sample(self, xt, cond_input=None, uncond_input=None, cf_guidance_scale=1.0, t_start=None, num_steps=None,
               z_fn=None, decode=True, noise_call=None):
        """Run timesteps t_start, t_start-1, ... (num_steps of them, default down to 0) from `xt`; returns
        (x_final, x0_estimate, images or None). The in-kernel noise is keyed by (scheduler.seed, call index): every
        call draws a fresh trajectory, as the reference's torch.randn does; pass `noise_call` to name the index (the
        same seed and index reproduce a run bit for bit)."""
        T = self.scheduler.num_timesteps
        t_start = T - 1 if t_start is None else t_start
        num_steps = t_start + 1 if num_steps is None else num_steps
        assert 0 < num_steps <= t_start + 1
        P = self.plan(xt, cond_input, uncond_input, cf_guidance_scale, inject_z=z_fn is not None)
        P.xt.copy_(xt)
        P.t.fill_(t_start)
        self._calls = getattr(self, "_calls", 0) + 1
        P.rng.copy_(torch.tensor([int(self.scheduler.seed), self._calls if noise_call is None else int(noise_call)],
                                 dtype=torch.int64))
        for k in range(num_steps):
            if z_fn is not None:
                P.z.copy_(z_fn(t_start - k))
            self.step(P)
        ims = None
        if decode and self.vae is not None and t_start - num_steps + 1 == 0:
            ims = self.vae.decode(P.xt)
        return P.xt.clone(), P.x0.clone(), ims


class _ParamSwap:
    """Temporarily binds every `module._parameters[name]` slot of a model to a stand-in tensor (same storage, fresh
    autograd leaf). Capture differentiates the stand-ins, never the real Parameters: their AccumulateGrad nodes (which
    DistributedDataParallel keeps alive, tied to the stream it was built on) stay out of the captured region."""

    def __init__(self, core, params):
        index = {id(p): i for i, p in enumerate(params)}
        self.slots = [(m, k, index[id(p)]) for m in core.modules() for k, p in m._parameters.items() if p is not None]
        self.standins = [p.detach().requires_grad_(p.requires_grad) for p in params]

    def __enter__(self):
        self.saved = [m._parameters[k] for m, k, _ in self.slots]
        for m, k, i in self.slots:
            m._parameters[k] = self.standins[i]

    def __exit__(self, *exc):
        for (m, k, _), p in zip(self.slots, self.saved):
            m._parameters[k] = p


class _ReplayFn(torch.autograd.Function):
    """forward = replay of the captured forward graph; backward = replay of the captured backward graph. Parameters are
    arguments only so that autograd routes the gradients to them (the graphs read the parameter storage in place)."""

    @staticmethod
    def forward(ctx, tg, x, *params):
        ctx.tg = tg
        tg.generation = ctx.generation = getattr(tg, "generation", 0) + 1
        tg.x.copy_(x)
        tg.fwd_graph.replay()
        rt.count_external(tg.fwd_launches)
        return tg.out.detach().clone()

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        tg = ctx.tg
        if tg.generation != ctx.generation:
            # the backward graph differentiates the activations the LAST forward replay left in the graph's pool
            raise rt.B200Error("backward of a captured training step after another forward of the same model replayed: "
                               "run two in-flight forwards (f(x1) + f(x2), distillation losses) with "
                               "sdb200.graphs.ENABLE_TRAIN_GRAPHS = False")
        tg.gout.copy_(g)
        tg.bwd_graph.replay()
        rt.count_external(tg.bwd_launches)
        return (None, tg.dx.clone() if tg.dx is not None else None) + tg.take_grads()


def arrival_slots(numels, shapes, trainable, order):
    """Slot layout of the flat gradient buffer: (offset, numel, shape) per parameter (None when not trainable), slots
    64-element aligned and laid out in `order` (parameter indices; trainable ones not listed follow at the end). With
    order = the sequence in which a backward pass produces the gradients, every prefix of the buffer is complete before
    the rest, so the exchange can start on the first bucket while the backward pass is still producing the later ones —
    in PARAMETER order the parameters used at the very start of the forward pass (time-embedding bank, conditioning
    projections) are scattered over all buckets and hold every one of them open until the end of the backward pass
    (measured: profiles/r2e_overlap_n2.md)."""
    slots, off = [None] * len(numels), 0
    seen = set()
    for i in list(order) + list(range(len(numels))):
        if i in seen or not trainable[i]:
            continue
        seen.add(i)
        slots[i] = (off, numels[i], tuple(shapes[i]))
        off += (numels[i] + 63) // 64 * 64
    return slots, off


def plan_buckets(slots, fired, bucket_elems):
    """Buckets of parameter indices for the in-graph gradient exchange: the parameters that receive a gradient (`fired`),
    walked in slot order (= gradient-arrival order, see arrival_slots), closed whenever at least `bucket_elems` elements
    have been collected. Every bucket is a run of consecutive used slots, so it owns one contiguous slice [bucket_range]
    of the flat gradient buffer. slots[i] = (offset, numel, shape) or None."""
    buckets, cur, size = [], [], 0
    for i in sorted(set(fired), key=lambda i: slots[i][0]):
        cur.append(i)
        size += slots[i][1]
        if size >= bucket_elems:
            buckets.append(cur)
            cur, size = [], 0
    if cur:
        buckets.append(cur)
    return buckets


def bucket_range(slots, idxs):
    """[lo, hi) of the flat buffer covered by a bucket (slots are 64-element aligned)."""
    lo = min(slots[i][0] for i in idxs)
    hi = max(slots[i][0] + (slots[i][1] + 63) // 64 * 64 for i in idxs)
    return lo, hi


class _BucketReducer:
    """Gradient mean over ranks, captured inside the backward graph. The flat buffer (slots in gradient-arrival order,
    arrival_slots) is cut into contiguous buckets; a tensor hook on every stand-in parameter counts its
    bucket down, and the hook that completes a bucket (i) copies the bucket's gradients that were not written in place
    into their slots, (ii) forks a side stream off the capturing stream and issues ONE in-place NCCL all-reduce (AVG) on
    the bucket's slice there. `finish()` reduces whatever is still open and joins the side stream, so the graph ends
    with every slot holding the mean. Element-wise the result is identical to one all-reduce of the whole buffer."""

    def __init__(self, tg, dev):
        import torch.distributed as dist
        self.dist, self.tg = dist, tg
        self.group, bucket_elems = tg.overlap
        self.side = torch.cuda.Stream(device=dev)
        self.index_of = {i: k for k, i in enumerate([i for i, sl in enumerate(tg.slots) if sl is not None])}
        self.buckets = plan_buckets(tg.slots, tg._fired, bucket_elems)
        self.bucket_of = {i: b for b, idxs in enumerate(self.buckets) for i in idxs}
        self.pending = [len(b) for b in self.buckets]
        self.done = [False] * len(self.buckets)
        self.grads, self.copied, self.handles = {}, set(), []
        self._keep = []       # sources of copies issued on the side stream: alive until finish() has joined it

    def arm(self):
        for i, k in self.index_of.items():
            if i in self.bucket_of:
                self.handles.append(self.tg._standins[k].register_hook(lambda g, i=i: self._landed(i, g)))

    def disarm(self):
        for h in self.handles:
            h.remove()
        self.handles = []

    def _landed(self, i, g):
        self.grads[i] = g
        b = self.bucket_of[i]
        self.pending[b] -= 1
        if self.pending[b] == 0:
            self._flush(b)

    def _flush(self, b):
        tg, idxs = self.tg, self.buckets[b]
        views, srcs = [], []
        for i in idxs:
            g = self.grads.pop(i, None)
            if g is None:
                continue
            off, n, shape = tg.slots[i]
            v = tg.flat[off:off + n].view(shape)
            self.copied.add(i)
            if g.dtype == torch.float32 and g.data_ptr() == v.data_ptr():
                continue
            views.append(v)
            srcs.append(g if g.dtype == torch.float32 else g.float())
        lo, hi = bucket_range(tg.slots, idxs)
        main = torch.cuda.current_stream()
        self.side.wait_stream(main)          # fork: the bucket's producers are ordered before the collective —
        if rt.Side.FORKED:
            self.side.wait_stream(rt.Side.stream(tg.flat.device))   # — those on the parameter-gradient branch too
        with torch.cuda.stream(self.side):
            if views:                        # (bias sums and other gradients not written in place: they may come from
                torch._foreach_copy_(views, srcs)   # the parameter-gradient branch, so the copy rides this stream)
                self._keep.extend(srcs)
            self.dist.all_reduce(tg.flat[lo:min(hi, tg.flat.numel())], op=self.dist.ReduceOp.AVG, group=self.group)
        self.done[b] = True

    def finish(self):
        for b in range(len(self.buckets)):
            if not self.done[b]:
                for i in self.buckets[b]:   # gradients whose hook did not fire during capture were copied by the caller
                    self.grads.pop(i, None)
                self._flush(b)
        torch.cuda.current_stream().wait_stream(self.side)   # join
        self._keep.clear()


# Version of "the module trees of this process": bumped whenever ANY module registers a sub-module, a parameter or a
# buffer (module surgery of cim_qn_train/progressive_qn_train.py:576-651, `layer.weight = nn.Parameter(...)`, ...). A
# captured step re-walks its model's tree only when this moved since it last looked.
STRUCTURE = [0]


def _bump_structure(*_a, **_k):
    STRUCTURE[0] += 1


torch.nn.modules.module.register_module_module_registration_hook(_bump_structure)
torch.nn.modules.module.register_module_parameter_registration_hook(_bump_structure)
torch.nn.modules.module.register_module_buffer_registration_hook(_bump_structure)


class TrainGraph:
    """One training step's forward and backward as two CUDA graphs: ~1200 kernel launches + ~1000 small torch ops per
    step cost ~55 ms of host time when issued one by one, several times the device time at the reference's per-GPU
    batch of 16 (train_ddpm_cond_celebhq_multi_gpu.py:341-362). Parameters are read in place on every replay (weight
    packing is inside the graph), so optimizer steps need no re-capture.

    The backward graph ends by gathering every parameter gradient into ONE flat fp32 buffer (`flat`, 256-byte aligned
    slots in parameter order). Each backward hands out views of a fresh copy of it, so `.grad` accumulation,
    in-place clipping and DistributedDataParallel's bucket copies see ordinary tensors; `grad_sync`, when set, is called
    with that flat copy before the views are handed out (one collective for the whole model, see parallel.py)."""

    WARMUP = 2

    def __init__(self, core, x, tt, cond_tensors, has_text, has_mask, overlap=None):
        """overlap = (process_group, bucket_elems) or None: when given (parallel.DataParallel on NCCL), the gradient mean
        over ranks is captured INSIDE the backward graph, bucket by bucket on a side stream, as soon as the last
        gradient of a bucket has been produced (the reference's DDP reducer does the same from autograd hooks,
        train_ddpm_cond_celebhq_multi_gpu.py:257-263,362) — in place on slices of the flat buffer, no bucket copies."""
        self.params = list(core.parameters())
        self.structure = STRUCTURE[0]
        self.key = self.make_key(core, x, tt, cond_tensors, overlap)
        self.grad_sync = None
        self.overlap = overlap
        self.reduced_in_graph = False
        self.n_buckets = 0
        dev = x.device
        swap = _ParamSwap(core, self.params)
        self.x = x.detach().clone().requires_grad_(x.requires_grad)
        self.tt = tt.clone()
        self.cond = [c.detach().clone() for c in cond_tensors]
        wrt = [t for t in [self.x] + swap.standins if t.requires_grad]
        if not wrt:
            raise rt.B200Error("training graph requested but nothing requires grad")

        self.bank = None

        def forward():
            if self.bank is not None:
                self.bank.refresh()
            ci, k = {}, 0
            if has_text:
                ci["text"], k = self.cond[k], k + 1
            if has_mask:
                ci["image"] = self.cond[k]
            with swap:
                core._weight_side(before_cond=True)     # all quantised leaves' weights: one launch per phase
                st = core.prepare_cond(ci, self.x.shape, dev) if core.cond else E.CondState()
                return as_nchw(core.run(self.x, self.tt, st))

        # slots of the flat gradient buffer: parameter order, or — when the gradient exchange is captured inside the
        # backward graph — the order in which a backward pass produces the gradients (one eager discovery pass)
        order = self._arrival_order(forward, wrt) if overlap is not None else []
        self.slots, off = arrival_slots([p.numel() for p in self.params], [p.shape for p in self.params],
                                        [p.requires_grad for p in self.params], order)
        self.flat = torch.zeros(max(off, 64), device=dev, dtype=torch.float32)
        # packed bf16 forms of every parameter in one launch per step; weight gradients written straight into `flat`
        self.bank = BK.WeightBank(swap.standins, self.flat, self.slots)
        BK.ACTIVE = self.bank
        try:
            self._build(forward, wrt, dev)
        finally:
            BK.ACTIVE = None

    def _arrival_order(self, forward, wrt):
        """Parameter indices in the order one (eager, bank-less) backward pass produces their gradients."""
        idx = [i for i, p in enumerate(self.params) if p.requires_grad]
        standins = [t for t in wrt if t is not self.x]
        fired = []
        with torch.autocast("cuda", enabled=False), torch.enable_grad():
            out = forward()
            hs = [t.register_hook(lambda g, i=i: fired.append(i)) for i, t in zip(idx, standins)]
            torch.autograd.grad((out,), wrt, (torch.ones_like(out),), allow_unused=True)
            for h in hs:
                h.remove()
        return fired

    def _build(self, forward, wrt, dev):
        rt.Side.DEFER = True     # parameter-gradient branches join at the end of a backward pass (runtime.Side)
        try:
            self._build_deferred(forward, wrt, dev)
        finally:
            rt.Side.DEFER = False
            rt.Side._HELD.clear()

    def _build_deferred(self, forward, wrt, dev):
        with torch.autocast("cuda", enabled=False), torch.enable_grad(), rt.training_pass():
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                self._fired = []   # parameter indices in the order their gradients are produced (last warm-up pass)
                self._standins = [t for t in wrt if t is not self.x]
                for _ in range(self.WARMUP):  # lazy kernel attributes, allocator warm-up, weight-bank discovery
                    self.bank.begin_pass()
                    out = forward()
                    fired, hs = [], []
                    if self.overlap is not None:
                        idx = [i for i, sl in enumerate(self.slots) if sl is not None]
                        for i, t in zip(idx, self._standins):
                            hs.append(t.register_hook(lambda g, i=i: fired.append(i)))
                    torch.autograd.grad((out,), wrt, (torch.ones_like(out),), allow_unused=True)
                    rt.Side.join_all(dev)
                    for h in hs:
                        h.remove()
                    self._fired = fired
                    self.bank.finalize()
                del out
            self.bank.begin_pass()
            torch.cuda.current_stream().wait_stream(side)
            pool = torch.cuda.graph_pool_handle()
            self.fwd_graph, self.bwd_graph = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
            before = rt.launch_count()
            with quiesced_capture(), torch.cuda.graph(self.fwd_graph, pool=pool, capture_error_mode="thread_local"):
                self.out = forward()
            self.fwd_launches = rt.launch_count() - before
            self.gout = torch.empty_like(self.out)
            before = rt.launch_count()
            red = _BucketReducer(self, dev) if self.overlap is not None and self._fired else None
            with quiesced_capture(), torch.cuda.graph(self.bwd_graph, pool=pool, capture_error_mode="thread_local"):
                if red is not None:
                    red.arm()
                grads = list(torch.autograd.grad((self.out,), wrt, (self.gout,), allow_unused=True))
                rt.Side.join_all(dev)      # the deferred parameter-gradient branches
                if red is not None:
                    red.disarm()
                self.dx = grads.pop(0) if self.x.requires_grad else None
                views, srcs, self.unused = [], [], set()
                it = iter(grads)
                for i, sl in enumerate(self.slots):
                    if sl is None:
                        continue
                    g = next(it)
                    if g is None:
                        self.unused.add(i)
                        continue
                    if red is not None and i in red.copied:
                        continue   # already in the flat buffer (and, with its bucket, on its way through NCCL)
                    v = self.flat[sl[0]:sl[0] + sl[1]].view(sl[2])
                    if g.dtype == torch.float32 and g.data_ptr() == v.data_ptr():
                        continue   # the weight-gradient kernel wrote it in place
                    views.append(v)
                    srcs.append(g if g.dtype == torch.float32 else g.float())
                if views:
                    torch._foreach_copy_(views, srcs)
                if red is not None:
                    red.finish()   # buckets still open (late or unhooked gradients), then join the side stream
                    self.reduced_in_graph = True
                    self.n_buckets = len(red.buckets)
            self.bwd_launches = rt.launch_count() - before
        self.launches = self.fwd_launches + self.bwd_launches

    def take_grads(self):
        flat = self.flat.clone()
        if self.grad_sync is not None:
            self.grad_sync(flat, reduced=self.reduced_in_graph)
        # one split over the slots in buffer order (one C++ call) instead of a slice + view per parameter: this runs
        # on the host once per step, between the loss read-back and the next replay (tools/host_profile_train.py)
        plan = self.__dict__.get("_grad_plan")
        if plan is None:
            live = sorted((sl[0], i) for i, sl in enumerate(self.slots) if sl is not None and i not in self.unused)
            sizes, owner, pos = [], [], 0
            for off, i in live:
                if off > pos:
                    sizes.append(off - pos), owner.append(None)      # alignment gap / unused slot
                sizes.append(self.slots[i][1]), owner.append(i)
                pos = off + self.slots[i][1]
            if pos < self.flat.numel():
                sizes.append(self.flat.numel() - pos), owner.append(None)
            plan = self.__dict__["_grad_plan"] = (sizes, owner)
        out = [None] * len(self.slots)
        for piece, i in zip(flat.split_with_sizes(plan[0]), plan[1]):
            if i is not None:
                out[i] = piece.view(self.slots[i][2])
        return tuple(out)

    @staticmethod
    def make_key(core, x, tt, cond_tensors, overlap=None, params=None):
        """Everything a captured step depends on. `params`: the parameter list of an existing graph — used while the
        module tree is known to be unchanged (STRUCTURE version, bumped by torch's global module / parameter
        registration hooks), which saves walking the tree (1.8 ms of host time per step for the 400-parameter UNet)."""
        ps = core.parameters() if params is None else params
        return (tuple(x.shape), x.requires_grad, tuple(tt.shape), None if overlap is None else (id(overlap[0]), overlap[1])) \
            + tuple((tuple(c.shape), c.dtype) for c in cond_tensors) \
            + tuple((p.data_ptr(), p.requires_grad) for p in ps)

    def still_valid(self, core, x, tt, cond_tensors, overlap=None):
        if self.structure != STRUCTURE[0]:
            if self.key != self.make_key(core, x, tt, cond_tensors, overlap):     # full walk: the tree may have changed
                return False
            self.structure = STRUCTURE[0]
            return True
        return self.key == self.make_key(core, x, tt, cond_tensors, overlap, params=self.params)

    def __call__(self, x, tt, cond_tensors):
        self.tt.copy_(tt)
        for s, c in zip(self.cond, cond_tensors):
            s.copy_(c)
        return _ReplayFn.apply(self, x, *self.params)

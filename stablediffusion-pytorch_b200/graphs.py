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
from . import runtime as rt

ENABLE_FORWARD_GRAPHS = True
ENABLE_TRAIN_GRAPHS = True   # forward AND backward of a training step replay captured graphs (see TrainGraph)


def _param_stamp(params):
    return tuple((p.data_ptr(), p._version) for p in params)


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
        self.graph = torch.cuda.CUDAGraph()
        before = rt.launch_count()
        with torch.cuda.graph(self.graph), torch.no_grad():
            self.out = model.run(self.x, self.t, st).to_nchw()
        self.launches = rt.launch_count() - before

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
        P.key = (tuple(xt.shape), id(cond_input), id(uncond_input), float(cfg_scale), inject_z,
                 _param_stamp(list(m.parameters())))
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
        seed = self.scheduler.seed

        def body():
            eps = m.run(P.xt, P.t, P.st).to_nchw()
            eps_u = m.run(P.xt, P.t, P.st_u).to_nchw() if P.st_u is not None else None
            rt.sched_prev_step(P.xt, eps, P.t[:1], tables, z=P.z, eps_uncond=eps_u, cfg_scale=cfg_scale,
                               seed=seed, offset=0, x_prev=P.xt, x0=P.x0)
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
            P.graph = torch.cuda.CUDAGraph()
            before = rt.launch_count()
            with torch.cuda.graph(P.graph), torch.no_grad():
                body()
            P.launches = rt.launch_count() - before
        return P

    def plan(self, xt, cond_input=None, uncond_input=None, cf_guidance_scale=1.0, inject_z=False):
        key = (tuple(xt.shape), id(cond_input), id(uncond_input), float(cf_guidance_scale), inject_z,
               _param_stamp(list(self.model.parameters())))
        if self._plan is None or self._plan.key != key:
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
               z_fn=None, decode=True):
        """Run timesteps t_start, t_start-1, ... (num_steps of them, default down to 0) from `xt`; returns
        (x_final, x0_estimate, images or None)."""
        T = self.scheduler.num_timesteps
        t_start = T - 1 if t_start is None else t_start
        num_steps = t_start + 1 if num_steps is None else num_steps
        assert 0 < num_steps <= t_start + 1
        P = self.plan(xt, cond_input, uncond_input, cf_guidance_scale, inject_z=z_fn is not None)
        P.xt.copy_(xt)
        P.t.fill_(t_start)
        for k in range(num_steps):
            if z_fn is not None:
                P.z.copy_(z_fn(t_start - k))
            self.step(P)
        ims = None
        if decode and self.vae is not None and t_start - num_steps + 1 == 0:
            ims = self.vae.decode(P.xt)
        return P.xt.clone(), P.x0.clone(), ims


class _TensorArgs(torch.nn.Module):
    """Tensor-only calling convention over Unet.run for torch.cuda.make_graphed_callables (which needs an nn.Module
    whose parameters it can differentiate and whose arguments are all tensors). The wrapped model is kept out of
    this module's registered children on purpose (it is reached through a plain attribute)."""

    def __init__(self, core, has_text, has_mask):
        super().__init__()
        object.__setattr__(self, "_core", core)
        self.has_text, self.has_mask = has_text, has_mask
        # same Parameter objects, registered here so make_graphed_callables sees them
        for i, p in enumerate(core.parameters()):
            self.register_parameter(f"p{i}", p)

    def forward(self, x, tt, *cond):
        core = self._core
        ci = {}
        k = 0
        if self.has_text:
            ci["text"] = cond[k]
            k += 1
        if self.has_mask:
            ci["image"] = cond[k]
        st = core.prepare_cond(ci, x.shape, x.device) if core.cond else E.CondState()
        return core.run(x, tt, st).to_nchw()


class TrainGraph:
    """One training step's forward and backward as two CUDA graphs (torch.cuda.make_graphed_callables over the
    model's kernel sequence): ~1200 kernel launches + ~1000 small torch ops per step cost ~55 ms of host time when
    issued one by one, several times the device time at the reference's per-GPU batch of 16. Parameters are read in
    place on every replay (weight packing is inside the graph), so optimizer steps need no re-capture. Parameter
    gradients surface when the backward graph finishes; DistributedDataParallel's hooks fire then."""

    def __init__(self, core, x, tt, cond_tensors, has_text, has_mask):
        self.params = list(core.parameters())
        self.key = self.make_key(core, x, tt, cond_tensors)
        mod = _TensorArgs(core, has_text, has_mask)
        sample = (x.detach().clone().requires_grad_(x.requires_grad), tt.clone()) + tuple(
            c.detach().clone() for c in cond_tensors)
        before = rt.launch_count()
        with torch.autocast("cuda", enabled=False):
            self.fn = torch.cuda.make_graphed_callables(mod, sample, allow_unused_input=True)
        # make_graphed_callables runs 3 warm-up forward+backward passes and captures one more
        self.launches = (rt.launch_count() - before) // 4

    @staticmethod
    def make_key(core, x, tt, cond_tensors):
        return (tuple(x.shape), x.requires_grad, tuple(tt.shape)) + tuple((tuple(c.shape), c.dtype) for c in cond_tensors) \
            + tuple((p.data_ptr(), p.requires_grad) for p in core.parameters())

    def __call__(self, x, tt, cond_tensors):
        rt.count_external(self.launches)   # library kernels replayed by the forward graph and, later, the backward graph
        return self.fn(x, tt, *cond_tensors)

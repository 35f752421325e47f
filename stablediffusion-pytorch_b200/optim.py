"""Optimizer side of the training step.

    from sdb200.optim import FusedAdam
    opt = FusedAdam(model.parameters(), lr=cfg.lr, max_grad_norm=1.0, ema_params=ema_model.parameters(), ema_decay=0.9999)
    ...
    loss.backward()
    opt.step()            # clip + finite check + Adam + EMA: two kernel launches
    grad_norm = opt.grad_norm   # 0-dim device tensor (pre-clip total norm), no host sync

replaces this block of the reference's trainers (train_ddpm_cond_celebhq_multi_gpu.py:362-378, Model_DiT_9L_train.py
same shape):

    grad_norm = torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
    if not torch.isfinite(grad_norm): optimizer.zero_grad(); continue      # host sync
    optimizer.step()                                                       # torch.optim.Adam
    for ema_param, param in zip(ema_model.parameters(), model.parameters()):
        ema_param.data.mul_(EMA_DECAY).add_(param.data, alpha=1 - EMA_DECAY)

It is a torch.optim.Optimizer: param_groups / lr schedulers / GradScaler.step / state_dict work as with torch.optim.Adam,
and the state_dict is interchangeable with torch.optim.Adam's (per-parameter `step`, `exp_avg`, `exp_avg_sq`).
Differences, on purpose: the clipped gradient is consumed, not written back to `.grad`; a non-finite gradient norm skips
the update on the device instead of through a host round trip.
"""
import torch

from . import runtime as rt


def _align64(n):
    return (n + 63) // 64 * 64


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, max_grad_norm=None,
                 skip_nonfinite=None, ema_params=None, ema_decay=0.9999):
        if lr < 0 or eps < 0 or not 0 <= betas[0] < 1 or not 0 <= betas[1] < 1 or weight_decay < 0:
            raise ValueError("FusedAdam: invalid hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay))
        self.max_grad_norm = max_grad_norm
        self.skip_nonfinite = (max_grad_norm is not None) if skip_nonfinite is None else bool(skip_nonfinite)
        self.ema_decay = float(ema_decay)
        self._all = [p for g in self.param_groups for p in g["params"]]
        if not self._all:
            raise ValueError("FusedAdam: no parameters")
        for p in self._all:
            if p.dtype != torch.float32 or not p.is_cuda or not p.is_contiguous():
                raise rt.B200Error("FusedAdam: parameters must be contiguous fp32 CUDA tensors")
        self._ema = None
        if ema_params is not None:
            self._ema = list(ema_params)
            if len(self._ema) != len(self._all) or any(e.shape != p.shape or e.dtype != torch.float32 or not e.is_cuda
                                                         or not e.is_contiguous()
                                                         for e, p in zip(self._ema, self._all)):
                raise rt.B200Error("FusedAdam: ema_params must mirror the parameters (shape, fp32, CUDA, contiguous)")
        dev = self._all[0].device
        self._dev = dev
        self._off, off = [], 0
        for p in self._all:
            self._off.append(off)
            off += _align64(p.numel())
        self._total = max(off, 64)
        self._m = torch.zeros(self._total, device=dev)
        self._v = torch.zeros(self._total, device=dev)
        self._g = None                                   # own flat gradient buffer, only if the grads are scattered
        self._scalars = torch.zeros(4, device=dev)       # norm, clip coefficient, finite flag, step
        self._plans = {}
        for p, o in zip(self._all, self._off):
            self.state[p] = {"step": torch.zeros((), dtype=torch.float32),
                             "exp_avg": self._m[o:o + p.numel()].view(p.shape),
                             "exp_avg_sq": self._v[o:o + p.numel()].view(p.shape)}

    # ------------------------------------------------------------------------------------------ bookkeeping
    @property
    def grad_norm(self):
        """Total gradient norm seen by the last step() (before clipping), 0-dim device tensor."""
        return self._scalars[0]

    @property
    def last_step_applied(self):
        """1.0 if the last step() updated the parameters, 0.0 if it was skipped (non-finite gradients)."""
        return self._scalars[2]

    def _table(self, present, groups, goffs):
        key = (present, goffs, tuple(p.data_ptr() for p in self._all), None if self._ema is None else
               tuple(e.data_ptr() for e in self._ema))
        plan = self._plans.get(key)
        if plan is not None:
            return plan
        ce = rt.opt_chunk_elems()
        plan = []
        index = {id(p): i for i, p in enumerate(self._all)}
        for g in groups:
            rows = []
            for p in g["params"]:
                i = index[id(p)]
                if not present[i]:
                    continue
                e = None if self._ema is None else self._ema[i]
                vec = 1 if p.data_ptr() % 16 == 0 and (e is None or e.data_ptr() % 16 == 0) else 0
                for c0 in range(0, p.numel(), ce):
                    rows.append((p.data_ptr() + 4 * c0, 0 if e is None else e.data_ptr() + 4 * c0, self._off[i] + c0,
                                 goffs[i] + c0, min(ce, p.numel() - c0), vec))
            if not rows:
                plan.append(None)
                continue
            arr = (rt.OptChunk * len(rows))()
            for k, (pp, ee, oo, go, nn, vv) in enumerate(rows):
                arr[k].p, arr[k].ema, arr[k].off, arr[k].goff, arr[k].n, arr[k].vec = pp, ee or None, oo, go, nn, vv
            tab = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(self._dev)
            plan.append((tab, len(rows)))
        # the norm runs over every group's chunks at once
        alltab = [t for t in plan if t is not None]
        if len(alltab) == 1:
            norm_tab = alltab[0]
        else:
            norm_tab = (torch.cat([t[0] for t in alltab]), sum(t[1] for t in alltab))
        partial = torch.empty(norm_tab[1], device=self._dev)
        self._plans.clear()       # pointers changed: older plans are stale
        self._plans[key] = (plan, norm_tab, partial)
        return self._plans[key]

    def _grad_base(self, present):
        """(address of a flat fp32 buffer, per-parameter element offsets into it) of the present gradients: the
        gradients' own storage when they are views of ONE buffer (graphs.TrainGraph hands them out that way, in its own
        slot order — gradient-arrival order when the data-parallel exchange overlaps the backward pass), otherwise a
        private buffer in this optimizer's slot order filled by one multi-tensor copy."""
        # fast path: the same layout as last step (a captured training step hands out views of a fresh clone of its
        # flat buffer every time: new base address, same offsets) — one data_ptr() per parameter instead of storage /
        # dtype / contiguity queries (0.6 ms of host time per step for 400 parameters, tools/host_profile_train.py)
        last = self.__dict__.get("_last_layout")
        if last is not None and last[0] == present:
            goffs, first = last[1], last[2]
            base = self._all[first].grad.data_ptr() - 4 * goffs[first]
            if base % 16 == 0 and all(p.grad.data_ptr() == base + 4 * go
                                      for p, go in zip(self._all, goffs) if go is not None):
                return base, goffs
        store, ok = None, True
        goffs = [None] * len(self._all)
        for i, (p, has) in enumerate(zip(self._all, present)):
            if not has:
                continue
            g = p.grad
            if g.dtype != torch.float32 or not g.is_contiguous() or g.is_sparse:
                ok = False
                break
            st = g.untyped_storage().data_ptr()
            d = g.data_ptr() - st
            if store is None:
                store = st
            if st != store or store % 16 or d % 16:
                ok = False
                break
            goffs[i] = d // 4
        if ok and store is not None:
            goffs = tuple(goffs)
            self.__dict__["_last_layout"] = (present, goffs, next(i for i, go in enumerate(goffs) if go is not None))
            return store, goffs
        self.__dict__.pop("_last_layout", None)
        if self._g is None:
            self._g = torch.zeros(self._total, device=self._dev)
        dst, src = [], []
        for p, o, has in zip(self._all, self._off, present):
            if has:
                dst.append(self._g[o:o + p.numel()].view(p.shape))
                src.append(p.grad if p.grad.dtype == torch.float32 else p.grad.float())
        torch._foreach_copy_(dst, src)
        return self._g.data_ptr(), tuple(o if has else None for o, has in zip(self._off, present))

    # ------------------------------------------------------------------------------------------ step
    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        present = tuple(p.grad is not None for p in self._all)
        if not any(present):
            return loss
        gbase, goffs = self._grad_base(present)
        plan, (norm_tab, n_norm), partial = self._table(present, self.param_groups, goffs)
        rt.grad_sqnorm(norm_tab, n_norm, gbase, self.max_grad_norm or 0.0, self.skip_nonfinite, partial, self._scalars)
        for g, ent in zip(self.param_groups, plan):
            if ent is None:
                continue
            tab, n = ent
            rt.adam_ema_step(tab, n, gbase, self._m, self._v, self._scalars, g["lr"], g["betas"][0], g["betas"][1],
                             g["eps"], g["weight_decay"], self.ema_decay)
        self._bump_versions(present)
        return loss

    def _bump_versions(self, present):
        """The update kernels write parameters and EMA shadows through raw pointers, which autograd's version counters
        do not see. Every inference-side cache (packed bf16 weights, captured forward graphs, hoisted conditioning) is
        stamped with (data_ptr, _version): bump the counters so an in-process validation / sampling pass after a
        training step re-derives them instead of silently running on the weights packed before the step."""
        written = [p for p, has in zip(self._all, present) if has]
        if self._ema is not None:
            written += [e for e, has in zip(self._ema, present) if has]
        torch.autograd.graph.increment_version(written)

    # ------------------------------------------------------------------------------------------ checkpoints
    def state_dict(self):
        step = self._scalars[3].item()
        for p in self._all:
            self.state[p]["step"] = torch.tensor(step, dtype=torch.float32)
        return super().state_dict()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._all = [p for g in self.param_groups for p in g["params"]]
        step = 0.0
        for p, o in zip(self._all, self._off):
            st = self.state.get(p, {})
            n = p.numel()
            if "exp_avg" in st:
                self._m[o:o + n].copy_(st["exp_avg"].reshape(-1))
                self._v[o:o + n].copy_(st["exp_avg_sq"].reshape(-1))
                step = max(step, float(st.get("step", 0.0)))
            self.state[p] = {"step": torch.tensor(step, dtype=torch.float32),
                             "exp_avg": self._m[o:o + n].view(p.shape), "exp_avg_sq": self._v[o:o + n].view(p.shape)}
        self._scalars[3] = step
        self._plans.clear()


class PercentOptimizerFP(FusedAdam):
    """Drop-in for the CIM trainers' `PercentOptimizerFP(params, lr, betas, eps)` (cim_layers/IBA_optimizer.py:80-143):
    Adam moments, then each tensor's update is rescaled so that its largest element equals `lr` times the tensor's
    largest weight (and never less than plain `lr`). The reference walks the parameters in Python (~14 tensor ops and
    two reductions per parameter); here the step is two table-driven launches (`b200sd_percent_opt_step`) over the same
    flat moment buffers FusedAdam uses, in the reference's fp32 operation order with correctly rounded arithmetic
    (moments bit-identical, parameters within 1e-6 relative of torch's CPU run, whose MKL sqrt is not correctly rounded). `state_dict()` has the reference's per-parameter `step`, `exp_avg`, `exp_avg_sq`."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=0.0, max_grad_norm=None,
                         skip_nonfinite=False)
        self._steps = [0] * len(self._all)      # per parameter, as the reference counts them (:101-105)
        self._ptabs = {}

    def _ptable(self, idxs, goffs):
        """(chunk table, tensor index per chunk, chunk count, per-tensor max buffer) for the parameters `idxs`."""
        key = (tuple(idxs), tuple(goffs[i] for i in idxs), tuple(self._all[i].data_ptr() for i in idxs))
        ent = self._ptabs.get(key)
        if ent is None:
            ce = rt.opt_chunk_elems()
            rows, tid = [], []
            for k, i in enumerate(idxs):
                p = self._all[i]
                for c0 in range(0, p.numel(), ce):
                    rows.append((p.data_ptr() + 4 * c0, self._off[i] + c0, goffs[i] + c0, min(ce, p.numel() - c0)))
                    tid.append(k)
            arr = (rt.OptChunk * len(rows))()
            for k, (pp, oo, go, nn) in enumerate(rows):
                arr[k].p, arr[k].ema, arr[k].off, arr[k].goff, arr[k].n, arr[k].vec = pp, None, oo, go, nn, 0
            tab = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(self._dev)
            ent = (tab, torch.tensor(tid, dtype=torch.int32, device=self._dev), len(rows),
                   torch.zeros(len(idxs), 2, dtype=torch.int32, device=self._dev))
            self._ptabs.clear()
            self._ptabs[key] = ent
        return ent

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        present = tuple(p.grad is not None for p in self._all)
        if not any(present):
            return loss
        gbase, goffs = self._grad_base(present)
        index = {id(p): i for i, p in enumerate(self._all)}
        for g in self.param_groups:
            lr, (b1, b2), eps = g["lr"], g["betas"], g["eps"]
            by_step = {}
            for p in g["params"]:
                i = index[id(p)]
                if present[i]:
                    self._steps[i] += 1
                    by_step.setdefault(self._steps[i], []).append(i)
            for step, idxs in sorted(by_step.items()):   # one entry unless some parameters skipped earlier steps
                bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step              # :113-114, in Python doubles as there
                adj, bc2s = (bc2 ** 0.5) / bc1, bc2 ** 0.5             # :117, :120
                tab, tid, n, tmax = self._ptable(idxs, goffs)
                tmax.zero_()
                rt.percent_opt_step(tab, tid, n, gbase, self._m, self._v, tmax, lr, b1, b2, 1 - b1, 1 - b2, adj, bc2s,
                                    eps)
        self._bump_versions(present)
        return loss

    def state_dict(self):
        for p, n in zip(self._all, self._steps):
            self.state[p]["step"] = n
        return torch.optim.Optimizer.state_dict(self)

    def load_state_dict(self, state_dict):
        torch.optim.Optimizer.load_state_dict(self, state_dict)
        self._all = [p for g in self.param_groups for p in g["params"]]
        for k, (p, o) in enumerate(zip(self._all, self._off)):
            st = self.state.get(p, {})
            n = p.numel()
            if "exp_avg" in st:
                self._m[o:o + n].copy_(st["exp_avg"].reshape(-1))
                self._v[o:o + n].copy_(st["exp_avg_sq"].reshape(-1))
            self._steps[k] = int(st.get("step", 0))
            self.state[p] = {"step": self._steps[k], "exp_avg": self._m[o:o + n].view(p.shape),
                             "exp_avg_sq": self._v[o:o + n].view(p.shape)}
        self._ptabs.clear()


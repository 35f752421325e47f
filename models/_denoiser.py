"""Forward dispatch shared by the drop-in denoisers (Unet variants, DIT): the reference call
`model(x, t, cond_input)` lands here and is routed to
  * a replay of the captured training step (autograd recording; graphs.TrainGraph),
  * a replay of the captured inference forward (torch.no_grad; graphs.GraphedForward) with the (x, t)-independent
    conditioning hoisted and cached, or
  * the eager kernel sequence (inside someone else's capture, class-conditioned variants, graphs disabled).
Sub-classes provide `run(x, tt, st)`, `prepare_cond(cond_input, x_shape, device)`, `_cond_params()` and the flags
`cond`, `text_cond`, `image_cond`, `class_cond`."""
import torch
import torch.nn as nn

import sdb200  # noqa: F401
from sdb200 import engine as E
from sdb200 import graphs as G
from sdb200 import runtime as rt


def validate_image_conditional_input(cond_input, x):
    assert 'image' in cond_input, \
        "Model initialized with image conditioning but cond_input has no image information"
    assert cond_input['image'].shape[0] == x.shape[0], \
        "Batch size mismatch of image condition and input"
    assert cond_input['image'].shape[2] % x.shape[2] == 0, \
        "Height/Width of image condition must be divisible by latent input"


class Denoiser(nn.Module):
    def _cached_cond(self, cond_input, x):
        """In inference (no grad) reuse the CondState while the caller keeps passing the same, unmodified condition
        tensors and the conditioning weights are untouched — which is what every sampling loop does."""
        if isinstance(cond_input, E.CondState):
            return cond_input
        if torch.is_grad_enabled() or not E.HOIST_CONDITIONING:
            return self.prepare_cond(cond_input, x.shape, x.device)
        # A hit needs the SAME tensor objects (weak references: a freed tensor can never match, and nothing is kept
        # alive), unmodified since the entry was made (_version), and unchanged conditioning weights. Addresses are
        # not part of the key: the caching allocator hands the next batch's same-shape tensor the address of the
        # previous one (batch_condition_image_generator.py builds its conditions with torch.cat(...).to(device) per batch).
        wstamp = tuple((p.data_ptr(), p._version) for p in self._cond_params())
        cache = self.__dict__.setdefault("_cond_cache", [])
        for ent in cache:
            if ent[0] == (tuple(x.shape), E.PARITY_MODE) and ent[2] == wstamp and G.cond_key_matches(ent[1], cond_input):
                return ent[3]
        st = self.prepare_cond(cond_input, x.shape, x.device)
        if st.ctx is not None:
            st.ctx.persistent = True
        # A state whose condition tensors are gone (the previous sampling job's) lends its storage to the new
        # condition: the forward graphs captured on it keep working (engine.rebind_cond).
        for k, ent in enumerate(cache):
            if ent[0] == (tuple(x.shape), E.PARITY_MODE) and not G.cond_key_alive(ent[1]) and E.rebind_cond(ent[3], st):
                st = ent[3]
                del cache[k]
                break
        cache[:] = cache[-3:]   # conditional + unconditional (CFG) and a little slack; dead entries wait to be rebound
        cache.append(((tuple(x.shape), E.PARITY_MODE), G.cond_key(cond_input), wstamp, st))
        return st

    def forward(self, x, t, cond_input=None):
        if not x.is_cuda:
            raise rt.B200Error("Unet.forward: CUDA tensors required (this package has no CPU implementation; "
                               "the reference itself is the CPU path)")
        if self.cond:
            assert cond_input is not None, "Model initialized with conditioning so cond_input cannot be None"
        tt = torch.as_tensor(t).to(device=x.device, dtype=torch.int64).reshape(-1)
        if tt.numel() not in (1, x.shape[0]):
            raise RuntimeError(f"t has {tt.numel()} entries for batch size {x.shape[0]}")
        if torch.is_grad_enabled() and G.ENABLE_TRAIN_GRAPHS and not torch.cuda.is_current_stream_capturing() \
                and not self.class_cond and not isinstance(cond_input, E.CondState) \
                and (x.requires_grad or any(p.requires_grad for p in self.parameters())):
            return self._forward_train_graphed(x, tt, cond_input)
        with torch.autocast("cuda", enabled=False):
            return self._forward_eager(x, tt, cond_input)

    def _forward_train_graphed(self, x, tt, cond_input):
        cond_t = []
        if self.text_cond:
            assert 'text' in cond_input, \
                "Model initialized with text conditioning but cond_input has no text information"
            cond_t.append(cond_input['text'].float().contiguous())
        if self.image_cond:
            validate_image_conditional_input(cond_input, x)
            cond_t.append(cond_input['image'].float().contiguous())
        x = x.float().contiguous()
        tg = self.__dict__.get("_train_graph")
        ov = self.__dict__.get("_grad_overlap")   # (process group, bucket elements): set by sdb200.parallel.DataParallel
        if tg is None or not tg.still_valid(self, x, tt, cond_t, ov):
            tg = G.TrainGraph(self, x, tt, cond_t, self.text_cond, self.image_cond, overlap=ov)
            self.__dict__["_train_graph"] = tg
        tg.grad_sync = self.__dict__.get("_grad_sync")   # set by sdb200.parallel.DataParallel
        return tg(x, tt, cond_t)

    def _forward_eager(self, x, tt, cond_input):
        # an unconditional model has ONE (empty) condition state: a fresh object per call would change the captured
        # forward's key on every call and re-capture the graph on every timestep of tools/sample_ddpm_vqvae.py's loop
        st = self._cached_cond(cond_input, x) if self.cond else self.__dict__.setdefault("_no_cond", E.CondState())
        if G.ENABLE_FORWARD_GRAPHS and not torch.is_grad_enabled() and not torch.cuda.is_current_stream_capturing():
            graphs = self.__dict__.setdefault("_graphed", {})
            key = (tuple(x.shape), tt.numel(), id(st), E.PARITY_MODE)
            gf = graphs.get(key)
            if gf is None or not gf.valid(st):
                if len(graphs) >= 4:
                    graphs.pop(next(iter(graphs)))
                gf = graphs[key] = G.GraphedForward(self, x, tt, st)
            return gf(x, tt)
        return G.as_nchw(self.run(x, tt, st))

    def _weight_side(self, before_cond=False):
        """Weight side of every cim_layers leaf of this model (LSQ quantiser + noise) as one launch per phase
        (sdb200.quant.prepare_weight_side). `run` calls it; a caller that also runs the conditioning leaves in the same
        pass (the captured training step) calls it first with before_cond=True, and `run` then skips its own."""
        from sdb200 import quant as Q
        if not before_cond and self.__dict__.pop("_wq_prepared", False):
            return
        Q.prepare_weight_side(self)
        if before_cond:
            self.__dict__["_wq_prepared"] = True

    def invalidate_plan(self):
        """Drop captured graphs / hoisted conditioning (call after swapping sub-modules in place)."""
        self.__dict__.pop("_graphed", None)
        self.__dict__.pop("_cond_cache", None)
        self.__dict__.pop("_train_graph", None)

    def train(self, mode=True):
        self.invalidate_plan()
        return super().train(mode)

    def _apply(self, fn, *a, **k):
        self.invalidate_plan()
        return super()._apply(fn, *a, **k)

"""Weight bank of a captured training step (graphs.TrainGraph).

Every optimizer step changes every fp32 parameter, so every step re-derives the bf16 K-major matrices the GEMM kernels
read: forward form, transposed form (Linear dgrad), flipped-tap / channel-swapped form (conv dgrad), sub-pixel phase
forms (transposed convs). Leaf by leaf with torch ops that is ~700 small kernels per step for the conditioned UNet.
The bank learns the set of (parameter, form) pairs during the warm-up pass, keeps one persistent bf16 arena entry per
pair, and refreshes ALL of them with one table-driven launch (b200sd_pack_weights) at the top of the captured forward.

It also hands the backward a parameter-layout fp32 view into the step's flat gradient buffer, so the weight-gradient
GEMM's split-K reduction (b200sd_wgrad_scatter) writes the gradient where the optimizer / all-reduce will read it —
no packed intermediate, no permute copy, no gather into the flat buffer.

Tensors that ARE parameter storage (matched by address) go through the bank, and so do the quantised (+ noisy) weights
of the cim_layers leaves: their producer (sdb200.quant.prepare_weight_side, one launch per phase for the whole model)
writes them into a buffer the bank owns (`derived_buffer`), and a SECOND job table, refreshed right after that
producer (`refresh_derived`), re-derives their packed forms — one launch instead of one or two per leaf and direction.
Other derived weights (concatenated time-embedding projections, slices) keep the per-use torch path, because their
producers run at arbitrary points inside the step.
"""
import torch

from . import runtime as rt

ACTIVE = None   # the bank of the TrainGraph being warmed up / captured, else None


def _pad64(c):
    return (c + 63) // 64 * 64


class WeightBank:
    def __init__(self, params, flat, slots):
        self.dev = flat.device
        self.index = {p.data_ptr(): i for i, p in enumerate(params) if p.numel() > 0}
        self.numel = [p.numel() for p in params]
        self.flat, self.slots = flat, slots
        self.entries = {}
        self.jobs = []
        self.table = None
        self.n_tiles = 0
        self.uses = {}          # parameter index -> gradient views handed out in the current pass
        self.shared = set()     # parameters used by more than one leaf: their gradients are summed by autograd instead
        self.dbuf = None        # fp32 buffer of the derived (quantised) weights, written inside the step
        self.derived = {}       # address -> numel of the derived weights living in dbuf
        self.djobs, self.dtable, self.dn_tiles = [], None, 0

    # ------------------------------------------------------------------------------------------ bookkeeping
    def knows(self, w):
        if w is None or w.dtype != torch.float32 or not w.is_contiguous():
            return False
        ptr = w.data_ptr()
        if self.derived.get(ptr) == w.numel():
            return True
        return self.index.get(ptr, -1) >= 0 and self.numel[self.index[ptr]] == w.numel()

    def derived_buffer(self, sizes):
        """Persistent fp32 storage for the derived weights of one pass (numel per weight, in order); the views of it
        become tensors the bank knows. One buffer per bank: a pass with a different set of sizes replaces it (and the
        forms registered on the old one, which only warm-up passes can do)."""
        total = sum(sizes)
        if self.dbuf is None or self.dbuf.numel() != total or list(self.derived.values()) != list(sizes):
            if torch.cuda.is_current_stream_capturing():
                raise rt.B200Error("weight bank: the set of quantised weights changed between warm-up and capture")
            self.dbuf = torch.empty(max(total, 1), device=self.dev, dtype=torch.float32)
            stale = set(self.derived)
            self.entries = {k: v for k, v in self.entries.items() if not (set(k[1:]) & stale)}
            self.djobs, self.dtable = [], None
            self.derived, o = {}, 0
            for n in sizes:
                self.derived[self.dbuf.data_ptr() + 4 * o] = n
                o += n
        return self.dbuf

    def _is_derived(self, src):
        b = self.dbuf
        return b is not None and b.data_ptr() <= src.data_ptr() < b.data_ptr() + 4 * b.numel()

    def grad_view(self, w):
        """fp32 view of the flat gradient buffer shaped like parameter `w`, or None."""
        if not self.knows(w) or w.data_ptr() not in self.index:
            return None
        i = self.index[w.data_ptr()]
        sl = self.slots[i]
        if sl is None or i in self.shared:
            return None
        self.uses[i] = self.uses.get(i, 0) + 1
        return self.flat[sl[0]:sl[0] + sl[1]].view(w.shape)

    def _job(self, src, dst, rows, cin, cp, taps, s_r, s_c, tap_off, src_off=0):
        j = rt.PackJob()
        j.src, j.dst = src.data_ptr() + 4 * src_off, dst.data_ptr()
        j.s_r, j.s_c, j.dst_ld = s_r, s_c, dst.stride(0)
        j.rows, j.cin, j.cp, j.taps = rows, cin, cp, taps
        for t, o in enumerate(tap_off):
            j.tap_off[t] = o
        j.tiles_c = cp // 32
        return j, ((rows + 31) // 32) * (cp // 32), self._is_derived(src)

    def _launch(self, jobs_with_tiles):
        arr = (rt.PackJob * len(jobs_with_tiles))()
        t0 = 0
        for i, (j, nt) in enumerate(jobs_with_tiles):
            j.tile0 = t0
            arr[i] = j
            t0 += nt
        raw = bytes(arr)
        table = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(self.dev)
        return table, len(jobs_with_tiles), t0

    def _add(self, key, build):
        """build() -> (result, [(job, tiles), ...]); registers the jobs, runs them once now."""
        hit = self.entries.get(key)
        if hit is not None:
            return hit
        if torch.cuda.is_current_stream_capturing():
            raise rt.B200Error("weight bank: a weight form first appeared during graph capture (warm-up missed it)")
        result, jobs = build()
        table, n, tiles = self._launch([(j, nt) for j, nt, _ in jobs])
        rt.pack_weights(table, n, tiles)
        for j, nt, der in jobs:
            if der:
                self.djobs.append((j, nt))
                self.dtable = None
            else:
                self.jobs.append((j, nt))
                self.table = None    # stale: finalize() rebuilds
        self.entries[key] = result
        return result

    def begin_pass(self):
        self.uses = {}

    def finalize(self):
        """End of a warm-up pass: freeze the job table; parameters that received two in-place gradients in one pass
        (weight sharing) fall back to autograd's summation from the next pass on."""
        self.shared |= {i for i, c in self.uses.items() if c > 1}
        if self.jobs and self.table is None:
            self.table, self.n_jobs, self.n_tiles = self._launch(self.jobs)
        if self.djobs and self.dtable is None:
            self.dtable, self.dn_jobs, self.dn_tiles = self._launch(self.djobs)

    def refresh(self):
        """Re-derive every registered form from the current parameter values: one launch."""
        if self.table is not None:
            rt.pack_weights(self.table, self.n_jobs, self.n_tiles)

    def refresh_derived(self):
        """Re-derive the forms of the derived weights; their producer calls this right after it has written them."""
        if self.dtable is not None:
            rt.pack_weights(self.dtable, self.dn_jobs, self.dn_tiles)

    # ------------------------------------------------------------------------------------------ forms
    def _alloc(self, rows, cols):
        return torch.empty(rows, cols, device=self.dev, dtype=torch.bfloat16)

    def linear(self, w, row0=0, rows=None):
        """[N, K] (rows row0 .. row0 + rows of it) -> [rows, K_pad64]."""
        k = w.numel() // w.shape[0]
        n = w.shape[0] - row0 if rows is None else rows

        def build():
            dst = self._alloc(n, _pad64(k))
            return dst, [self._job(w, dst, n, k, _pad64(k), 1, k, 1, [0], src_off=row0 * k)]
        return self._add(("lin", w.data_ptr(), row0, n), build)

    def linear_t(self, w, row0=0, rows=None):
        """[N, K] (rows row0 .. row0 + rows of it) -> transposed [K, rows_pad64] (Linear / 1x1-conv dgrad)."""
        k = w.numel() // w.shape[0]
        n = w.shape[0] - row0 if rows is None else rows

        def build():
            dst = self._alloc(k, _pad64(n))
            return dst, [self._job(w, dst, k, n, _pad64(n), 1, 1, k, [0], src_off=row0 * k)]
        return self._add(("linT", w.data_ptr(), row0, n), build)

    def block(self, w, c0, cn, r0, rn):
        """Rows c0 .. c0 + cn, columns r0 .. r0 + rn of [N, K] (one mapped array block of a compute-in-memory layer)
        -> [cn, rn_pad64]."""
        k = w.shape[1]

        def build():
            dst = self._alloc(cn, _pad64(rn))
            return dst, [self._job(w, dst, cn, rn, _pad64(rn), 1, k, 1, [0], src_off=c0 * k + r0)]
        return self._add(("blk", w.data_ptr(), c0, cn, r0, rn), build)

    def block_t(self, w, c0, cn, r0, rn):
        """The same block transposed -> [rn, cn_pad64] (its input-gradient GEMM)."""
        k = w.shape[1]

        def build():
            dst = self._alloc(rn, _pad64(cn))
            return dst, [self._job(w, dst, rn, cn, _pad64(cn), 1, 1, k, [0], src_off=c0 * k + r0)]
        return self._add(("blkT", w.data_ptr(), c0, cn, r0, rn), build)

    def conv(self, w, w2=None):
        """Conv2d [co, ci, kh, kw] -> [co, kh*kw*ci_pad] (+ a fused 1x1 shortcut's [co, ci2, 1, 1] appended along K)."""
        co, ci, kh, kw = w.shape
        kk, cp = kh * kw, _pad64(ci)
        ci2 = 0 if w2 is None else w2.shape[1]
        cols = kk * cp + (_pad64(ci2) if w2 is not None else 0)

        def build():
            dst = self._alloc(co, cols)
            jobs = [self._job(w, dst, co, ci, cp, kk, ci * kk, kk, list(range(kk)))]
            if w2 is not None:
                jobs.append(self._job(w2, dst[:, kk * cp:], co, ci2, _pad64(ci2), 1, ci2, 1, [0]))
            return dst, jobs
        return self._add(("conv", w.data_ptr(), 0 if w2 is None else w2.data_ptr()), build)

    def conv_dgrad(self, w):
        """Conv2d [co, ci, kh, kw] -> the dgrad convolution's weights: taps flipped, channels swapped: [ci, kk*co_pad]."""
        co, ci, kh, kw = w.shape
        kk, cp = kh * kw, _pad64(co)

        def build():
            dst = self._alloc(ci, kk * cp)
            return dst, [self._job(w, dst, ci, co, cp, kk, kk, ci * kk, [kk - 1 - t for t in range(kk)])]
        return self._add(("convD", w.data_ptr()), build)

    def convT_phases(self, w, taps_of_phase):
        """ConvTranspose2d(4, 2, 1) weight [ci, co, 4, 4] -> four sub-pixel phase matrices [co, 4*ci_pad] (also the
        dgrad of Conv2d(4, 2, 1) with the same tensor read as [in = co_conv, out = ci_conv])."""
        ci, co = w.shape[0], w.shape[1]
        cp = _pad64(ci)

        def build():
            outs, jobs = [], []
            for taps, segs in taps_of_phase:
                dst = self._alloc(co, 4 * cp)
                jobs.append(self._job(w, dst, co, ci, cp, 4, 16, co * 16, [ky * 4 + kx for ky, kx in taps]))
                outs.append((dst, segs))
            return outs, jobs
        return self._add(("convT", w.data_ptr()), build)

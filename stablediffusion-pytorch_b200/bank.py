"""Weight bank of a captured training step (graphs.TrainGraph).

Every optimizer step changes every fp32 parameter, so every step re-derives the bf16 K-major matrices the GEMM kernels
read: forward form, transposed form (Linear dgrad), flipped-tap / channel-swapped form (conv dgrad), sub-pixel phase
forms (transposed convs). Leaf by leaf with torch ops that is ~700 small kernels per step for the conditioned UNet.
The bank learns the set of (parameter, form) pairs during the warm-up pass, keeps one persistent bf16 arena entry per
pair, and refreshes ALL of them with one table-driven launch (b200sd_pack_weights) at the top of the captured forward.

It also hands the backward a parameter-layout fp32 view into the step's flat gradient buffer, so the weight-gradient
GEMM's split-K reduction (b200sd_wgrad_scatter) writes the gradient where the optimizer / all-reduce will read it —
no packed intermediate, no permute copy, no gather into the flat buffer.

Only tensors that ARE parameter storage (matched by address) go through the bank; derived weights (concatenated
time-embedding projections, quantised weights, slices) keep the per-use torch path, because their producers run inside
the step.
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

    # ------------------------------------------------------------------------------------------ bookkeeping
    def knows(self, w):
        return w is not None and w.dtype == torch.float32 and w.is_contiguous() and \
            self.index.get(w.data_ptr(), -1) >= 0 and self.numel[self.index[w.data_ptr()]] == w.numel()

    def grad_view(self, w):
        """fp32 view of the flat gradient buffer shaped like parameter `w`, or None."""
        if not self.knows(w):
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
        return j, ((rows + 31) // 32) * (cp // 32)

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
        self.jobs.extend(jobs)
        table, n, tiles = self._launch([(j, nt) for j, nt in jobs])
        rt.pack_weights(table, n, tiles)
        self.entries[key] = result
        self.table = None    # stale: finalize() rebuilds
        return result

    def begin_pass(self):
        self.uses = {}

    def finalize(self):
        """End of a warm-up pass: freeze the job table; parameters that received two in-place gradients in one pass
        (weight sharing) fall back to autograd's summation from the next pass on."""
        self.shared |= {i for i, c in self.uses.items() if c > 1}
        if self.jobs and self.table is None:
            self.table, self.n_jobs, self.n_tiles = self._launch(self.jobs)

    def refresh(self):
        """Re-derive every registered form from the current parameter values: one launch."""
        if self.table is not None:
            rt.pack_weights(self.table, self.n_jobs, self.n_tiles)

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

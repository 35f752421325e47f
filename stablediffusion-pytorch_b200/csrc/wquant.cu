// Weight side of the cim_layers drop-ins as four kernels per layer (two forward, two backward) instead of ~50 small
// torch kernels (profile of the LSQ training leg: 9,948 launches per step, most of them these):
//
//   forward   weight_quant_noise (cim_layers/layers_utils_lsq.py:48-64) =
//               lazy step init      if step == 1: step = 1 / (1 / max|w| * Qp)                 (:40-42, init_step_size :31-34)
//               weight_quant_lsq    s' = grad_scale(step, 1/sqrt(Qp numel)); q = round(clamp(w / s', -Qp, Qp));
//                                   wq = q s'   (isint: q s' / s')                              (quant_noise_utils.py:160-179)
//               add_noise           out = wq + (max wq - min wq) * n_scale * N(0, 1)            (quant_noise_utils.py:92-98)
//             kernel 1 reduces max|w|, min w, max w (clamp + round are monotone, so the extrema of wq are the quantised
//             extrema of w) and performs the lazy init in its last CTA; kernel 2 is elementwise and counts the elements
//             tied at the two extreme codes (torch's max() / min() backward spreads the gradient evenly over ties).
//   backward  of exactly that graph: straight-through round, clamp mask, the step-size gradient with its 1/sqrt(Qp numel)
//             scale, and the noise term's gradient through max / min. Kernel 3 reduces T = sum(g * unit); kernel 4 is
//             elementwise and leaves per-CTA partials of d(step), folded in fixed order by its last CTA.
// All arithmetic is fp32 in the reference's operation order with explicitly rounded intrinsics where a contraction
// would change a code.
#include "host.h"

namespace b200 {

constexpr int WQ_THREADS = 256;

__device__ __forceinline__ float wq_effective_step(float s, float g) {
    const float y = __fmul_rn(s, g);
    return __fadd_rn(__fsub_rn(s, y), y);
}

__device__ __forceinline__ bool wq_last_cta(unsigned int* counter, unsigned int total) {
    __shared__ unsigned int s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int t = atomicAdd(counter, 1u);
        s_last = (t == total - 1u) ? 1u : 0u;
        if (s_last) *counter = 0u;
    }
    __syncthreads();
    if (s_last) __threadfence();
    return s_last != 0u;
}

// Fold of the per-CTA partials by ALL threads of the last CTA (a single thread walking a few hundred __ldcg loads took
// ~10 us per launch — longer than the elementwise pass over an 83 K-element weight). Fixed order: thread t takes
// partials t, t + 256, ...; warp shuffle tree; warp results in order.
template <int OP>   // 0: sum, 1: max, 2: min
__device__ __forceinline__ float wq_combine(float a, float b) {
    return OP == 0 ? a + b : (OP == 1 ? fmaxf(a, b) : fminf(a, b));
}
template <int OP>
__device__ __forceinline__ float wq_fold(const float* part, int stride, unsigned int count, float identity, float* red) {
    float v = identity;
    for (unsigned int k = threadIdx.x; k < count; k += WQ_THREADS) v = wq_combine<OP>(v, __ldcg(part + k * stride));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = wq_combine<OP>(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = red[0];
#pragma unroll
    for (int k = 1; k < WQ_THREADS / 32; ++k) r = wq_combine<OP>(r, red[k]);
    return r;
}

// scratch layout (floats unless noted): [0] max|w|  [1] min w  [2] max w  [3] T  [4] d(step)  [5..6] uint tie counts
// (max, min)  [7] uint arrival counter  [8 ..) per-CTA partials (3 per CTA in kernel 1, 1 per CTA in kernels 3 / 4)
struct WqScratch {
    float absmax, wmin, wmax, T, dstep;
    unsigned int cnt_max, cnt_min, arrive;
    float part[1];
};

// Every kernel body takes (cta, nctas) instead of reading blockIdx / gridDim: the single-layer kernels pass their own
// grid, the table-driven kernels (one launch for ALL quantised layers of a model) pass a layer's slice of the grid.
__device__ __forceinline__ void wq_stats_body(const float* __restrict__ w, long long n, float* __restrict__ step, float qp,
                                              int lazy_init, WqScratch* __restrict__ sc, unsigned int cta,
                                              unsigned int nctas) {
    __shared__ float red[3][WQ_THREADS / 32];
    float a = 0.f, lo = INFINITY, hi = -INFINITY;
    for (long long i = static_cast<long long>(cta) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(nctas) * blockDim.x) {
        const float v = w[i];
        a = fmaxf(a, fabsf(v)), lo = fminf(lo, v), hi = fmaxf(hi, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a = fmaxf(a, __shfl_xor_sync(0xffffffffu, a, o));
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) red[0][threadIdx.x >> 5] = a, red[1][threadIdx.x >> 5] = lo, red[2][threadIdx.x >> 5] = hi;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < WQ_THREADS / 32; ++k) a = fmaxf(a, red[0][k]), lo = fminf(lo, red[1][k]), hi = fmaxf(hi, red[2][k]);
        sc->part[3 * cta] = a, sc->part[3 * cta + 1] = lo, sc->part[3 * cta + 2] = hi;
    }
    if (wq_last_cta(&sc->arrive, nctas)) {
        a = wq_fold<1>(sc->part, 3, nctas, 0.f, red[0]);
        lo = wq_fold<2>(sc->part + 1, 3, nctas, INFINITY, red[1]);
        hi = wq_fold<1>(sc->part + 2, 3, nctas, -INFINITY, red[2]);
        if (threadIdx.x == 0) {
            sc->absmax = a, sc->wmin = lo, sc->wmax = hi;
            sc->cnt_max = 0u, sc->cnt_min = 0u;
            if (lazy_init && *step == 1.0f && a != 0.f)   // init_step_size: 1 / (1 / data_range * Qp), in that order
                *step = __fdiv_rn(1.0f, __fmul_rn(__fdiv_rn(1.0f, a), qp));
        }
    }
}
__global__ void __launch_bounds__(WQ_THREADS)
    wq_stats_kernel(const float* __restrict__ w, long long n, float* __restrict__ step, float qp, int lazy_init,
                    WqScratch* __restrict__ sc) {
    pdl_enter();
    wq_stats_body(w, n, step, qp, lazy_init, sc, blockIdx.x, gridDim.x);
}

__device__ __forceinline__ float wq_code(float v, float s, float qp) {
    return rintf(fminf(fmaxf(__fdiv_rn(v, s), -qp), qp));
}
__device__ __forceinline__ float wq_value(float q, float s, int isint) {
    const float v = __fmul_rn(q, s);
    return isint ? __fdiv_rn(v, s) : v;
}

__device__ __forceinline__ void wq_apply_body(const float* __restrict__ w, long long n, const float* __restrict__ step,
                                              float qp, float g, int isint, float n_scale, const float* __restrict__ unit,
                                              float* __restrict__ out, WqScratch* __restrict__ sc, unsigned int cta,
                                              unsigned int nctas) {
    __shared__ unsigned int s_cnt[2];
    if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0u;
    __syncthreads();
    const float s = wq_effective_step(*step, g);
    const float q_hi = wq_code(sc->wmax, s, qp), q_lo = wq_code(sc->wmin, s, qp);
    // (wq.max() - wq.min()) * n_scale, each operation rounded as the chain of torch ops does
    const float amp = __fmul_rn(__fsub_rn(wq_value(q_hi, s, isint), wq_value(q_lo, s, isint)), n_scale);
    unsigned int c_hi = 0u, c_lo = 0u;
    for (long long i = static_cast<long long>(cta) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(nctas) * blockDim.x) {
        const float q = wq_code(w[i], s, qp);
        c_hi += q == q_hi, c_lo += q == q_lo;
        float v = wq_value(q, s, isint);
        if (unit) v = __fadd_rn(v, __fmul_rn(amp, unit[i]));
        out[i] = v;
    }
    if (unit) {
        c_hi = __reduce_add_sync(0xffffffffu, c_hi), c_lo = __reduce_add_sync(0xffffffffu, c_lo);
        if ((threadIdx.x & 31) == 0) atomicAdd(&s_cnt[0], c_hi), atomicAdd(&s_cnt[1], c_lo);
        __syncthreads();
        if (threadIdx.x == 0) atomicAdd(&sc->cnt_max, s_cnt[0]), atomicAdd(&sc->cnt_min, s_cnt[1]);
    }
}
__global__ void __launch_bounds__(WQ_THREADS)
    wq_apply_kernel(const float* __restrict__ w, long long n, const float* __restrict__ step, float qp, float g, int isint,
                    float n_scale, const float* __restrict__ unit, float* __restrict__ out, WqScratch* __restrict__ sc) {
    pdl_enter();
    wq_apply_body(w, n, step, qp, g, isint, n_scale, unit, out, sc, blockIdx.x, gridDim.x);
}

// T = sum_i g_i * unit_i (fixed order: per-CTA partials, folded by the last CTA)
__device__ __forceinline__ void wq_bwd_dot_body(const float* __restrict__ gout, const float* __restrict__ unit,
                                                long long n, WqScratch* __restrict__ sc, unsigned int cta,
                                                unsigned int nctas) {
    __shared__ float red[WQ_THREADS / 32];
    float a = 0.f;
    for (long long i = static_cast<long long>(cta) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(nctas) * blockDim.x)
        a = fmaf(gout[i], unit[i], a);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < WQ_THREADS / 32; ++k) a += red[k];
        sc->part[cta] = a;
    }
    if (wq_last_cta(&sc->arrive, nctas)) {
        const float t = wq_fold<0>(sc->part, 1, nctas, 0.f, red);
        if (threadIdx.x == 0) sc->T = t;
    }
}
__global__ void __launch_bounds__(WQ_THREADS)
    wq_bwd_dot_kernel(const float* __restrict__ gout, const float* __restrict__ unit, long long n, WqScratch* __restrict__ sc) {
    pdl_enter();
    wq_bwd_dot_body(gout, unit, n, sc, blockIdx.x, gridDim.x);
}

__device__ __forceinline__ void wq_bwd_apply_body(const float* __restrict__ gout, const float* __restrict__ w, long long n,
                                                  const float* __restrict__ step, float qp, float g, int isint,
                                                  float n_scale, int noisy, float* __restrict__ dw,
                                                  float* __restrict__ dstep, WqScratch* __restrict__ sc, unsigned int cta,
                                                  unsigned int nctas) {
    __shared__ float red[WQ_THREADS / 32];
    const float s = wq_effective_step(*step, g);
    const float inv_s = __fdiv_rn(1.0f, s);
    const float q_hi = wq_code(sc->wmax, s, qp), q_lo = wq_code(sc->wmin, s, qp);
    float t_hi = 0.f, t_lo = 0.f;
    if (noisy) {   // d out / d max(wq) = n_scale * T, spread evenly over the tied elements; min likewise with the other sign
        const float t = n_scale * sc->T;
        t_hi = t / static_cast<float>(sc->cnt_max), t_lo = t / static_cast<float>(sc->cnt_min);
    }
    float acc = 0.f;
    for (long long i = static_cast<long long>(cta) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(nctas) * blockDim.x) {
        const float u = __fdiv_rn(w[i], s);
        const bool in = (u >= -qp) && (u <= qp);
        const float q = rintf(fminf(fmaxf(u, -qp), qp));
        float gg = gout[i];
        if (noisy) gg += (q == q_hi ? t_hi : 0.f) - (q == q_lo ? t_lo : 0.f);
        acc += gg * (q - (in ? u : 0.f));
        dw[i] = in ? (isint ? gg * inv_s : gg) : 0.f;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < WQ_THREADS / 32; ++k) acc += red[k];
        sc->part[cta] = acc;
    }
    if (wq_last_cta(&sc->arrive, nctas)) {
        const float t = wq_fold<0>(sc->part, 1, nctas, 0.f, red);
        if (threadIdx.x == 0) *dstep = (isint ? t * inv_s : t) * g;   // grad_scale's backward: the 1/sqrt(Qp numel) factor
    }
}
__global__ void __launch_bounds__(WQ_THREADS)
    wq_bwd_apply_kernel(const float* __restrict__ gout, const float* __restrict__ w, long long n,
                        const float* __restrict__ step, float qp, float g, int isint, float n_scale, int noisy,
                        float* __restrict__ dw, float* __restrict__ dstep, WqScratch* __restrict__ sc) {
    pdl_enter();
    wq_bwd_apply_body(gout, w, n, step, qp, g, isint, n_scale, noisy, dw, dstep, sc, blockIdx.x, gridDim.x);
}

// ---- table-driven launches: one kernel per phase for up to WQ_MULTI layers, the table passed BY VALUE in the kernel
// parameters (no device-side table to upload or keep alive, so the same call is CUDA-graph capturable and an eager
// step may hand in fresh gradient pointers every time). CTA c serves the layer whose [cta0, cta0 + nctas) holds c.
constexpr int WQ_MULTI = 64;
struct WqJob {
    const float* w;
    float* step;
    const float* unit;   // N(0,1) draw or null (no noise)
    float* out;          // forward: quantised (+ noisy) weight; backward: d weight
    const float* gout;   // backward: gradient of `out`
    float* dstep;
    WqScratch* sc;
    long long n;
    float qp, g, n_scale;
    int flags;           // bit 0: isint, bit 1: lazy_init
    unsigned int cta0, nctas;
};
struct WqTable {
    WqJob job[WQ_MULTI];
    int n_jobs;
};

template <int PHASE>   // 0 stats, 1 apply, 2 backward dot, 3 backward apply
__global__ void __launch_bounds__(WQ_THREADS) wq_multi_kernel(const __grid_constant__ WqTable t) {
    pdl_enter();
    int j = 0;
    while (j + 1 < t.n_jobs && blockIdx.x >= t.job[j + 1].cta0) ++j;
    const WqJob& q = t.job[j];
    const unsigned int cta = blockIdx.x - q.cta0;
    if (PHASE == 0) wq_stats_body(q.w, q.n, q.step, q.qp, (q.flags >> 1) & 1, q.sc, cta, q.nctas);
    if (PHASE == 1) wq_apply_body(q.w, q.n, q.step, q.qp, q.g, q.flags & 1, q.n_scale, q.unit, q.out, q.sc, cta, q.nctas);
    if (PHASE == 2) wq_bwd_dot_body(q.gout, q.unit, q.n, q.sc, cta, q.nctas);
    if (PHASE == 3)
        wq_bwd_apply_body(q.gout, q.w, q.n, q.step, q.qp, q.g, q.flags & 1, q.n_scale, q.unit != nullptr, q.out, q.dstep,
                          q.sc, cta, q.nctas);
}

static int wq_grid(long long n) {
    long long b = (n + 4 * WQ_THREADS - 1) / (4 * WQ_THREADS);
    const long long cap = 2ll * sm_count();
    if (b > cap) b = cap;
    return b < 1 ? 1 : static_cast<int>(b);
}

}  // namespace b200

using namespace b200;

extern "C" int b200sd_wq_scratch_floats(void) { return 8 + 3 * 2 * 148 * 2; }

template <int PHASE>
static int wq_launch_chunks(const b200sd_wq_job* jobs, int n_jobs, bool noisy_only, cudaStream_t st, const char* what) {
    WqTable t;
    int k = 0;
    unsigned int ctas = 0;
    auto flush = [&]() -> int {
        if (k == 0) return B200SD_OK;
        t.n_jobs = k;
        launch_k(wq_multi_kernel<PHASE>, ctas, WQ_THREADS, 0, st, t);
        B200_CHECK_LAUNCH(what);
        k = 0, ctas = 0;
        return B200SD_OK;
    };
    for (int i = 0; i < n_jobs; ++i) {
        const b200sd_wq_job& a = jobs[i];
        if (noisy_only && !(a.noise_scale != 0.f && a.unit)) continue;
        WqJob& q = t.job[k];
        q.w = a.w, q.step = a.step, q.unit = a.noise_scale != 0.f ? a.unit : nullptr, q.out = a.out, q.gout = a.gout;
        q.dstep = a.dstep, q.sc = reinterpret_cast<WqScratch*>(a.scratch), q.n = a.n;
        q.qp = a.qp, q.g = a.grad_scale, q.n_scale = a.noise_scale;
        q.flags = (a.isint ? 1 : 0) | (a.lazy_init ? 2 : 0);
        q.cta0 = ctas, q.nctas = static_cast<unsigned int>(wq_grid(a.n));
        ctas += q.nctas;
        if (++k == WQ_MULTI) {
            const int rc = flush();
            if (rc != B200SD_OK) return rc;
        }
    }
    return flush();
}

extern "C" int b200sd_weight_qn_multi(const b200sd_wq_job* jobs, int32_t n_jobs, int32_t backward, void* stream) {
    B200_REQUIRE(jobs && n_jobs > 0, "weight_qn_multi: bad arguments");
    for (int i = 0; i < n_jobs; ++i) {
        const b200sd_wq_job& a = jobs[i];
        B200_REQUIRE(a.w && a.step && a.out && a.scratch && a.n > 0 && a.qp >= 1.f, "weight_qn_multi: bad table row");
        B200_REQUIRE(a.noise_scale == 0.f || a.unit, "weight_qn_multi: noise needs the N(0,1) draw");
        B200_REQUIRE(!backward || (a.gout && a.dstep), "weight_qn_multi: the backward pass needs gout and dstep");
    }
    cudaStream_t st = as_stream(stream);
    int rc;
    if (!backward) {
        if ((rc = wq_launch_chunks<0>(jobs, n_jobs, false, st, "weight_qn_multi(stats)")) != B200SD_OK) return rc;
        return wq_launch_chunks<1>(jobs, n_jobs, false, st, "weight_qn_multi(apply)");
    }
    if ((rc = wq_launch_chunks<2>(jobs, n_jobs, true, st, "weight_qn_multi(dot)")) != B200SD_OK) return rc;
    return wq_launch_chunks<3>(jobs, n_jobs, false, st, "weight_qn_multi(bwd)");
}

extern "C" int b200sd_weight_qn(const float* w, int64_t n, float* step, float qp, float grad_scale, int32_t isint,
                                int32_t lazy_init, float noise_scale, const float* unit, float* out, float* scratch,
                                void* stream) {
    B200_REQUIRE(w && step && out && scratch && n > 0 && qp >= 1.f, "weight_qn: bad arguments");
    B200_REQUIRE(noise_scale == 0.f || unit, "weight_qn: noise needs the N(0,1) draw");
    cudaStream_t st = as_stream(stream);
    const int grid = wq_grid(n);
    WqScratch* sc = reinterpret_cast<WqScratch*>(scratch);
    launch_k(wq_stats_kernel, grid, WQ_THREADS, 0, st, w, (long long)n, step, qp, (int)lazy_init, sc);
    B200_CHECK_LAUNCH("weight_qn_stats");
    launch_k(wq_apply_kernel, grid, WQ_THREADS, 0, st, w, (long long)n, static_cast<const float*>(step), qp, grad_scale,
             (int)isint, noise_scale, noise_scale != 0.f ? unit : static_cast<const float*>(nullptr), out, sc);
    B200_CHECK_LAUNCH("weight_qn_apply");
    return B200SD_OK;
}

extern "C" int b200sd_weight_qn_bwd(const float* gout, const float* w, int64_t n, const float* step, float qp,
                                    float grad_scale, int32_t isint, float noise_scale, const float* unit, float* dw,
                                    float* dstep, float* scratch, void* stream) {
    B200_REQUIRE(gout && w && step && dw && dstep && scratch && n > 0 && qp >= 1.f, "weight_qn_bwd: bad arguments");
    B200_REQUIRE(noise_scale == 0.f || unit, "weight_qn_bwd: noise needs the N(0,1) draw of the forward");
    cudaStream_t st = as_stream(stream);
    const int grid = wq_grid(n);
    WqScratch* sc = reinterpret_cast<WqScratch*>(scratch);
    const int noisy = noise_scale != 0.f;
    if (noisy) {
        launch_k(wq_bwd_dot_kernel, grid, WQ_THREADS, 0, st, gout, unit, (long long)n, sc);
        B200_CHECK_LAUNCH("weight_qn_bwd_dot");
    }
    launch_k(wq_bwd_apply_kernel, grid, WQ_THREADS, 0, st, gout, w, (long long)n, step, qp, grad_scale, (int)isint,
             noise_scale, noisy, dw, dstep, sc);
    B200_CHECK_LAUNCH("weight_qn_bwd_apply");
    return B200SD_OK;
}

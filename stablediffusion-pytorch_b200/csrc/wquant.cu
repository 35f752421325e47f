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

__global__ void __launch_bounds__(WQ_THREADS)
    wq_stats_kernel(const float* __restrict__ w, long long n, float* __restrict__ step, float qp, int lazy_init,
                    WqScratch* __restrict__ sc) {
    pdl_enter();
    __shared__ float red[3][WQ_THREADS / 32];
    float a = 0.f, lo = INFINITY, hi = -INFINITY;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
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
        sc->part[3 * blockIdx.x] = a, sc->part[3 * blockIdx.x + 1] = lo, sc->part[3 * blockIdx.x + 2] = hi;
    }
    if (wq_last_cta(&sc->arrive, gridDim.x)) {
        a = wq_fold<1>(sc->part, 3, gridDim.x, 0.f, red[0]);
        lo = wq_fold<2>(sc->part + 1, 3, gridDim.x, INFINITY, red[1]);
        hi = wq_fold<1>(sc->part + 2, 3, gridDim.x, -INFINITY, red[2]);
        if (threadIdx.x == 0) {
            sc->absmax = a, sc->wmin = lo, sc->wmax = hi;
            sc->cnt_max = 0u, sc->cnt_min = 0u;
            if (lazy_init && *step == 1.0f && a != 0.f)   // init_step_size: 1 / (1 / data_range * Qp), in that order
                *step = __fdiv_rn(1.0f, __fmul_rn(__fdiv_rn(1.0f, a), qp));
        }
    }
}

__device__ __forceinline__ float wq_code(float v, float s, float qp) {
    return rintf(fminf(fmaxf(__fdiv_rn(v, s), -qp), qp));
}
__device__ __forceinline__ float wq_value(float q, float s, int isint) {
    const float v = __fmul_rn(q, s);
    return isint ? __fdiv_rn(v, s) : v;
}

__global__ void __launch_bounds__(WQ_THREADS)
    wq_apply_kernel(const float* __restrict__ w, long long n, const float* __restrict__ step, float qp, float g, int isint,
                    float n_scale, const float* __restrict__ unit, float* __restrict__ out, WqScratch* __restrict__ sc) {
    pdl_enter();
    __shared__ unsigned int s_cnt[2];
    if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0u;
    __syncthreads();
    const float s = wq_effective_step(*step, g);
    const float q_hi = wq_code(sc->wmax, s, qp), q_lo = wq_code(sc->wmin, s, qp);
    // (wq.max() - wq.min()) * n_scale, each operation rounded as the chain of torch ops does
    const float amp = __fmul_rn(__fsub_rn(wq_value(q_hi, s, isint), wq_value(q_lo, s, isint)), n_scale);
    unsigned int c_hi = 0u, c_lo = 0u;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
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

// T = sum_i g_i * unit_i (fixed order: per-CTA partials, folded by the last CTA)
__global__ void __launch_bounds__(WQ_THREADS)
    wq_bwd_dot_kernel(const float* __restrict__ gout, const float* __restrict__ unit, long long n, WqScratch* __restrict__ sc) {
    pdl_enter();
    __shared__ float red[WQ_THREADS / 32];
    float a = 0.f;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x)
        a = fmaf(gout[i], unit[i], a);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < WQ_THREADS / 32; ++k) a += red[k];
        sc->part[blockIdx.x] = a;
    }
    if (wq_last_cta(&sc->arrive, gridDim.x)) {
        const float t = wq_fold<0>(sc->part, 1, gridDim.x, 0.f, red);
        if (threadIdx.x == 0) sc->T = t;
    }
}

__global__ void __launch_bounds__(WQ_THREADS)
    wq_bwd_apply_kernel(const float* __restrict__ gout, const float* __restrict__ w, long long n,
                        const float* __restrict__ step, float qp, float g, int isint, float n_scale, int noisy,
                        float* __restrict__ dw, float* __restrict__ dstep, WqScratch* __restrict__ sc) {
    pdl_enter();
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
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
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
        sc->part[blockIdx.x] = acc;
    }
    if (wq_last_cta(&sc->arrive, gridDim.x)) {
        const float t = wq_fold<0>(sc->part, 1, gridDim.x, 0.f, red);
        if (threadIdx.x == 0) *dstep = (isint ? t * inv_s : t) * g;   // grad_scale's backward: the 1/sqrt(Qp numel) factor
    }
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

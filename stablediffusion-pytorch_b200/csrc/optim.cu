// Optimizer side of a training step: what the reference does with ~20 multi-tensor passes plus a
// per-parameter Python loop (train_ddpm_cond_celebhq_multi_gpu.py:362-378: clip_grad_norm_(1.0), isfinite check,
// Adam.step, EMA mul_/add_ per parameter) as two launches over a chunk table:
//
//   grad_sqnorm    sum of squares of every gradient element (fixed-order partials, last-CTA tail) -> total norm, clip
//                  coefficient min(1, max_norm / (norm + 1e-6)), a finite flag, and the step counter (+1 when finite)
//   adam_ema_step  g' = coef * g (+ wd * p);  m, v moments;  p -= lr/bc1 * m / (sqrt(v)/sqrt(bc2) + eps);
//                  ema = decay * ema + (1 - decay) * p   — one pass: 36 B/parameter of HBM traffic,
//                  skipped entirely (moments untouched) when the norm is not finite, as the reference script does.
//
// Gradients and moments live in flat fp32 buffers of 64-element-aligned slots, addressed by offset — the moments in
// parameter order, the gradients in whatever order the captured training step laid its buffer out (graphs.TrainGraph
// orders it by gradient arrival when the data-parallel exchange overlaps the backward pass), hence two offsets per
// chunk; parameters and EMA shadows stay where PyTorch allocated them, addressed by pointer.
#include "host.h"

namespace b200 {

constexpr int OPT_CHUNK = 32768;   // elements per table entry (python side splits tensors at this size)

struct OptChunk {                  // = b200sd_opt_chunk
    float* p;
    float* ema;
    long long off;                 // offset of the run in the moment buffers
    long long goff;                // offset of the run in the gradient buffer (its own layout: see graphs.TrainGraph)
    int n;
    int vec;                       // 1: p/ema 16-byte aligned, float4 path
};
static_assert(sizeof(OptChunk) == sizeof(b200sd_opt_chunk), "OptChunk must match the C-ABI struct");

__device__ __forceinline__ float block_sum(float v, float* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) red[w] = v;
    __syncthreads();
    float s = 0.f;
    if (threadIdx.x == 0)
        for (int i = 0; i < (blockDim.x >> 5); ++i) s += red[i];
    return s;   // valid in thread 0
}

__global__ void __launch_bounds__(256)
    grad_sqnorm_kernel(const OptChunk* __restrict__ tab, int n_chunks, const float* __restrict__ g, float max_norm,
                       int check_finite, float* __restrict__ partial, float* __restrict__ scalars,
                       unsigned int* counter) {
    pdl_enter();
    __shared__ float red[8];
    __shared__ unsigned int s_last;
    const OptChunk c = tab[blockIdx.x];
    const float* gp = g + c.goff;
    float s = 0.f;
    const int n4 = c.n >> 2;
    for (int i = threadIdx.x; i < n4; i += 256) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(gp) + i);
        s += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    for (int i = (n4 << 2) + threadIdx.x; i < c.n; i += 256) s += gp[i] * gp[i];
    s = block_sum(s, red);
    if (threadIdx.x == 0) {
        partial[blockIdx.x] = s;
        __threadfence();
        const unsigned int t = atomicAdd(counter, 1u);
        s_last = (t == gridDim.x - 1u) ? 1u : 0u;
        if (s_last) *counter = 0u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    float t = 0.f;
    for (int i = threadIdx.x; i < n_chunks; i += 256) t += __ldcg(partial + i);   // fixed assignment, fixed order
    t = block_sum(t, red);
    if (threadIdx.x == 0) {
        const float norm = sqrtf(t);
        const bool finite = !check_finite || (norm == norm && norm < 3.0e38f);
        float coef = 1.f;
        if (max_norm > 0.f) {
            coef = max_norm / (norm + 1e-6f);
            if (!(coef < 1.f)) coef = 1.f;
        }
        scalars[0] = norm;
        scalars[1] = coef;
        scalars[2] = finite ? 1.f : 0.f;
        if (finite) scalars[3] += 1.f;
    }
}

__global__ void __launch_bounds__(256)
    adam_ema_kernel(const OptChunk* __restrict__ tab, const float* __restrict__ g, float* __restrict__ m,
                    float* __restrict__ v, const float* __restrict__ scalars, float lr, float beta1, float beta2,
                    float eps, float wd, float ema_decay) {
    pdl_enter();
    if (scalars[2] == 0.f) return;   // non-finite gradient norm: the whole step is skipped
    const float coef = scalars[1];
    const double step = static_cast<double>(scalars[3]);
    const float bc1 = static_cast<float>(1.0 - pow(static_cast<double>(beta1), step));
    const float bc2_sqrt = static_cast<float>(sqrt(1.0 - pow(static_cast<double>(beta2), step)));
    const float step_size = lr / bc1;
    const float w1 = 1.f - beta1, w2 = 1.f - beta2, we = 1.f - ema_decay;
    const OptChunk c = tab[blockIdx.x];
    const float* gp = g + c.goff;
    float* mp = m + c.off;
    float* vp = v + c.off;

    auto upd = [&](float gv, float& mv, float& vv, float& pv, float& ev, bool has_ema) {
        gv *= coef;
        if (wd != 0.f) gv = gv + wd * pv;
        mv = mv + w1 * (gv - mv);                       // lerp_(grad, 1 - beta1)
        vv = vv * beta2 + w2 * (gv * gv);               // mul_(beta2).addcmul_(grad, grad, 1 - beta2)
        const float denom = sqrtf(vv) / bc2_sqrt + eps;
        pv = pv - step_size * (mv / denom);             // addcdiv_(exp_avg, denom, -step_size)
        if (has_ema) ev = ev * ema_decay + we * pv;     // mul_(decay).add_(param, alpha = 1 - decay)
    };

    const bool has_ema = c.ema != nullptr;
    int done = 0;
    if (c.vec) {
        const int n4 = c.n >> 2;
        for (int i = threadIdx.x; i < n4; i += 256) {
            const float4 g4 = __ldcs(reinterpret_cast<const float4*>(gp) + i);
            float4 m4 = reinterpret_cast<float4*>(mp)[i], v4 = reinterpret_cast<float4*>(vp)[i];
            float4 p4 = reinterpret_cast<float4*>(c.p)[i];
            float4 e4 = has_ema ? reinterpret_cast<float4*>(c.ema)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
            upd(g4.x, m4.x, v4.x, p4.x, e4.x, has_ema);
            upd(g4.y, m4.y, v4.y, p4.y, e4.y, has_ema);
            upd(g4.z, m4.z, v4.z, p4.z, e4.z, has_ema);
            upd(g4.w, m4.w, v4.w, p4.w, e4.w, has_ema);
            reinterpret_cast<float4*>(mp)[i] = m4;
            reinterpret_cast<float4*>(vp)[i] = v4;
            reinterpret_cast<float4*>(c.p)[i] = p4;
            if (has_ema) reinterpret_cast<float4*>(c.ema)[i] = e4;
        }
        done = n4 << 2;
    }
    for (int i = done + threadIdx.x; i < c.n; i += 256) {
        float mv = mp[i], vv = vp[i], pv = c.p[i], ev = has_ema ? c.ema[i] : 0.f;
        upd(gp[i], mv, vv, pv, ev, has_ema);
        mp[i] = mv, vp[i] = vv, c.p[i] = pv;
        if (has_ema) c.ema[i] = ev;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// PercentOptimizerFP (cim_layers/IBA_optimizer.py:80-143): Adam moments, then the whole tensor's update is rescaled so
// that its largest element is `lr` times the tensor's largest weight (never below plain `lr`):
//   m = b1*m + (1-b1)*g;  v = b2*v + (1-b2)*g*g;  u = (adj*m) / (sqrt(v)/sqrt(bc2) + eps)          (:107-123)
//   scale = max(lr * max|p| / (max|u| + eps), lr);  p -= u * scale                                   (:128-139)
// The reference walks the parameters in Python with ~14 tensor ops and two host-visible reductions each. Here:
// percent_moments (moments + per-tensor max|u|, max|p| by integer atomicMax on the float bits — exact and order
// independent) and percent_apply (u recomputed from the stored moments — same bits — then the scaled update).
// Every product/sum is an explicitly rounded intrinsic in the reference's operation order: no FMA contraction.
__device__ __forceinline__ float percent_update(float mv, float vv, float adj, float bc2s, float eps) {
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(vv), bc2s), eps);
    return __fdiv_rn(__fmul_rn(adj, mv), denom);
}

__global__ void __launch_bounds__(256)
    percent_moments_kernel(const OptChunk* __restrict__ tab, const int* __restrict__ tid, const float* __restrict__ g,
                           float* __restrict__ m, float* __restrict__ v, unsigned int* __restrict__ tmax, float beta1,
                           float beta2, float w1, float w2, float adj, float bc2s, float eps) {
    pdl_enter();
    __shared__ float red_u[8], red_p[8];
    const OptChunk c = tab[blockIdx.x];
    const float* gp = g + c.goff;
    float* mp = m + c.off;
    float* vp = v + c.off;
    float mu = 0.f, mw = 0.f;
    for (int i = threadIdx.x; i < c.n; i += 256) {
        const float gv = gp[i];
        const float mv = __fadd_rn(__fmul_rn(beta1, mp[i]), __fmul_rn(w1, gv));
        const float vv = __fadd_rn(__fmul_rn(beta2, vp[i]), __fmul_rn(w2, __fmul_rn(gv, gv)));
        mp[i] = mv, vp[i] = vv;
        mu = fmaxf(mu, fabsf(percent_update(mv, vv, adj, bc2s, eps)));
        mw = fmaxf(mw, fabsf(c.p[i]));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mu = fmaxf(mu, __shfl_xor_sync(0xffffffffu, mu, o));
        mw = fmaxf(mw, __shfl_xor_sync(0xffffffffu, mw, o));
    }
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) red_u[w] = mu, red_p[w] = mw;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < 8; ++i) mu = fmaxf(mu, red_u[i]), mw = fmaxf(mw, red_p[i]);
        const int t = tid[blockIdx.x];
        atomicMax(tmax + 2 * t, __float_as_uint(mu));       // non-negative floats order like their bit patterns
        atomicMax(tmax + 2 * t + 1, __float_as_uint(mw));
    }
}

__global__ void __launch_bounds__(256)
    percent_apply_kernel(const OptChunk* __restrict__ tab, const int* __restrict__ tid, const float* __restrict__ m,
                         const float* __restrict__ v, const unsigned int* __restrict__ tmax, float lr, float adj,
                         float bc2s, float eps) {
    pdl_enter();
    const OptChunk c = tab[blockIdx.x];
    const int t = tid[blockIdx.x];
    const float max_u = __uint_as_float(tmax[2 * t]), max_w = __uint_as_float(tmax[2 * t + 1]);
    float scale = __fdiv_rn(__fmul_rn(lr, max_w), __fadd_rn(max_u, eps));
    if (lr > scale) scale = lr;   // Python max(tensor, lr): lr only when it is strictly larger
    const float* mp = m + c.off;
    const float* vp = v + c.off;
    for (int i = threadIdx.x; i < c.n; i += 256) {
        const float u = percent_update(mp[i], vp[i], adj, bc2s, eps);
        c.p[i] = __fsub_rn(c.p[i], __fmul_rn(u, scale));
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200sd_percent_opt_step(const void* chunks, const int32_t* tensor_of_chunk, int32_t n_chunks,
                                       const float* g_base, float* m_base, float* v_base, uint32_t* tensor_max,
                                       float lr, float beta1, float beta2, float w1, float w2, float adj, float bc2_sqrt,
                                       float eps, void* stream) {
    B200_REQUIRE(chunks && tensor_of_chunk && g_base && m_base && v_base && tensor_max && n_chunks > 0,
                 "percent_opt_step: bad arguments");
    launch_k(percent_moments_kernel, n_chunks, 256, 0, as_stream(stream), static_cast<const OptChunk*>(chunks),
             tensor_of_chunk, g_base, m_base, v_base, tensor_max, beta1, beta2, w1, w2, adj, bc2_sqrt, eps);
    B200_CHECK_LAUNCH("percent_moments");
    launch_k(percent_apply_kernel, n_chunks, 256, 0, as_stream(stream), static_cast<const OptChunk*>(chunks),
             tensor_of_chunk, m_base, v_base, tensor_max, lr, adj, bc2_sqrt, eps);
    B200_CHECK_LAUNCH("percent_apply");
    return B200SD_OK;
}

extern "C" int b200sd_opt_chunk_size(void) { return static_cast<int>(sizeof(OptChunk)); }
extern "C" int b200sd_opt_chunk_elems(void) { return OPT_CHUNK; }

extern "C" int b200sd_grad_sqnorm(const void* chunks, int32_t n_chunks, const float* g_base, float max_norm,
                                  int32_t check_finite, float* partial, float* scalars, uint32_t* counter,
                                  void* stream) {
    B200_REQUIRE(chunks && g_base && partial && scalars && counter && n_chunks > 0, "grad_sqnorm: bad arguments");
    launch_k(grad_sqnorm_kernel, n_chunks, 256, 0, as_stream(stream), static_cast<const OptChunk*>(chunks), n_chunks,
             g_base, max_norm, check_finite, partial, scalars, counter);
    B200_CHECK_LAUNCH("grad_sqnorm");
    return B200SD_OK;
}

extern "C" int b200sd_adam_ema_step(const void* chunks, int32_t n_chunks, const float* g_base, float* m_base,
                                    float* v_base, const float* scalars, float lr, float beta1, float beta2, float eps,
                                    float weight_decay, float ema_decay, void* stream) {
    B200_REQUIRE(chunks && g_base && m_base && v_base && scalars && n_chunks > 0, "adam_ema_step: bad arguments");
    B200_REQUIRE(beta1 >= 0.f && beta1 < 1.f && beta2 >= 0.f && beta2 < 1.f && eps >= 0.f, "adam_ema_step: bad hyper-parameters");
    launch_k(adam_ema_kernel, n_chunks, 256, 0, as_stream(stream), static_cast<const OptChunk*>(chunks), g_base, m_base,
             v_base, scalars, lr, beta1, beta2, eps, weight_decay, ema_decay);
    B200_CHECK_LAUNCH("adam_ema_step");
    return B200SD_OK;
}

// DDPM scheduler kernels and the sinusoidal time embedding (a4).
// HBM-bound elementwise work: one launch per call, float4 accesses, grid = k * SM count (grid-stride).
// Arithmetic is written with explicitly rounded intrinsics (__fmul_rn, __fadd_rn, __fdiv_rn, __fsqrt_rn) so the
// compiler cannot contract a*b+c into an FMA: the reference evaluates every op as a separate fp32 torch kernel
// (scheduler/linear_noise_scheduler.py:47-48,59-78) and we match it bit for bit.
#include "host.h"
#include <cuda_bf16.h>
#include <curand_kernel.h>

namespace b200 {

struct PrevCoef {
    float s1m, sqrt_acp, beta, sqrt_alpha, sigma;
    int t;
};

__device__ __forceinline__ PrevCoef prev_coef(const float* betas, const float* alphas, const float* acp,
                                              const float* s1m, int t) {
    PrevCoef c;
    c.t = t;
    c.s1m = s1m[t];
    c.sqrt_acp = __fsqrt_rn(acp[t]);
    c.beta = betas[t];
    c.sqrt_alpha = __fsqrt_rn(alphas[t]);
    c.sigma = 0.f;
    if (t > 0) {
        float var = __fdiv_rn(__fsub_rn(1.f, acp[t - 1]), __fsub_rn(1.f, acp[t]));
        var = __fmul_rn(var, c.beta);
        c.sigma = __fsqrt_rn(var);  // torch: variance ** 0.5 lowers to sqrt
    }
    return c;
}

__device__ __forceinline__ void prev_one(const PrevCoef& c, float xt, float e, float z, float& xp, float& x0) {
    float a = __fdiv_rn(__fsub_rn(xt, __fmul_rn(c.s1m, e)), c.sqrt_acp);
    x0 = fminf(fmaxf(a, -1.f), 1.f);
    float m = __fsub_rn(xt, __fdiv_rn(__fmul_rn(c.beta, e), c.s1m));
    m = __fdiv_rn(m, c.sqrt_alpha);
    xp = (c.t == 0) ? m : __fadd_rn(m, __fmul_rn(c.sigma, z));
}

template <bool VEC>
__global__ void __launch_bounds__(256) add_noise_kernel(const float* __restrict__ x0, const float* __restrict__ nz,
                                                        const int64_t* __restrict__ t, int64_t t_len,
                                                        const float* __restrict__ sa, const float* __restrict__ sb,
                                                        int T, float* __restrict__ out, int64_t batch,
                                                        int64_t per_sample) {
    pdl_enter();
    const int64_t total = batch * per_sample;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    if (VEC) {
        const int64_t nv = total >> 2, pv = per_sample >> 2;
        for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nv; i += stride) {
            int64_t b = i / pv;
            int64_t tt = t[t_len == 1 ? 0 : b];
            tt = tt < 0 ? tt + T : tt;  // torch indexing wraps negatives
            float a = sa[tt], c = sb[tt];
            float4 x = reinterpret_cast<const float4*>(x0)[i];
            float4 n = reinterpret_cast<const float4*>(nz)[i];
            float4 o;
            o.x = __fadd_rn(__fmul_rn(a, x.x), __fmul_rn(c, n.x));
            o.y = __fadd_rn(__fmul_rn(a, x.y), __fmul_rn(c, n.y));
            o.z = __fadd_rn(__fmul_rn(a, x.z), __fmul_rn(c, n.z));
            o.w = __fadd_rn(__fmul_rn(a, x.w), __fmul_rn(c, n.w));
            reinterpret_cast<float4*>(out)[i] = o;
        }
    } else {
        for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
            int64_t b = i / per_sample;
            int64_t tt = t[t_len == 1 ? 0 : b];
            tt = tt < 0 ? tt + T : tt;
            out[i] = __fadd_rn(__fmul_rn(sa[tt], x0[i]), __fmul_rn(sb[tt], nz[i]));
        }
    }
}

template <bool VEC>
__global__ void __launch_bounds__(256)
    prev_step_kernel(const float* __restrict__ xt, const float* __restrict__ eps, const float* __restrict__ eps_u,
                     float cfg, const float* __restrict__ z, const int64_t* __restrict__ t_dev, int64_t t_host,
                     const float* __restrict__ betas, const float* __restrict__ alphas,
                     const float* __restrict__ acp, const float* __restrict__ s1m, int T,
                     float* __restrict__ x_prev, float* __restrict__ x0, int64_t n, uint64_t seed,
                     uint64_t offset, const uint64_t* __restrict__ rng_state) {
    pdl_enter();
    int64_t tt = t_dev ? *t_dev : t_host;
    tt = tt < 0 ? tt + T : tt;
    if (t_dev) tt = tt < 0 ? 0 : (tt >= T ? T - 1 : tt);   // a device-side t is not range-checked by the host: clamp
    if (rng_state) {
        // launch arguments are frozen inside a CUDA graph: the Philox key of a replayed loop comes from device memory,
        // {seed, call index}; each sampling call gets its own key, hence its own noise trajectory
        seed = rng_state[0] + 0x9E3779B97F4A7C15ull * rng_state[1];
    }
    const PrevCoef c = prev_coef(betas, alphas, acp, s1m, static_cast<int>(tt));
    const int64_t gid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    curandStatePhilox4_32_10_t rng;
    const bool draw = (z == nullptr) && c.t > 0;
    if (draw) {
        // when t is read from the device the launch arguments are frozen inside a CUDA graph, so the stream position
        // is derived from the timestep: every replay draws a fresh, non-overlapping block of the Philox sequence
        // 32-bit outputs one thread consumes per launch: a float4 path iteration draws 4 (curand_normal4), a scalar
        // iteration 1 on average (curand_normal draws 2 and keeps the second Box-Muller value) but 2 at worst
        const uint64_t iters = static_cast<uint64_t>(VEC ? ((n >> 2) + stride - 1) / stride : (n + stride - 1) / stride);
        const uint64_t per_call = (VEC ? 4ull : 2ull) * iters + 4ull;
        const uint64_t step = t_dev ? static_cast<uint64_t>(T - 1 - c.t) : 0ull;
        curand_init(seed, gid, offset + step * per_call, &rng);
    }
    if (VEC) {
        const int64_t nv = n >> 2;
        for (int64_t i = gid; i < nv; i += stride) {
            float4 x = reinterpret_cast<const float4*>(xt)[i];
            float4 e = reinterpret_cast<const float4*>(eps)[i];
            if (eps_u) {
                float4 u = reinterpret_cast<const float4*>(eps_u)[i];
                e.x = __fadd_rn(u.x, __fmul_rn(cfg, __fsub_rn(e.x, u.x)));
                e.y = __fadd_rn(u.y, __fmul_rn(cfg, __fsub_rn(e.y, u.y)));
                e.z = __fadd_rn(u.z, __fmul_rn(cfg, __fsub_rn(e.z, u.z)));
                e.w = __fadd_rn(u.w, __fmul_rn(cfg, __fsub_rn(e.w, u.w)));
            }
            float4 zz = make_float4(0.f, 0.f, 0.f, 0.f);
            if (z)
                zz = reinterpret_cast<const float4*>(z)[i];
            else if (draw)
                zz = curand_normal4(&rng);
            float4 p, o;
            prev_one(c, x.x, e.x, zz.x, p.x, o.x);
            prev_one(c, x.y, e.y, zz.y, p.y, o.y);
            prev_one(c, x.z, e.z, zz.z, p.z, o.z);
            prev_one(c, x.w, e.w, zz.w, p.w, o.w);
            reinterpret_cast<float4*>(x_prev)[i] = p;
            reinterpret_cast<float4*>(x0)[i] = o;
        }
    } else {
        for (int64_t i = gid; i < n; i += stride) {
            float e = eps[i];
            if (eps_u) e = __fadd_rn(eps_u[i], __fmul_rn(cfg, __fsub_rn(e, eps_u[i])));
            float zz = z ? z[i] : (draw ? curand_normal(&rng) : 0.f);
            float p, o;
            prev_one(c, xt[i], e, zz, p, o);
            x_prev[i] = p;
            x0[i] = o;
        }
    }
}

template <typename OutT>
__global__ void time_embedding_kernel(const int64_t* __restrict__ t, int64_t t_len, OutT* __restrict__ out,
                                      int64_t batch, int dim) {
    pdl_enter();
    const int half = dim >> 1;
    const int64_t total = batch * half;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int64_t b = i / half;
        const int j = static_cast<int>(i - b * half);
        // blocks.py:16-18: factor = 10000 ** (arange(half) / half), fp32 throughout
        const float factor = powf(10000.f, __fdiv_rn(static_cast<float>(j), static_cast<float>(half)));
        const float arg = __fdiv_rn(static_cast<float>(t[t_len == 1 ? 0 : b]), factor);
        out[b * dim + j] = static_cast<OutT>(sinf(arg));
        out[b * dim + half + j] = static_cast<OutT>(cosf(arg));
    }
}

__global__ void advance_t_kernel(int64_t* t, int64_t n, int64_t delta) {
    pdl_enter();
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) t[i] += delta;
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

static inline int ew_grid(int64_t work_items, int threads) {
    int64_t blocks = (work_items + threads - 1) / threads;
    int64_t cap = static_cast<int64_t>(sm_count()) * 8;  // 8 resident 256-thread CTAs per SM
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

// DDIM / DDPM sampler update (scheduler/linear_noise_scheduler.py:106-131 DDPMSampler, :185-204 DDIMSampler):
// out = (c1 * x + c2 * eps) + sigma * z, every product and sum rounded separately in the reference's order (no fma),
// so that the result is bit-identical to the reference's chain of elementwise torch kernels.
__global__ void sampler_step_kernel(const float* __restrict__ x, const float* __restrict__ eps,
                                    const float* __restrict__ z, float c1, float c2, float sigma,
                                    float* __restrict__ out, long long n) {
    pdl_enter();
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
        float r = __fadd_rn(__fmul_rn(c1, x[i]), __fmul_rn(c2, eps[i]));
        if (z) r = __fadd_rn(r, __fmul_rn(sigma, z[i]));
        out[i] = r;
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200sd_sched_add_noise(const float* x0, const float* noise, const int64_t* t, int64_t t_len,
                                      const float* sqrt_acp, const float* sqrt_1m_acp, int num_timesteps,
                                      float* out, int64_t batch, int64_t per_sample, void* stream) {
    B200_REQUIRE(x0 && noise && t && sqrt_acp && sqrt_1m_acp && out, "add_noise: null pointer");
    B200_REQUIRE(batch >= 0 && per_sample >= 0, "add_noise: negative size");
    B200_REQUIRE(t_len == 1 || t_len == batch, "add_noise: t has %lld entries for batch %lld", (long long)t_len,
                 (long long)batch);
    if (batch * per_sample == 0) return B200SD_OK;
    const bool vec = (per_sample % 4 == 0) && aligned16(x0) && aligned16(noise) && aligned16(out);
    const int64_t items = vec ? batch * per_sample / 4 : batch * per_sample;
    const int grid = ew_grid(items, 256);
    if (vec)
        launch_k(add_noise_kernel<true>, grid, 256, 0, as_stream(stream), x0, noise, t, t_len, sqrt_acp, sqrt_1m_acp,
                                                                      num_timesteps, out, batch, per_sample);
    else
        launch_k(add_noise_kernel<false>, grid, 256, 0, as_stream(stream), x0, noise, t, t_len, sqrt_acp, sqrt_1m_acp,
                                                                       num_timesteps, out, batch, per_sample);
    B200_CHECK_LAUNCH("add_noise");
    return B200SD_OK;
}

extern "C" int b200sd_sched_prev_step(const float* xt, const float* eps, const float* eps_uncond, float cfg_scale,
                                      const float* z, const int64_t* t_dev, int64_t t_host, const float* betas,
                                      const float* alphas, const float* acp, const float* sqrt_1m_acp,
                                      int num_timesteps, float* x_prev, float* x0, int64_t n, uint64_t seed,
                                      uint64_t offset, const uint64_t* rng_state, void* stream) {
    B200_REQUIRE(xt && eps && betas && alphas && acp && sqrt_1m_acp && x_prev && x0, "prev_step: null pointer");
    B200_REQUIRE(n >= 0, "prev_step: negative size");
    if (!t_dev)
        B200_REQUIRE(t_host >= -num_timesteps && t_host < num_timesteps, "prev_step: t=%lld out of range [0,%d)",
                     (long long)t_host, num_timesteps);
    if (n == 0) return B200SD_OK;
    const bool vec = (n % 4 == 0) && aligned16(xt) && aligned16(eps) && aligned16(x_prev) && aligned16(x0) &&
                     (!z || aligned16(z)) && (!eps_uncond || aligned16(eps_uncond));
    const int64_t items = vec ? n / 4 : n;
    const int grid = ew_grid(items, 256);
    if (vec)
        launch_k(prev_step_kernel<true>, grid, 256, 0, as_stream(stream), xt, eps, eps_uncond, cfg_scale, z, t_dev, t_host,
                                                                      betas, alphas, acp, sqrt_1m_acp, num_timesteps,
                                                                      x_prev, x0, n, seed, offset, rng_state);
    else
        launch_k(prev_step_kernel<false>, grid, 256, 0, as_stream(stream), xt, eps, eps_uncond, cfg_scale, z, t_dev,
                                                                       t_host, betas, alphas, acp, sqrt_1m_acp,
                                                                       num_timesteps, x_prev, x0, n, seed, offset, rng_state);
    B200_CHECK_LAUNCH("prev_step");
    return B200SD_OK;
}

extern "C" int b200sd_time_embedding(const int64_t* t, int64_t t_len, void* out, int out_dtype, int64_t batch,
                                     int dim, void* stream) {
    B200_REQUIRE(t && out, "time_embedding: null pointer");
    B200_REQUIRE(dim > 0 && dim % 2 == 0, "time embedding dimension must be divisible by 2");
    B200_REQUIRE(t_len == 1 || t_len == batch, "time_embedding: t has %lld entries for batch %lld",
                 (long long)t_len, (long long)batch);
    if (batch == 0) return B200SD_OK;
    const int grid = ew_grid(batch * (dim / 2), 128);
    if (out_dtype == B200SD_DT_BF16)
        launch_k(time_embedding_kernel<__nv_bfloat16>, grid, 128, 0, as_stream(stream), t, t_len, static_cast<__nv_bfloat16*>(out), batch, dim);
    else if (out_dtype == B200SD_DT_F32)
        launch_k(time_embedding_kernel<float>, grid, 128, 0, as_stream(stream), t, t_len, static_cast<float*>(out), batch, dim);
    else {
        set_error("time_embedding: bad dtype %d", out_dtype);
        return B200SD_ERR_ARG;
    }
    B200_CHECK_LAUNCH("time_embedding");
    return B200SD_OK;
}

extern "C" int b200sd_advance_t(int64_t* t, int64_t t_len, int64_t delta, void* stream) {
    B200_REQUIRE(t && t_len >= 0, "advance_t: bad arguments");
    if (t_len == 0) return B200SD_OK;
    launch_k(advance_t_kernel, static_cast<int>((t_len + 127) / 128), 128, 0, as_stream(stream), t, t_len, delta);
    B200_CHECK_LAUNCH("advance_t");
    return B200SD_OK;
}

extern "C" int b200sd_sampler_step(const float* x, const float* eps, const float* z, float c1, float c2, float sigma,
                                   float* out, int64_t n, void* stream) {
    B200_REQUIRE(x && eps && out && n >= 0, "sampler_step: bad arguments");
    if (n == 0) return B200SD_OK;
    launch_k(sampler_step_kernel, ew_grid(n, 256), 256, 0, as_stream(stream), x, eps, z, c1, c2, sigma, out,
             static_cast<long long>(n));
    B200_CHECK_LAUNCH("sampler_step");
    return B200SD_OK;
}

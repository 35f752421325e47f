// GroupNorm (+SiLU/ReLU) over NHWC bf16 activations. HBM-bound: one read + one write per element.
// A CTA owns one image and a slab of `gpc` adjacent groups (cw = gpc * C/G channels); the slab is staged in shared
// memory when it fits so that the three passes (mean, centred variance, normalise) touch HBM once for the read
// and once for the write. Mean/variance use the two-pass form in fp32 (no E[x^2]-E[x]^2 cancellation).
#include "host.h"
#include "ptx.cuh"
#include <cstdlib>
#include <cuda_bf16.h>

namespace b200 {

__device__ __forceinline__ float act_apply(float v, int act) {
    if (act == B200SD_ACT_SILU) return silu_mufu(v);
    if (act == B200SD_ACT_RELU) return fmaxf(v, 0.f);
    return v;
}

// VEC = channels per thread item (2 -> bf16x2 accesses; 1 -> scalar, for odd channels-per-group)
template <int VEC>
__global__ void __launch_bounds__(512)
    groupnorm_kernel(const __nv_bfloat16* __restrict__ x, long long ld_x, __nv_bfloat16* __restrict__ y,
                     long long ld_y, const float* __restrict__ gamma, const float* __restrict__ beta, long long hw,
                     int c, int cpg, int gpc, float eps, int act, int cached, int rows_per_pass) {
    pdl_enter();
    extern __shared__ __align__(16) uint8_t gn_smem[];
    __shared__ float s_sum[64];
    __shared__ float s_mean[64];
    __shared__ float s_rstd[64];

    const int cw = gpc * cpg;
    const int vpr = cw / VEC;  // vector items per pixel row of the slab
    const int chunks = c / cw;
    const int img = blockIdx.x / chunks;
    const int c0 = (blockIdx.x % chunks) * cw;
    const int tid = threadIdx.x;
    const bool active = tid < vpr * rows_per_pass;
    const int v = tid % vpr;           // constant per thread: blockDim is used in whole rows
    const int row0 = tid / vpr;
    const int g_local = (v * VEC) / cpg;  // VEC divides cpg, so an item never straddles groups
    const __nv_bfloat16* xs = x + (long long)img * hw * ld_x + c0 + v * VEC;
    __nv_bfloat16* ys = y + (long long)img * hw * ld_y + c0 + v * VEC;
    __nv_bfloat16* cache = reinterpret_cast<__nv_bfloat16*>(gn_smem);
    const float inv_cnt = 1.f / (float)(hw * cpg);

    if (tid < 64) s_sum[tid] = 0.f;
    __syncthreads();

    // pass 1: sum (and fill the cache)
    float acc = 0.f;
    if (active) {
        for (long long r = row0; r < hw; r += rows_per_pass) {
            if (VEC == 2) {
                __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(xs + r * ld_x);
                if (cached) *reinterpret_cast<__nv_bfloat162*>(cache + r * cw + v * 2) = h;
                float2 f = __bfloat1622float2(h);
                acc += f.x + f.y;
            } else {
                __nv_bfloat16 h = xs[r * ld_x];
                if (cached) cache[r * cw + v] = h;
                acc += __bfloat162float(h);
            }
        }
        atomicAdd(&s_sum[g_local], acc);
    }
    __syncthreads();
    if (tid < gpc) s_mean[tid] = s_sum[tid] * inv_cnt;
    __syncthreads();
    if (tid < 64) s_sum[tid] = 0.f;
    __syncthreads();

    // pass 2: centred sum of squares
    if (active) {
        const float m = s_mean[g_local];
        acc = 0.f;
        for (long long r = row0; r < hw; r += rows_per_pass) {
            if (VEC == 2) {
                __nv_bfloat162 h = cached ? *reinterpret_cast<const __nv_bfloat162*>(cache + r * cw + v * 2)
                                          : *reinterpret_cast<const __nv_bfloat162*>(xs + r * ld_x);
                float2 f = __bfloat1622float2(h);
                acc += (f.x - m) * (f.x - m) + (f.y - m) * (f.y - m);
            } else {
                float f = __bfloat162float(cached ? cache[r * cw + v] : xs[r * ld_x]);
                acc += (f - m) * (f - m);
            }
        }
        atomicAdd(&s_sum[g_local], acc);
    }
    __syncthreads();
    if (tid < gpc) s_rstd[tid] = rsqrtf(s_sum[tid] * inv_cnt + eps);
    __syncthreads();

    // pass 3: normalise, affine, activation, store
    if (active) {
        const float m = s_mean[g_local], rs = s_rstd[g_local];
        const int ch = c0 + v * VEC;
        const float g0 = gamma ? gamma[ch] : 1.f, b0 = beta ? beta[ch] : 0.f;
        const float sc0 = rs * g0, sh0 = b0 - m * rs * g0;
        float sc1 = 0.f, sh1 = 0.f;
        if (VEC == 2) {
            const float g1 = gamma ? gamma[ch + 1] : 1.f, b1 = beta ? beta[ch + 1] : 0.f;
            sc1 = rs * g1, sh1 = b1 - m * rs * g1;
        }
        for (long long r = row0; r < hw; r += rows_per_pass) {
            if (VEC == 2) {
                __nv_bfloat162 h = cached ? *reinterpret_cast<const __nv_bfloat162*>(cache + r * cw + v * 2)
                                          : *reinterpret_cast<const __nv_bfloat162*>(xs + r * ld_x);
                float2 f = __bfloat1622float2(h);
                f.x = act_apply(f.x * sc0 + sh0, act);
                f.y = act_apply(f.y * sc1 + sh1, act);
                *reinterpret_cast<__nv_bfloat162*>(ys + r * ld_y) = __floats2bfloat162_rn(f.x, f.y);
            } else {
                float f = __bfloat162float(cached ? cache[r * cw + v] : xs[r * ld_x]);
                ys[r * ld_y] = __float2bfloat16_rn(act_apply(f * sc0 + sh0, act));
            }
        }
    }
}


// ---------------------------------------------------------------------------------------------------------------
// GroupNorm APPLY on precomputed partial statistics (written by the producing igemm's epilogue, b200sd.h `stats`):
// the tensor is read once and written once, fully coalesced 16-byte accesses, any number of CTAs per image.
// Prologue per CTA: fold the P partial (sum, sumsq) slots of every channel in fixed order (bit-reproducible), reduce
// channels to groups, turn (mean, rstd, gamma, beta) into a per-channel scale/shift held in registers.
struct StatSeg {
    const float* ptr;
    int p, ld;
};

// SiLU / ReLU with the activation fixed at compile time. SiLU = v * rcp(1 + ex2(-v log2 e)): two MUFU ops and three
// FP32 ops per element (silu_mufu, ptx.cuh; an IEEE division here was ~20 instructions per element and made the kernel
// issue-bound, and even __fdividef / __expf cost 3.5 more than the bare MUFU pair for the same bits).
// GN_ACT_SILU_TANH: SiLU = h + h * tanh(h), h = v / 2 — ONE MUFU op (tanh.approx.f32, relative error 2^-11 on tanh,
// i.e. at most |v| * 2.4e-4 absolute on the result: below the bf16 rounding of the stored output for |v| < 8) instead of
// two (ex2 + rcp).
constexpr int GN_ACT_SILU_TANH = 100;
template <int ACT>
__device__ __forceinline__ float act_fast(float v) {
    if (ACT == GN_ACT_SILU_TANH) {
        const float h = 0.5f * v;
        float t;
        asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
        return fmaf(h, t, h);
    }
    if (ACT == B200SD_ACT_SILU) return silu_mufu(v);
    if (ACT == B200SD_ACT_RELU) return fmaxf(v, 0.f);
    return v;
}

constexpr int GN_U = 4;  // 16-byte loads each thread keeps in flight per batch

template <int ACT>
__global__ void __launch_bounds__(256, 4)
    gn_apply_kernel(const __nv_bfloat16* __restrict__ x, long long ld_x, __nv_bfloat16* __restrict__ y, long long ld_y,
                    StatSeg s0, int c_split, StatSeg s1, const float* __restrict__ gamma,
                    const float* __restrict__ beta, long long hw, int c, int groups, float eps,
                    int rows_per_block, int rpp) {
    pdl_enter();
    extern __shared__ float gn_sm[];
    float* sc = gn_sm;          // [c]  channel sums, then scale
    float* sh = gn_sm + c;      // [c]  channel sums of squares, then shift
    float* gs = gn_sm + 2 * c;  // [groups][2] mean, rstd
    const int img = blockIdx.y;
    const int tid = threadIdx.x, nt = blockDim.x;
    const int cpg = c / groups;
    const int cv = c >> 3;
    const int v = tid % cv, r0 = tid / cv;
    const bool active = r0 < rpp;
    const long long row_begin = static_cast<long long>(blockIdx.x) * rows_per_block;
    const int nrows = static_cast<int>(hw - row_begin < rows_per_block ? hw - row_begin : rows_per_block);
    const __nv_bfloat16* px = x + (static_cast<long long>(img) * hw + row_begin + r0) * ld_x + v * 8;
    __nv_bfloat16* py = y + (static_cast<long long>(img) * hw + row_begin + r0) * ld_y + v * 8;
    const long long sx = static_cast<long long>(rpp) * ld_x, sy = static_cast<long long>(rpp) * ld_y;

    // first batch of rows goes in flight before the statistics prologue: its HBM latency hides the fold below
    uint4 cur[GN_U];
    int i = r0;  // row inside this CTA's block
#pragma unroll
    for (int j = 0; j < GN_U; ++j)
        if (active && i + j * rpp < nrows) cur[j] = *reinterpret_cast<const uint4*>(px + j * sx);

    for (int ch = tid; ch < c; ch += nt) {
        const bool first = ch < c_split;
        const float* sp = first ? s0.ptr : s1.ptr;
        const int np = first ? s0.p : s1.p, sld = first ? s0.ld : s1.ld;
        const int cc = first ? ch : ch - c_split;
        const float* p = sp + (static_cast<long long>(img) * np * sld + cc) * 2;
        float a = 0.f, b = 0.f;
        for (int k = 0; k < np; ++k) {
            const float2 t = *reinterpret_cast<const float2*>(p + static_cast<long long>(k) * sld * 2);
            a += t.x, b += t.y;
        }
        sc[ch] = a, sh[ch] = b;
    }
    __syncthreads();
    const float inv_cnt = 1.f / (static_cast<float>(hw) * cpg);
    for (int g = tid; g < groups; g += nt) {
        float a = 0.f, b = 0.f;
        for (int k = 0; k < cpg; ++k) a += sc[g * cpg + k], b += sh[g * cpg + k];
        const float mean = a * inv_cnt;
        const float var = fmaxf(b * inv_cnt - mean * mean, 0.f);
        gs[2 * g] = mean, gs[2 * g + 1] = rsqrtf(var + eps);
    }
    __syncthreads();
    for (int ch = tid; ch < c; ch += nt) {
        const int g = ch / cpg;
        const float ga = gamma ? gamma[ch] : 1.f, be = beta ? beta[ch] : 0.f;
        const float scale = gs[2 * g + 1] * ga;
        sc[ch] = scale, sh[ch] = be - gs[2 * g] * scale;
    }
    __syncthreads();
    if (!active) return;
    float scl[8], sft[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) scl[j] = sc[v * 8 + j], sft[j] = sh[v * 8 + j];
    const int step = GN_U * rpp;
    for (;;) {
#pragma unroll
        for (int j = 0; j < GN_U; ++j) {
            if (i + j * rpp < nrows) {
                const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&cur[j]);
                uint4 o;
                __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    float2 f = __bfloat1622float2(h[e]);
                    f.x = act_fast<ACT>(fmaf(f.x, scl[2 * e], sft[2 * e]));
                    f.y = act_fast<ACT>(fmaf(f.y, scl[2 * e + 1], sft[2 * e + 1]));
                    oh[e] = __floats2bfloat162_rn(f.x, f.y);
                }
                *reinterpret_cast<uint4*>(py + j * sy) = o;
            }
        }
        i += step, px += GN_U * sx, py += GN_U * sy;
        if (i >= nrows) break;
#pragma unroll
        for (int j = 0; j < GN_U; ++j)
            if (i + j * rpp < nrows) cur[j] = *reinterpret_cast<const uint4*>(px + j * sx);
    }
}


// ---------------------------------------------------------------------------------------------------------------
// GroupNorm (+activation) BACKWARD. With xhat = (x - mean) * rstd, z = xhat * gamma + beta, y = act(z):
//   dz = dy * act'(z);   A[c] = sum_rows dz,  B[c] = sum_rows dz * xhat   (per image; dbeta = sum_n A, dgamma = sum_n B)
//   S1[g] = sum_{c in g} gamma[c] A[c],  S2[g] = sum_{c in g} gamma[c] B[c]
//   dx = rstd * (gamma * dz - (S1 + xhat * S2) / (hw * cpg))  (+ an optional extra gradient of x, e.g. the shortcut's)
// Pass 1 (gn_bwd_reduce) writes per-(image, row block) partial A/B; pass 2 (gn_bwd_apply) folds them in fixed order
// and streams dx. Mean / rstd are rebuilt from the forward's partial statistics exactly as gn_apply does.
__device__ __forceinline__ float act_grad(float z, int act) {
    if (act == B200SD_ACT_SILU) {
        const float s = __fdividef(1.f, 1.f + __expf(-z));
        return s * (1.f + z * (1.f - s));
    }
    if (act == B200SD_ACT_RELU) return z > 0.f ? 1.f : 0.f;
    return 1.f;
}

// shared prologue: per-channel mean/rstd-derived (scale = rstd*gamma, shift = beta - mean*scale) plus per-channel
// mean and rstd themselves, from the forward's partial statistics
__device__ __forceinline__ void gn_channel_params(float* sc, float* sh, float* mu, float* rs, float* gs, StatSeg s0,
                                                  int c_split, StatSeg s1, const float* gamma, const float* beta,
                                                  long long hw, int c, int groups, float eps, int img) {
    const int tid = threadIdx.x, nt = blockDim.x;
    const int cpg = c / groups;
    for (int ch = tid; ch < c; ch += nt) {
        const bool first = ch < c_split;
        const StatSeg sg = first ? s0 : s1;
        const int cc = first ? ch : ch - c_split;
        const float* p = sg.ptr + (static_cast<long long>(img) * sg.p * sg.ld + cc) * 2;
        float a = 0.f, b = 0.f;
        for (int k = 0; k < sg.p; ++k) {
            const float2 v = *reinterpret_cast<const float2*>(p + static_cast<long long>(k) * sg.ld * 2);
            a += v.x, b += v.y;
        }
        sc[ch] = a, sh[ch] = b;
    }
    __syncthreads();
    const float inv_cnt = 1.f / (static_cast<float>(hw) * cpg);
    for (int g = tid; g < groups; g += nt) {
        float a = 0.f, b = 0.f;
        for (int k = 0; k < cpg; ++k) a += sc[g * cpg + k], b += sh[g * cpg + k];
        const float mean = a * inv_cnt;
        const float var = fmaxf(b * inv_cnt - mean * mean, 0.f);
        gs[2 * g] = mean, gs[2 * g + 1] = rsqrtf(var + eps);
    }
    __syncthreads();
    for (int ch = tid; ch < c; ch += nt) {
        const int g = ch / cpg;
        const float ga = gamma ? gamma[ch] : 1.f, be = beta ? beta[ch] : 0.f;
        const float scale = gs[2 * g + 1] * ga;
        sc[ch] = scale, sh[ch] = be - gs[2 * g] * scale;
        mu[ch] = gs[2 * g], rs[ch] = gs[2 * g + 1];
    }
    __syncthreads();
}

__global__ void __launch_bounds__(256)
    gn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ x, long long ld_x, const __nv_bfloat16* __restrict__ dy,
                         long long ld_dy, StatSeg s0, int c_split, StatSeg s1, const float* __restrict__ gamma,
                         const float* __restrict__ beta, long long hw, int c, int groups, float eps, int act,
                         int rows_per_block, int rpp, float* __restrict__ part /* [n][gridDim.x][c][2] */) {
    pdl_enter();
    extern __shared__ float gn_sm[];
    float* sc = gn_sm;
    float* sh = sc + c;
    float* mu = sh + c;
    float* rs = mu + c;
    float* gs = rs + c;                 // [groups][2]
    float* red = gs + 2 * groups;       // [rpp][c][2]
    const int img = blockIdx.y, tid = threadIdx.x;
    gn_channel_params(sc, sh, mu, rs, gs, s0, c_split, s1, gamma, beta, hw, c, groups, eps, img);
    const int cv = c >> 3;
    const int v = tid % cv, r0 = tid / cv;
    float a[8], b[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = b[j] = 0.f;
    if (r0 < rpp) {
        float scl[8], sft[8], m[8], r[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) scl[j] = sc[v * 8 + j], sft[j] = sh[v * 8 + j], m[j] = mu[v * 8 + j], r[j] = rs[v * 8 + j];
        const long long row_begin = static_cast<long long>(blockIdx.x) * rows_per_block;
        long long row_end = row_begin + rows_per_block;
        if (row_end > hw) row_end = hw;
        const __nv_bfloat16* xs = x + static_cast<long long>(img) * hw * ld_x + v * 8;
        const __nv_bfloat16* ds = dy + static_cast<long long>(img) * hw * ld_dy + v * 8;
        for (long long rr = row_begin + r0; rr < row_end; rr += rpp) {
            const uint4 ux = *reinterpret_cast<const uint4*>(xs + rr * ld_x);
            const uint4 ud = *reinterpret_cast<const uint4*>(ds + rr * ld_dy);
            const __nv_bfloat162* hx = reinterpret_cast<const __nv_bfloat162*>(&ux);
            const __nv_bfloat162* hd = reinterpret_cast<const __nv_bfloat162*>(&ud);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 fx = __bfloat1622float2(hx[j]), fd = __bfloat1622float2(hd[j]);
                const float z0 = fx.x * scl[2 * j] + sft[2 * j], z1 = fx.y * scl[2 * j + 1] + sft[2 * j + 1];
                const float d0 = fd.x * act_grad(z0, act), d1 = fd.y * act_grad(z1, act);
                a[2 * j] += d0, a[2 * j + 1] += d1;
                b[2 * j] += d0 * (fx.x - m[2 * j]) * r[2 * j], b[2 * j + 1] += d1 * (fx.y - m[2 * j + 1]) * r[2 * j + 1];
            }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) red[(r0 * c + v * 8 + j) * 2] = a[j], red[(r0 * c + v * 8 + j) * 2 + 1] = b[j];
    }
    __syncthreads();
    float* dst = part + (static_cast<long long>(img) * gridDim.x + blockIdx.x) * c * 2;
    for (int i = tid; i < 2 * c; i += blockDim.x) {
        float s = 0.f;
        for (int k = 0; k < rpp; ++k) s += red[k * c * 2 + i];
        dst[i] = s;
    }
}

// Fixed-order tail reduction by the LAST CTA of a launch (no second kernel, no atomics on the data): every CTA
// publishes its partials, bumps `counter`; whichever CTA sees total-1 owns the tail and leaves the counter at 0 for the
// next launch. The summation order does not depend on which CTA that is.
__device__ __forceinline__ bool last_cta_arrives(unsigned int* counter, unsigned int total) {
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

// out[b][c] = sum over r of in[b][r][c * col_stride], r ascending in four interleaved chains (fixed order)
__device__ __forceinline__ void tail_reduce(const float* in, long long batch, long long rows, long long cols,
                                            long long col_stride, float* __restrict__ out) {
    const long long total = batch * cols;
    for (long long i = threadIdx.x; i < total; i += blockDim.x) {
        const long long b = i / cols, cidx = i - b * cols;
        const float* p = in + b * rows * cols * col_stride + cidx * col_stride;
        const long long step = cols * col_stride;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        long long r = 0;
        for (; r + 4 <= rows; r += 4) {
            s0 += __ldcg(p + r * step), s1 += __ldcg(p + (r + 1) * step);
            s2 += __ldcg(p + (r + 2) * step), s3 += __ldcg(p + (r + 3) * step);
        }
        for (; r < rows; ++r) s0 += __ldcg(p + r * step);
        out[i] = (s0 + s1) + (s2 + s3);
    }
}

__global__ void __launch_bounds__(256)
    gn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ x, long long ld_x, const __nv_bfloat16* __restrict__ dy,
                        long long ld_dy, const __nv_bfloat16* __restrict__ add, long long ld_add,
                        __nv_bfloat16* __restrict__ dx, long long ld_dx, StatSeg s0, int c_split, StatSeg s1,
                        const float* __restrict__ gamma, const float* __restrict__ beta, long long hw, int c, int groups,
                        float eps, int act, int rows_per_block, int rpp, const float* __restrict__ part, int pb,
                        float* __restrict__ ab /* [n][c][2] or null */, float* __restrict__ dgamma,
                        float* __restrict__ dbeta, unsigned int* counter) {
    pdl_enter();
    extern __shared__ float gn_sm[];
    float* sc = gn_sm;
    float* sh = sc + c;
    float* mu = sh + c;
    float* rs = mu + c;
    float* gs = rs + c;               // [groups][2]
    float* ca = gs + 2 * groups;      // [c] gamma*A then unused
    float* cb = ca + c;               // [c] gamma*B
    float* g12 = cb + c;              // [groups][2] S1, S2 scaled by 1/(hw*cpg)
    const int img = blockIdx.y, tid = threadIdx.x, nt = blockDim.x;
    const int cpg = c / groups;
    gn_channel_params(sc, sh, mu, rs, gs, s0, c_split, s1, gamma, beta, hw, c, groups, eps, img);
    for (int ch = tid; ch < c; ch += nt) {
        const float* p = part + (static_cast<long long>(img) * pb * c + ch) * 2;
        float a = 0.f, b = 0.f;
        for (int k = 0; k < pb; ++k) a += p[static_cast<long long>(k) * c * 2], b += p[static_cast<long long>(k) * c * 2 + 1];
        if (ab && blockIdx.x == 0) ab[(static_cast<long long>(img) * c + ch) * 2] = a, ab[(static_cast<long long>(img) * c + ch) * 2 + 1] = b;
        const float ga = gamma ? gamma[ch] : 1.f;
        ca[ch] = ga * a, cb[ch] = ga * b;
    }
    __syncthreads();
    const float inv_cnt = 1.f / (static_cast<float>(hw) * cpg);
    for (int g = tid; g < groups; g += nt) {
        float a = 0.f, b = 0.f;
        for (int k = 0; k < cpg; ++k) a += ca[g * cpg + k], b += cb[g * cpg + k];
        g12[2 * g] = a * inv_cnt, g12[2 * g + 1] = b * inv_cnt;
    }
    __syncthreads();
    if (counter && blockIdx.x == 0 && last_cta_arrives(counter, gridDim.y)) {   // dgamma / dbeta over the batch
        tail_reduce(ab + 1, 1, gridDim.y, c, 2, dgamma);
        tail_reduce(ab, 1, gridDim.y, c, 2, dbeta);
    }
    const int cv = c >> 3;
    const int v = tid % cv, r0 = tid / cv;
    if (r0 >= rpp) return;
    float scl[8], sft[8], m[8], r[8], t1[8], t2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int ch = v * 8 + j, g = ch / cpg;
        scl[j] = sc[ch], sft[j] = sh[ch], m[j] = mu[ch], r[j] = rs[ch], t1[j] = g12[2 * g], t2[j] = g12[2 * g + 1];
    }
    const long long row_begin = static_cast<long long>(blockIdx.x) * rows_per_block;
    long long row_end = row_begin + rows_per_block;
    if (row_end > hw) row_end = hw;
    const long long base = static_cast<long long>(img) * hw;
    for (long long rr = row_begin + r0; rr < row_end; rr += rpp) {
        const uint4 ux = *reinterpret_cast<const uint4*>(x + (base + rr) * ld_x + v * 8);
        const uint4 ud = *reinterpret_cast<const uint4*>(dy + (base + rr) * ld_dy + v * 8);
        uint4 ua = make_uint4(0u, 0u, 0u, 0u);
        if (add) ua = *reinterpret_cast<const uint4*>(add + (base + rr) * ld_add + v * 8);
        const __nv_bfloat162* hx = reinterpret_cast<const __nv_bfloat162*>(&ux);
        const __nv_bfloat162* hd = reinterpret_cast<const __nv_bfloat162*>(&ud);
        const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&ua);
        uint4 o;
        __nv_bfloat162* ho = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 fx = __bfloat1622float2(hx[j]), fd = __bfloat1622float2(hd[j]), fa = __bfloat1622float2(ha[j]);
            float res[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int jj = 2 * j + e;
                const float xv = e ? fx.y : fx.x, dv = e ? fd.y : fd.x, av = e ? fa.y : fa.x;
                const float z = xv * scl[jj] + sft[jj];
                const float dz = dv * act_grad(z, act);
                const float xh = (xv - m[jj]) * r[jj];
                const float gam = scl[jj] / r[jj];  // gamma = scale / rstd
                res[e] = r[jj] * (gam * dz - t1[jj] - xh * t2[jj]) + av;
            }
            ho[j] = __floats2bfloat162_rn(res[0], res[1]);
        }
        *reinterpret_cast<uint4*>(dx + (base + rr) * ld_dx + v * 8) = o;
    }
}

// ---- one-pass variant: a thread-block CLUSTER of K CTAs owns G whole groups of ONE image (cc = G * c/groups channels,
// chosen >= 64 where the layer has them so that a row segment is at least 128 contiguous bytes); CTA `rank` keeps rows
// [rank * rows_per_cta, ...) of that [hw, cc] slab of x and dy in shared memory between the reduction and the apply
// phase, so x and dy are read from memory ONCE (two-pass: twice, plus a partials round trip and a second launch whose
// every CTA repeats the statistics prologue for ALL channels). The per-channel sums A / B are exchanged through
// distributed shared memory: every CTA adds the K partial vectors in rank order. Each thread re-reads exactly the
// slab entries it copied itself (cp.async, issued before the statistics prologue so that the prologue's dependent loads
// hide under the slab's flight). The host picks G, K so that the slab fits (gn_fused_plan). Same arithmetic per element as the two-pass kernels; A / B
// are summed per thread over rows r0, r0 + rpp, ..., over r0 in four interleaved chains, then over ranks (fixed order).
__device__ __forceinline__ float ld_cluster_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
    return v;
}

// act'(z) with the activation fixed at compile time. SiLU: s = rcp(1 + ex2(-z log2 e)) as two bare MUFU operations
// (__expf / __fdividef wrap theirs in range-scaling sequences, ~8 more instructions per element in a kernel that is
// bound by instruction issue), act' = s * (1 + z * (1 - s)). z -> -inf: ex2 overflows to +inf, rcp gives 0, the product
// is -0 — the limit.
template <int ACT>
__device__ __forceinline__ float act_grad_fast(float z) {
    if (ACT == B200SD_ACT_SILU) {
        float e, sg;
        asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(z * -1.4426950408889634f));
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(sg) : "f"(1.f + e));
        return sg * fmaf(z, 1.f - sg, 1.f);
    }
    if (ACT == B200SD_ACT_RELU) return z > 0.f ? 1.f : 0.f;
    return 1.f;
}

__device__ __forceinline__ void gn_cp_async16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}

template <int ACT>
__global__ void __launch_bounds__(256, 3)
    gn_bwd_fused_kernel(const __nv_bfloat16* __restrict__ x, long long ld_x, const __nv_bfloat16* __restrict__ dy,
                        long long ld_dy, const __nv_bfloat16* __restrict__ add, long long ld_add,
                        __nv_bfloat16* __restrict__ dx, long long ld_dx, StatSeg s0, int c_split, StatSeg s1,
                        const float* __restrict__ gamma, const float* __restrict__ beta, int hw, int c, int groups,
                        float eps, int G, int rpp, int K, int rows_per_cta,
                        float* __restrict__ ab /* [n][c][2] or null */, float* __restrict__ dgamma,
                        float* __restrict__ dbeta, unsigned int* counter /* one per channel chunk */) {
    pdl_enter();
    constexpr int V = 8, H = 4;
    extern __shared__ __align__(16) unsigned char gnf_raw[];
    const int rank = K > 1 ? static_cast<int>(cluster_ctarank()) : 0;
    const int chunk = blockIdx.x / K;
    const int cpg = c / groups, cc = G * cpg, c0 = chunk * cc, img = blockIdx.y;
    const int tid = threadIdx.x, nt = blockDim.x;
    const int row_begin = rank * rows_per_cta;
    const int nrows = min(hw - row_begin, rows_per_cta);
    float* sc = reinterpret_cast<float*>(gnf_raw);   // scale = rstd * gamma
    float* sh = sc + cc;                             // shift = beta - mean * scale
    float* mu = sh + cc;
    float* rs = mu + cc;
    float* gm = rs + cc;                             // gamma
    float* fa = gm + cc;                             // [cc][2] this CTA's share of A[c] = sum dz, B[c] = sum dz * xhat
    float* ta = fa + 2 * cc;                         // A, B over the whole image
    float* tb = ta + cc;
    float* gs = tb + cc;                             // [G][2] mean, rstd
    float* g12 = gs + 2 * G;                         // [G][2] S1, S2 scaled by 1 / (hw * cpg)
    float* red = g12 + 2 * G;                        // [rpp][cc][2]
    const int cv = cc / V;
    const size_t slab_off = (static_cast<size_t>(9 * cc + 4 * G + 2 * rpp * cc) * sizeof(float) + 15) & ~size_t(15);
    uint4* xs = reinterpret_cast<uint4*>(gnf_raw + slab_off);
    uint4* ds = xs + static_cast<size_t>(rows_per_cta) * cv;
    uint4* as = ds + static_cast<size_t>(rows_per_cta) * cv;      // the extra gradient's slab (only when add != null)

    // ---- the slab starts flowing in before anything else: each thread copies the entries it will work on itself
    const int v = tid % cv, r0 = tid / cv;
    const bool active = r0 < rpp;
    const long long base = static_cast<long long>(img) * hw + row_begin;
    if (active) {
        const __nv_bfloat16* xp = x + (base + r0) * ld_x + c0 + v * V;
        const __nv_bfloat16* dp = dy + (base + r0) * ld_dy + c0 + v * V;
        const long long sx = rpp * ld_x, sd = rpp * ld_dy;
        const uint32_t slab_b = static_cast<uint32_t>(rows_per_cta) * cv * 16, step_b = static_cast<uint32_t>(rpp) * cv * 16;
        uint32_t sa = smem_u32(&xs[r0 * cv + v]);
        for (int rr = r0; rr < nrows; rr += rpp, xp += sx, dp += sd, sa += step_b) {
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(xp) : "memory");
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa + slab_b), "l"(dp) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        if (add) {                        // needed in phase 2 only: a second group, waited for there
            const __nv_bfloat16* ap = add + (base + r0) * ld_add + c0 + v * V;
            const long long sadd = rpp * ld_add;
            sa = smem_u32(&as[r0 * cv + v]);
            for (int rr = r0; rr < nrows; rr += rpp, ap += sadd, sa += step_b)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(ap) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }

    // ---- statistics of the cluster's channels, summed exactly like gn_channel_params / gn_apply
    for (int ch = tid; ch < cc; ch += nt) {
        const int gch = c0 + ch;
        const bool first = gch < c_split;
        const StatSeg sg = first ? s0 : s1;
        const int idx = first ? gch : gch - c_split;
        const float* p = sg.ptr + (static_cast<long long>(img) * sg.p * sg.ld + idx) * 2;
        float a = 0.f, b = 0.f;
        for (int k = 0; k < sg.p; ++k) {
            const float2 t = *reinterpret_cast<const float2*>(p + static_cast<long long>(k) * sg.ld * 2);
            a += t.x, b += t.y;
        }
        sc[ch] = a, sh[ch] = b;
        gm[ch] = gamma ? gamma[gch] : 1.f, mu[ch] = beta ? beta[gch] : 0.f;      // (mu: beta until the next step)
    }
    __syncthreads();
    const float inv_cnt = 1.f / (static_cast<float>(hw) * cpg);
    for (int g = tid; g < G; g += nt) {
        float a = 0.f, b = 0.f;
        for (int k = 0; k < cpg; ++k) a += sc[g * cpg + k], b += sh[g * cpg + k];
        const float mean = a * inv_cnt;
        const float var = fmaxf(b * inv_cnt - mean * mean, 0.f);
        gs[2 * g] = mean, gs[2 * g + 1] = rsqrtf(var + eps);
    }
    __syncthreads();
    for (int ch = tid; ch < cc; ch += nt) {
        const int g = ch / cpg;
        const float scale = gs[2 * g + 1] * gm[ch];
        sc[ch] = scale, sh[ch] = mu[ch] - gs[2 * g] * scale;
        mu[ch] = gs[2 * g], rs[ch] = gs[2 * g + 1];
    }
    __syncthreads();

    // ---- phase 1: per-thread partial A / B over the thread's own slab entries
    float scl[V], sft[V], m[V], r[V];
    const int stride_v = rpp * cv;                    // slab entries between a thread's consecutive rows
    if (active) {
        float mr[V];                                  // xhat = x * rstd - mean * rstd
#pragma unroll
        for (int j = 0; j < V; ++j) {
            scl[j] = sc[v * V + j], sft[j] = sh[v * V + j], m[j] = mu[v * V + j], r[j] = rs[v * V + j];
            mr[j] = m[j] * r[j];
        }
        float a[V], b[V];
#pragma unroll
        for (int j = 0; j < V; ++j) a[j] = b[j] = 0.f;
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        const uint4* px = xs + r0 * cv + v;
        const uint4* pd = ds + r0 * cv + v;
#pragma unroll 2
        for (int rr = r0; rr < nrows; rr += rpp, px += stride_v, pd += stride_v) {
            const uint4 ux = *px, ud = *pd;
            const __nv_bfloat162* hx = reinterpret_cast<const __nv_bfloat162*>(&ux);
            const __nv_bfloat162* hd = reinterpret_cast<const __nv_bfloat162*>(&ud);
#pragma unroll
            for (int j = 0; j < H; ++j) {
                const float2 fx = __bfloat1622float2(hx[j]), fd = __bfloat1622float2(hd[j]);
                const float z0 = fmaf(fx.x, scl[2 * j], sft[2 * j]), z1 = fmaf(fx.y, scl[2 * j + 1], sft[2 * j + 1]);
                const float d0 = fd.x * act_grad_fast<ACT>(z0), d1 = fd.y * act_grad_fast<ACT>(z1);
                a[2 * j] += d0, a[2 * j + 1] += d1;
                b[2 * j] = fmaf(d0, fmaf(fx.x, r[2 * j], -mr[2 * j]), b[2 * j]);
                b[2 * j + 1] = fmaf(d1, fmaf(fx.y, r[2 * j + 1], -mr[2 * j + 1]), b[2 * j + 1]);
            }
        }
        // red[(j, which)][r0][v]: consecutive threads write consecutive words
#pragma unroll
        for (int j = 0; j < V; ++j)
            red[((2 * j) * rpp + r0) * cv + v] = a[j], red[((2 * j + 1) * rpp + r0) * cv + v] = b[j];
    }
    __syncthreads();
    for (int i = tid; i < 2 * cc; i += nt) {
        const int ch = i >> 1;
        const float* p = red + ((2 * (ch % V) + (i & 1)) * rpp) * cv + ch / V;
        float t0 = 0.f, t1 = 0.f, t2 = 0.f, t3 = 0.f;
        int k = 0;
        for (; k + 4 <= rpp; k += 4) {
            t0 += p[k * cv], t1 += p[(k + 1) * cv];
            t2 += p[(k + 2) * cv], t3 += p[(k + 3) * cv];
        }
        for (; k < rpp; ++k) t0 += p[k * cv];
        fa[i] = (t0 + t1) + (t2 + t3);
    }
    if (K > 1) {
        cluster_sync_all();              // every CTA's partial vector is in its shared memory
        for (int i = tid; i < 2 * cc; i += nt) {
            float part[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) part[k] = k < K ? ld_cluster_f32(mapa_rank(&fa[i], k)) : 0.f;   // all in flight
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < 8; ++k) t += part[k];
            if (i & 1) tb[i >> 1] = t; else ta[i >> 1] = t;
        }
    } else {
        __syncthreads();
        for (int i = tid; i < 2 * cc; i += nt)
            if (i & 1) tb[i >> 1] = fa[i]; else ta[i >> 1] = fa[i];
    }
    __syncthreads();
    if (ab && rank == 0)
        for (int i = tid; i < 2 * cc; i += nt)
            ab[(static_cast<long long>(img) * c + c0) * 2 + i] = (i & 1) ? tb[i >> 1] : ta[i >> 1];
    for (int g = tid; g < G; g += nt) {
        float a = 0.f, b = 0.f;
        for (int k = 0; k < cpg; ++k) a += gm[g * cpg + k] * ta[g * cpg + k], b += gm[g * cpg + k] * tb[g * cpg + k];
        g12[2 * g] = a * inv_cnt, g12[2 * g + 1] = b * inv_cnt;
    }
    // arrival for the dgamma / dbeta tail HERE, while the only stores in flight are the few ab words: the fence inside
    // would otherwise wait for all of phase 2's dx stores (the tail itself runs after phase 2)
    bool own_tail = false;
    if (counter && rank == 0) own_tail = last_cta_arrives(counter + chunk, gridDim.y);
    __syncthreads();

    // ---- phase 2: dx from the slab. With c2 = rstd^2 * S2 and c1 = rstd * S1 - mean * c2 (per channel):
    //   dx = rstd * (gamma * dz - S1 - xhat * S2) = scale * dz - c1 - x * c2      (scale = rstd * gamma)
    if (active) {
        float c1[V], c2[V];
#pragma unroll
        for (int j = 0; j < V; ++j) {
            const int g = (v * V + j) / cpg;
            c2[j] = r[j] * r[j] * g12[2 * g + 1];
            c1[j] = r[j] * g12[2 * g] - m[j] * c2[j];
        }
        __nv_bfloat16* op = dx + (base + r0) * ld_dx + c0 + v * V;
        const long long so = rpp * ld_dx;
        asm volatile("cp.async.wait_all;" ::: "memory");
        const uint4* px = xs + r0 * cv + v;
        const uint4* pd = ds + r0 * cv + v;
        const uint4* pa = as + r0 * cv + v;
#pragma unroll 2
        for (int rr = r0; rr < nrows; rr += rpp, px += stride_v, pd += stride_v, pa += stride_v, op += so) {
            const uint4 ux = *px, ud = *pd;
            uint4 ua = make_uint4(0u, 0u, 0u, 0u);
            if (add) ua = *pa;
            const __nv_bfloat162* hx = reinterpret_cast<const __nv_bfloat162*>(&ux);
            const __nv_bfloat162* hd = reinterpret_cast<const __nv_bfloat162*>(&ud);
            const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&ua);
            uint4 o;
            __nv_bfloat162* ho = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
            for (int j = 0; j < H; ++j) {
                const float2 fx = __bfloat1622float2(hx[j]), fd = __bfloat1622float2(hd[j]), fa2 = __bfloat1622float2(ha[j]);
                float res[2];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int jj = 2 * j + e;
                    const float xv = e ? fx.y : fx.x, dv = e ? fd.y : fd.x, av = e ? fa2.y : fa2.x;
                    const float z = fmaf(xv, scl[jj], sft[jj]);
                    const float dz = dv * act_grad_fast<ACT>(z);
                    res[e] = fmaf(-xv, c2[jj], fmaf(scl[jj], dz, -c1[jj])) + av;
                }
                ho[j] = __floats2bfloat162_rn(res[0], res[1]);
            }
            *reinterpret_cast<uint4*>(op) = o;
        }
    }
    // ---- dgamma / dbeta of this channel chunk over the batch: whichever image's rank-0 CTA arrives last sums the chunk's
    // columns of ab in tail_reduce's order (four interleaved chains over ascending images) — one short tail per chunk
    // running beside the other clusters' streaming instead of one long tail over all channels at the end of the launch
    if (own_tail) {
        const long long step = static_cast<long long>(c) * 2;
        const int nimg = gridDim.y;
        for (int i = tid; i < 2 * cc; i += nt) {
            const float* p = ab + static_cast<long long>(c0) * 2 + i;
            float q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
            int k = 0;
            for (; k + 4 <= nimg; k += 4) {
                q0 += __ldcg(p + k * step), q1 += __ldcg(p + (k + 1) * step);
                q2 += __ldcg(p + (k + 2) * step), q3 += __ldcg(p + (k + 3) * step);
            }
            for (; k < nimg; ++k) q0 += __ldcg(p + k * step);
            ((i & 1) ? dgamma : dbeta)[c0 + (i >> 1)] = (q0 + q1) + (q2 + q3);
        }
    }
    if (K > 1) cluster_sync_all();       // my partial vector stays readable until every CTA of the cluster has read it
}

// out[b][c] (+)= sum over r of in[b][r][c]  (fp32; bias / GroupNorm-affine / time-embedding gradients)
__global__ void reduce_f32_kernel(const float* __restrict__ in, long long batch, long long rows, long long cols,
                                  long long col_stride, float* __restrict__ out, int accumulate) {
    pdl_enter();
    const long long total = batch * cols;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long b = i / cols, cidx = i - b * cols;
        const float* p = in + b * rows * cols * col_stride + cidx * col_stride;
        float s = 0.f;
        for (long long r = 0; r < rows; ++r) s += p[r * cols * col_stride];
        out[i] = accumulate ? out[i] + s : s;
    }
}

// Few outputs, many rows (the step-size / ADC-gain gradients of the cim_layers: ONE output summed over hundreds of
// per-CTA partials): one CTA per output, 256 strided partial sums and a shared-memory tree — fixed order, and ~10 us of a
// single thread's serial loop becomes ~2 us (the LSQ training leg made 229 such launches per step: 2.3 ms).
__global__ void __launch_bounds__(256)
    reduce_f32_tall_kernel(const float* __restrict__ in, long long rows, long long cols, long long col_stride,
                           float* __restrict__ out, int accumulate) {
    pdl_enter();
    __shared__ float red[256];
    const long long o = blockIdx.x;                 // output index = b * cols + cidx
    const long long b = o / cols, cidx = o - b * cols;
    const float* p = in + b * rows * cols * col_stride + cidx * col_stride;
    const long long step = cols * col_stride;
    float s = 0.f;
    for (long long r = threadIdx.x; r < rows; r += 256) s += p[r * step];
    red[threadIdx.x] = s;
    __syncthreads();
#pragma unroll
    for (int w = 128; w > 0; w >>= 1) {
        if (threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[o] = accumulate ? out[o] + red[0] : red[0];
}

// Column sums of a bf16 [n, hw, ld] tensor. CTA = (64-channel slab, row block, image): 8 lanes x 16 B cover the
// slab's 128 B of a row, 32 row lanes stride down the block. part[n][pb][c] holds the per-block sums; with `counter`
// (one zeroed uint32 PER SLAB) the last CTA of each slab folds its 64 columns over the row blocks, so the tails of a
// launch run in parallel, one per slab, each over a few dozen partial rows.
__global__ void __launch_bounds__(256)
    colsum_kernel(const __nv_bfloat16* __restrict__ x, long long ld_x, long long hw, int c, int rows_per_block,
                  float* __restrict__ part, float* __restrict__ out, int per_image, unsigned int* counter) {
    pdl_enter();
    __shared__ float red[32][65];
    const int slab = blockIdx.x, rb = blockIdx.y, img = blockIdx.z, pb = gridDim.y;
    const int tid = threadIdx.x, v = tid & 7, lane_r = tid >> 3;
    const int ch = slab * 64 + v * 8;
    float a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = 0.f;
    if (ch < c) {
        const long long row_begin = static_cast<long long>(rb) * rows_per_block;
        long long row_end = row_begin + rows_per_block;
        if (row_end > hw) row_end = hw;
        const __nv_bfloat16* xs = x + static_cast<long long>(img) * hw * ld_x + ch;
        for (long long rr = row_begin + lane_r; rr < row_end; rr += 32) {
            const uint4 u = *reinterpret_cast<const uint4*>(xs + rr * ld_x);
            const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 f = __bfloat1622float2(h[j]);
                a[2 * j] += f.x, a[2 * j + 1] += f.y;
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) red[lane_r][v * 8 + j] = a[j];
    __syncthreads();
    const int ncol = min(64, c - slab * 64);
    if (tid < ncol) {
        float s = 0.f;
#pragma unroll 8
        for (int k = 0; k < 32; ++k) s += red[k][tid];
        part[(static_cast<long long>(img) * pb + rb) * c + slab * 64 + tid] = s;
    }
    if (counter && last_cta_arrives(counter + slab, gridDim.y * gridDim.z)) {
        const float* p0 = part + slab * 64;
        if (per_image) {
            for (int i = tid; i < static_cast<int>(gridDim.z) * ncol; i += 256) {
                const int im = i / ncol, col = i - im * ncol;
                const float* p = p0 + static_cast<long long>(im) * pb * c + col;
                float s = 0.f;
                for (int k = 0; k < pb; ++k) s += __ldcg(p + static_cast<long long>(k) * c);
                out[static_cast<long long>(im) * c + slab * 64 + col] = s;
            }
        } else {
            // 4 row-block lanes per column, folded in fixed order
            const int col = tid & 63, q = tid >> 6;
            const long long rows = static_cast<long long>(pb) * gridDim.z;
            float s = 0.f;
            if (col < ncol)
                for (long long k = q; k < rows; k += 4) s += __ldcg(p0 + k * c + col);
            red[q][col] = s;
            __syncthreads();
            if (tid < ncol) out[slab * 64 + tid] = (red[0][tid] + red[1][tid]) + (red[2][tid] + red[3][tid]);
        }
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200sd_groupnorm(const void* x, int64_t ld_x, void* y, int64_t ld_y, const float* gamma,
                                const float* beta, int64_t n, int64_t hw, int c, int groups, float eps, int act,
                                void* stream) {
    B200_REQUIRE(x && y, "groupnorm: null pointer");
    B200_REQUIRE(groups > 0 && c > 0 && c % groups == 0, "groupnorm: %d channels not divisible by %d groups", c,
                 groups);
    B200_REQUIRE(ld_x >= c && ld_y >= c, "groupnorm: leading dimension smaller than channel count");
    if (n == 0 || hw == 0) return B200SD_OK;
    const int cpg = c / groups;
    const bool vec2 = (cpg % 2 == 0) && (ld_x % 2 == 0) && (ld_y % 2 == 0) &&
                      ((reinterpret_cast<uintptr_t>(x) & 3) == 0) && ((reinterpret_cast<uintptr_t>(y) & 3) == 0);
    const int VEC = vec2 ? 2 : 1;
    // groups per CTA: widen the slab towards 64 B of contiguous channels per pixel while it still fits a 64 KB
    // stage and while the grid keeps >= 2 CTAs per SM
    int gpc = 1;
    const long long min_ctas = 2ll * sm_count();
    while (true) {
        int next = gpc * 2;
        if (groups % next != 0) break;
        long long cw = (long long)next * cpg;
        if (cw > 128 || next > 64) break;
        if (cw / VEC > 256) break;
        if (hw * cw * 2 > 65536 && hw * (long long)gpc * cpg * 2 <= 65536) break;
        if ((long long)gpc * cpg * 2 >= 64 && n * (groups / next) < min_ctas) break;
        gpc = next;
    }
    const int cw = gpc * cpg;
    const int vpr = cw / VEC;
    B200_REQUIRE(vpr <= 512, "groupnorm: %d channels per group slab is too wide", cw);
    int rows_per_pass = 512 / vpr;
    if (rows_per_pass > hw) rows_per_pass = (int)hw;
    if (rows_per_pass < 1) rows_per_pass = 1;
    int threads = ((vpr * rows_per_pass + 31) / 32) * 32;
    const long long slab = hw * cw * 2;
    const int cached = slab <= 65536 ? 1 : 0;
    const size_t smem = cached ? (size_t)slab : 0;
    const long long grid = n * (c / cw);
    B200_REQUIRE(grid < (1ll << 31), "groupnorm: grid too large");
    cudaStream_t st = as_stream(stream);
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(groupnorm_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
        cudaFuncSetAttribute(groupnorm_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
        configured = true;
    }
    if (VEC == 2)
        launch_k(groupnorm_kernel<2>, (int)grid, threads, smem, st, 
            static_cast<const __nv_bfloat16*>(x), ld_x, static_cast<__nv_bfloat16*>(y), ld_y, gamma, beta, hw, c, cpg,
            gpc, eps, act, cached, rows_per_pass);
    else
        launch_k(groupnorm_kernel<1>, (int)grid, threads, smem, st, 
            static_cast<const __nv_bfloat16*>(x), ld_x, static_cast<__nv_bfloat16*>(y), ld_y, gamma, beta, hw, c, cpg,
            gpc, eps, act, cached, rows_per_pass);
    B200_CHECK_LAUNCH("groupnorm");
    return B200SD_OK;
}

extern "C" int b200sd_groupnorm_apply(const void* x, int64_t ld_x, void* y, int64_t ld_y, const float* stats0, int p0,
                                      int ld_s0, int c_split, const float* stats1, int p1, int ld_s1,
                                      const float* gamma, const float* beta, int64_t n, int64_t hw, int c, int groups,
                                      float eps, int act, void* stream) {
    B200_REQUIRE(x && y && stats0, "groupnorm_apply: null pointer");
    B200_REQUIRE(groups > 0 && c > 0 && c % groups == 0, "groupnorm_apply: %d channels not divisible by %d groups", c,
                 groups);
    B200_REQUIRE(c % 8 == 0 && c <= 2048 && ld_x % 8 == 0 && ld_y % 8 == 0 && ld_x >= c && ld_y >= c,
                 "groupnorm_apply: needs channels %% 8 == 0, <= 2048 and 16-byte aligned row pitches");
    B200_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15) == 0,
                 "groupnorm_apply: pointers unaligned");
    B200_REQUIRE(c_split >= 0 && c_split <= c && (c_split == c || stats1), "groupnorm_apply: bad statistics split");
    B200_REQUIRE(p0 >= 1 && ld_s0 >= c_split && (c_split == c || (p1 >= 1 && ld_s1 >= c - c_split)),
                 "groupnorm_apply: bad statistics geometry");
    B200_REQUIRE(n <= 65535, "groupnorm_apply: batch exceeds grid limit");
    if (n == 0 || hw == 0) return B200SD_OK;
    const int cv = c / 8;
    int rpp = 256 / cv;  // rows a CTA covers per pass
    if (rpp > hw) rpp = (int)hw;
    const int threads = ((cv * rpp + 31) / 32) * 32;
    // ~8 CTAs per SM overall = two resident waves of 4 (64 B in flight per thread; other CTAs cover the compute phase), each at least two load batches deep so the statistics
    // prologue is amortised
    long long blocks_per_img = (8ll * sm_count() + n - 1) / n;
    long long max_bpi = (hw + 2ll * GN_U * rpp - 1) / (2ll * GN_U * rpp);
    if (blocks_per_img > max_bpi) blocks_per_img = max_bpi;
    if (blocks_per_img < 1) blocks_per_img = 1;
    long long rows_per_block = (hw + blocks_per_img - 1) / blocks_per_img;
    rows_per_block = (rows_per_block + rpp - 1) / rpp * rpp;
    blocks_per_img = (hw + rows_per_block - 1) / rows_per_block;
    const size_t smem = (size_t)(2 * c + 2 * groups) * sizeof(float);
    StatSeg s0{stats0, p0, ld_s0}, s1{stats1, p1, ld_s1};
    dim3 grid((unsigned)blocks_per_img, (unsigned)n);
    auto go = [&](auto kern) {
        launch_k(kern, grid, threads, smem, as_stream(stream), static_cast<const __nv_bfloat16*>(x), ld_x,
                 static_cast<__nv_bfloat16*>(y), ld_y, s0, c_split, s1, gamma, beta, hw, c, groups, eps,
                 (int)rows_per_block, rpp);
    };
    static const bool silu_tanh = [] { const char* f = getenv("B200SD_GN_SILU_TANH"); return f && atoi(f) != 0; }();
    if (act == B200SD_ACT_SILU && silu_tanh) go(gn_apply_kernel<GN_ACT_SILU_TANH>);
    else if (act == B200SD_ACT_SILU) go(gn_apply_kernel<B200SD_ACT_SILU>);
    else if (act == B200SD_ACT_RELU) go(gn_apply_kernel<B200SD_ACT_RELU>);
    else go(gn_apply_kernel<B200SD_ACT_NONE>);
    B200_CHECK_LAUNCH("groupnorm_apply");
    return B200SD_OK;
}

// row-block geometry shared by the backward / column-sum kernels
struct RowGeom {
    int cv, rpp, threads, rows_per_block, blocks_per_img;
};
static RowGeom row_geom(int64_t n, int64_t hw, int c) {
    RowGeom g;
    g.cv = c / 8;
    g.rpp = 256 / g.cv;
    if (g.rpp > hw) g.rpp = (int)hw;
    if (g.rpp < 1) g.rpp = 1;
    g.threads = ((g.cv * g.rpp + 31) / 32) * 32;
    long long bpi = (4ll * sm_count() + n - 1) / n;
    long long max_bpi = (hw + 8ll * g.rpp - 1) / (8ll * g.rpp);
    if (bpi > max_bpi) bpi = max_bpi;
    if (bpi < 1) bpi = 1;
    long long rpb = (hw + bpi - 1) / bpi;
    rpb = (rpb + g.rpp - 1) / g.rpp * g.rpp;
    g.rows_per_block = (int)rpb;
    g.blocks_per_img = (int)((hw + rpb - 1) / rpb);
    return g;
}

extern "C" int b200sd_rowblocks(int64_t n, int64_t hw, int c) {
    if (n <= 0 || hw <= 0 || c <= 0 || c % 8) return 0;
    return row_geom(n, hw, c).blocks_per_img;
}

// One-pass GroupNorm backward (gn_bwd_fused_kernel): groups per cluster (G: the fewest that give >= 64 channels in whole
// 16-byte vectors, else narrower vectors), CTAs per cluster (K: the fewest whose slab share fits beside a second CTA on
// the SM, doubled while the launch has fewer CTAs than two per SM and a CTA would keep >= 32 rows), rows per pass,
// shared memory. false -> the two-pass kernels. B200SD_GN_BWD_FUSED=0 (read per call) forces the two-pass route.
constexpr int GN_FUSED_SMEM_MAX = 200 * 1024;
struct GnFusedPlan {
    int V, G, K, rows_per_cta, rpp, threads;
    size_t smem;
};
static bool gn_fused_plan(int64_t n, int64_t hw, int c, int groups, bool with_add, GnFusedPlan& out) {
    const char* e = getenv("B200SD_GN_BWD_FUSED");
    if (e && e[0] == '0') return false;
    if (hw <= 0 || hw > (1 << 20)) return false;
    const int cpg = c / groups, V = 8;
    const char* em = getenv("B200SD_GN_MIN_CH");
    const int min_ch = em && atoi(em) >= 8 ? atoi(em) : 64;
    int G = 0;
    for (int g = 1; g <= groups; ++g)   // (g == groups always qualifies: c is a multiple of 8 and at most 2048)
        if (groups % g == 0 && (g * cpg) % V == 0 && (g * cpg >= min_ch || g == groups) && g * cpg / V <= 256 &&
            groups / g <= 64) {
            G = g;
            break;
        }
    if (!G) return false;
    const int cc = G * cpg, cv = cc / V;
    auto smem_of = [&](int K, int& rows, int& rpp) {
        rows = (int)((hw + K - 1) / K);
        rpp = 256 / cv;
        if (rpp > rows) rpp = rows;
        const size_t head = ((size_t)(9 * cc + 4 * G + 2 * rpp * cc) * sizeof(float) + 15) & ~size_t(15);
        return head + (with_add ? 3 : 2) * (size_t)rows * cc * 2;
    };
    const size_t limits[3] = {72 * 1024, 108 * 1024, (size_t)GN_FUSED_SMEM_MAX};   // three, two, one CTA per SM
    for (int pass = 0; pass < 3; ++pass) {
        for (int K = 1; K <= 8; K *= 2) {
            int rows, rpp;
            if ((int64_t)(K - 1) * ((hw + K - 1) / K) >= hw) break;       // a rank without rows
            if (smem_of(K, rows, rpp) > limits[pass]) continue;
            while (K < 8 && n * (groups / G) * K < 2ll * sm_count() && (hw + 2 * K - 1) / (2 * K) >= 32 &&
                   (int64_t)(2 * K - 1) * ((hw + 2 * K - 1) / (2 * K)) < hw)
                K *= 2;
            out.smem = smem_of(K, rows, rpp);
            out.V = V, out.G = G, out.K = K, out.rows_per_cta = rows, out.rpp = rpp;
            out.threads = ((cv * rpp + 31) / 32) * 32;
            return true;
        }
    }
    return false;
}

extern "C" int b200sd_groupnorm_bwd_plan(int64_t n, int64_t hw, int c, int groups, int with_add, int32_t* out6) {
    GnFusedPlan fp;
    if (n <= 0 || c <= 0 || groups <= 0 || c % groups || c % 8 || c > 2048 || !gn_fused_plan(n, hw, c, groups, with_add != 0, fp))
        return 0;
    if (out6) {
        out6[0] = fp.G, out6[1] = fp.K, out6[2] = fp.rows_per_cta, out6[3] = fp.rpp, out6[4] = fp.threads;
        out6[5] = static_cast<int32_t>(fp.smem);
    }
    return 1;
}

extern "C" int b200sd_groupnorm_bwd(const void* x, int64_t ld_x, const void* dy, int64_t ld_dy, const void* add,
                                    int64_t ld_add, void* dx, int64_t ld_dx, const float* stats0, int p0, int ld_s0,
                                    int c_split, const float* stats1, int p1, int ld_s1, const float* gamma,
                                    const float* beta, int64_t n, int64_t hw, int c, int groups, float eps, int act,
                                    float* part, float* ab, float* dgamma, float* dbeta, uint32_t* counter,
                                    void* stream) {
    B200_REQUIRE(x && dy && dx && stats0 && part, "groupnorm_bwd: null pointer");
    B200_REQUIRE(!counter || (ab && dgamma && dbeta), "groupnorm_bwd: fused dgamma/dbeta needs ab, dgamma and dbeta");
    B200_REQUIRE(groups > 0 && c > 0 && c % groups == 0 && c % 8 == 0 && c <= 2048,
                 "groupnorm_bwd: channels must be divisible by groups and by 8, <= 2048");
    B200_REQUIRE(ld_x % 8 == 0 && ld_dy % 8 == 0 && ld_dx % 8 == 0 && (!add || ld_add % 8 == 0),
                 "groupnorm_bwd: row pitches must be multiples of 8 elements");
    B200_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(dx) |
                   reinterpret_cast<uintptr_t>(add)) & 15) == 0, "groupnorm_bwd: pointers unaligned");
    B200_REQUIRE(c_split >= 0 && c_split <= c && (c_split == c || stats1), "groupnorm_bwd: bad statistics split");
    B200_REQUIRE(n <= 65535, "groupnorm_bwd: batch exceeds grid limit");
    if (n == 0 || hw == 0) return B200SD_OK;
    StatSeg s0{stats0, p0, ld_s0}, s1{stats1, p1, ld_s1};
    cudaStream_t st = as_stream(stream);
    GnFusedPlan fp;
    if (gn_fused_plan(n, hw, c, groups, add != nullptr, fp)) {
        static bool fused_configured = false;
        if (!fused_configured) {
            // the plan sizes slabs for two or three CTAs per SM: that needs the full shared-memory carve-out
#define GN_FUSED_CFG(A)                                                                                                  \
    cudaFuncSetAttribute(gn_bwd_fused_kernel<A>, cudaFuncAttributeMaxDynamicSharedMemorySize, GN_FUSED_SMEM_MAX);         \
    cudaFuncSetAttribute(gn_bwd_fused_kernel<A>, cudaFuncAttributePreferredSharedMemoryCarveout,                          \
                         cudaSharedmemCarveoutMaxShared)
            GN_FUSED_CFG(B200SD_ACT_NONE);
            GN_FUSED_CFG(B200SD_ACT_SILU);
            GN_FUSED_CFG(B200SD_ACT_RELU);
#undef GN_FUSED_CFG
            fused_configured = true;
        }
        dim3 fgrid((unsigned)(groups / fp.G * fp.K), (unsigned)n);
#define GN_FUSED_LAUNCH(A)                                                                                               \
    launch_k_cluster(gn_bwd_fused_kernel<A>, fgrid, fp.threads, fp.smem, st, (unsigned)fp.K,                              \
                     static_cast<const __nv_bfloat16*>(x), ld_x, static_cast<const __nv_bfloat16*>(dy), ld_dy,            \
                     static_cast<const __nv_bfloat16*>(add), ld_add, static_cast<__nv_bfloat16*>(dx), ld_dx, s0, c_split, \
                     s1, gamma, beta, (int)hw, c, groups, eps, fp.G, fp.rpp, fp.K, fp.rows_per_cta, ab, dgamma, dbeta,    \
                     counter)
        if (act == B200SD_ACT_SILU) GN_FUSED_LAUNCH(B200SD_ACT_SILU);
        else if (act == B200SD_ACT_RELU) GN_FUSED_LAUNCH(B200SD_ACT_RELU);
        else GN_FUSED_LAUNCH(B200SD_ACT_NONE);
#undef GN_FUSED_LAUNCH
        B200_CHECK_LAUNCH("groupnorm_bwd_fused");
        return B200SD_OK;
    }
    const RowGeom g = row_geom(n, hw, c);
    dim3 grid((unsigned)g.blocks_per_img, (unsigned)n);
    const size_t sm1 = (size_t)(4 * c + 2 * groups + 2 * g.rpp * c) * sizeof(float);
    const size_t sm2 = (size_t)(6 * c + 4 * groups) * sizeof(float);
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(gn_bwd_reduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
        cudaFuncSetAttribute(gn_bwd_apply_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
        configured = true;
    }
    B200_REQUIRE(sm1 <= 160 * 1024, "groupnorm_bwd: reduction scratch too large");
    launch_k(gn_bwd_reduce_kernel, grid, g.threads, sm1, st, 
        static_cast<const __nv_bfloat16*>(x), ld_x, static_cast<const __nv_bfloat16*>(dy), ld_dy, s0, c_split, s1, gamma,
        beta, hw, c, groups, eps, act, g.rows_per_block, g.rpp, part);
    B200_CHECK_LAUNCH("groupnorm_bwd_reduce");
    launch_k(gn_bwd_apply_kernel, grid, g.threads, sm2, st, 
        static_cast<const __nv_bfloat16*>(x), ld_x, static_cast<const __nv_bfloat16*>(dy), ld_dy,
        static_cast<const __nv_bfloat16*>(add), ld_add, static_cast<__nv_bfloat16*>(dx), ld_dx, s0, c_split, s1, gamma,
        beta, hw, c, groups, eps, act, g.rows_per_block, g.rpp, part, g.blocks_per_img, ab, dgamma, dbeta, counter);
    B200_CHECK_LAUNCH("groupnorm_bwd_apply");
    return B200SD_OK;
}

static int colsum_blocks(int64_t n, int64_t hw, int c) {
    const long long slabs = (c + 63) / 64;
    long long pb = (3ll * sm_count() + n * slabs - 1) / (n * slabs);
    const long long max_pb = (hw + 63) / 64;          // at least two passes of the 32 row lanes per CTA
    if (pb > max_pb) pb = max_pb;
    if (pb > 65535) pb = 65535;
    if (pb < 1) pb = 1;
    return static_cast<int>(pb);
}

extern "C" int b200sd_colsum_blocks(int64_t n, int64_t hw, int c) {
    if (n <= 0 || hw <= 0 || c <= 0 || c % 8) return 0;
    return colsum_blocks(n, hw, c);
}

extern "C" int b200sd_colsum(const void* x, int64_t ld_x, int64_t n, int64_t hw, int c, float* part, float* out,
                             int per_image, uint32_t* counter, void* stream) {
    B200_REQUIRE(x && part, "colsum: null pointer");
    B200_REQUIRE(!counter || out, "colsum: fused total needs an output");
    B200_REQUIRE(c > 0 && c % 8 == 0 && c <= 4096 && ld_x % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0,
                 "colsum: needs channels %% 8 == 0 (<= 4096) and 16-byte aligned rows");
    B200_REQUIRE(n <= 65535, "colsum: batch exceeds grid limit");
    if (n == 0 || hw == 0) return B200SD_OK;
    const int pb = colsum_blocks(n, hw, c);
    const long long rpb = (hw + pb - 1) / pb;
    dim3 grid((unsigned)((c + 63) / 64), (unsigned)pb, (unsigned)n);
    launch_k(colsum_kernel, grid, 256, 0, as_stream(stream), static_cast<const __nv_bfloat16*>(x), ld_x, hw, c, (int)rpb, part,
                                                       out, per_image, counter);
    B200_CHECK_LAUNCH("colsum");
    return B200SD_OK;
}

extern "C" int b200sd_reduce_f32(const float* in, int64_t batch, int64_t rows, int64_t cols, int64_t col_stride,
                                 float* out, int accumulate, void* stream) {
    B200_REQUIRE(in && out && batch >= 0 && rows >= 0 && cols >= 0 && col_stride >= 1, "reduce_f32: bad arguments");
    if (batch * cols == 0) return B200SD_OK;
    if (batch * cols <= 64 && rows >= 64) {
        launch_k(reduce_f32_tall_kernel, (int)(batch * cols), 256, 0, as_stream(stream), in, rows, cols, col_stride, out,
                 accumulate);
        B200_CHECK_LAUNCH("reduce_f32");
        return B200SD_OK;
    }
    long long blocks = (batch * cols + 127) / 128;
    if (blocks > 4ll * sm_count()) blocks = 4ll * sm_count();
    launch_k(reduce_f32_kernel, (int)blocks, 128, 0, as_stream(stream), in, batch, rows, cols, col_stride, out, accumulate);
    B200_CHECK_LAUNCH("reduce_f32");
    return B200SD_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// b200sd_colsum_static — the same column sums for callers that may not allocate (CUDA-graph capture, serving loops,
// harnesses that forbid cudaMalloc). No counterpart in the reference.
// The row-block partials and the per-slab tail counters live in a file-scope __device__ array (held by every GPU that
// loads the library until exit). Host code may not use the array's NAME as an address (it is a host-side stand-in):
// the device address is fetched once per device with cudaGetSymbolAddress and passed to the kernel.
// All calls share the array: two calls that overlap on the GPU (different streams, or concurrent graph branches) race.
// ---------------------------------------------------------------------------------------------------------------
#include <mutex>

namespace b200 {
constexpr long long COLSUM_STATIC_MAX_FLOATS = 512ll * 1024;   // 2 MiB of partials: n * row blocks * c <= this
constexpr int COLSUM_STATIC_COUNTERS = 64;                     // one per 64-channel slab (c <= 4096)
struct ColsumStaticScratch {
    float part[COLSUM_STATIC_MAX_FLOATS];
    unsigned int counter[COLSUM_STATIC_COUNTERS];
};
static __device__ __align__(256) ColsumStaticScratch g_colsum_static;
constexpr int MAX_DEVICES = 64;
static ColsumStaticScratch* g_colsum_static_addr[MAX_DEVICES] = {};
static std::mutex g_colsum_static_mu;
}  // namespace b200

extern "C" int64_t b200sd_colsum_static_max_floats(void) { return COLSUM_STATIC_MAX_FLOATS; }

extern "C" int b200sd_colsum_static(const void* x, int64_t ld_x, int64_t n, int64_t hw, int c, float* out,
                                    int per_image, void* stream) {
    B200_REQUIRE(x && out, "colsum_static: null pointer");
    B200_REQUIRE(c > 0 && c % 8 == 0 && c <= 4096 && ld_x % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0,
                 "colsum_static: needs channels %% 8 == 0 (<= 4096) and 16-byte aligned rows");
    B200_REQUIRE(n >= 0 && n <= 65535 && hw >= 0, "colsum_static: batch exceeds grid limit");
    const int pb = (n == 0 || hw == 0) ? 0 : colsum_blocks(n, hw, c);
    const long long need = static_cast<long long>(n) * pb * c;
    B200_REQUIRE(need <= COLSUM_STATIC_MAX_FLOATS,
                 "colsum_static: %lld floats of partials (n %lld x %d row blocks x c %d) exceed the static scratch of "
                 "%lld floats; use b200sd_colsum with caller-owned scratch",
                 need, static_cast<long long>(n), pb, c, COLSUM_STATIC_MAX_FLOATS);
    if (n == 0 || hw == 0) return B200SD_OK;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) {
        b200::set_error("colsum_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        return B200SD_ERR_CUDA;
    }
    ColsumStaticScratch* scratch;
    {
        std::lock_guard<std::mutex> lk(g_colsum_static_mu);
        if (!g_colsum_static_addr[dev]) {   // set-up, once per device, with that device current
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_colsum_static);
            if (e != cudaSuccess || !p) {
                b200::set_error("colsum_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
                return B200SD_ERR_CUDA;
            }
            g_colsum_static_addr[dev] = static_cast<ColsumStaticScratch*>(p);
        }
        scratch = g_colsum_static_addr[dev];
    }
    cudaStream_t st = as_stream(stream);
    const unsigned slabs = (unsigned)((c + 63) / 64);
    // the counters must start at zero; a launch leaves them there, but an aborted one might not: clear them (no allocation)
    e = cudaMemsetAsync(scratch->counter, 0, sizeof(unsigned int) * slabs, st);
    if (e != cudaSuccess) {
        b200::set_error("colsum_static: cudaMemsetAsync: %s", cudaGetErrorString(e));
        return B200SD_ERR_CUDA;
    }
    const long long rpb = (hw + pb - 1) / pb;
    dim3 grid(slabs, (unsigned)pb, (unsigned)n);
    launch_k(colsum_kernel, grid, 256, 0, st, static_cast<const __nv_bfloat16*>(x), ld_x, hw, c, (int)rpb, scratch->part, out,
             per_image, scratch->counter);
    B200_CHECK_LAUNCH("colsum_static");
    return B200SD_OK;
}

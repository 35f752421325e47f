// cim_layers quantisers on activations (reference cim_layers/quant_noise_utils.py:101-120 data_quant_lsq,
// cim_layers/layers_utils_lsq.py:31-82 lazy step-size initialisation, cim_layers/layers_utils_adda.py:161-185
// bit_split_tensor / :144-159 bit_concat_tensor, cim_layers/layers_qn_lsq_adda_cim.py:464-480 the ADC).
//
// Step sizes and the ADC scale are nn.Parameters: every kernel reads them through DEVICE pointers, so a forward makes
// no host synchronisation (the reference does three `step_size == 1` .item()-style checks per layer per forward).
// Arithmetic that decides an integer code uses the same IEEE operations in the same order as the reference's fp32
// torch ops (true division, clamp, round-half-even), so codes are bit-exact for identical inputs. All kernels are
// HBM-bound streams with 16-byte accesses; reductions are fixed-order partials (no atomics on floats).
#include <cstdlib>
#include "host.h"
#include <cuda_bf16.h>

namespace b200 {

__device__ __forceinline__ void q_ld8(const __nv_bfloat16* p, float* f) {
    const uint4 u = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float2 v = __bfloat1622float2(h[j]);
        f[2 * j] = v.x, f[2 * j + 1] = v.y;
    }
}
__device__ __forceinline__ void q_ld8(const float* p, float* f) {
    const float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
    f[0] = a.x, f[1] = a.y, f[2] = a.z, f[3] = a.w, f[4] = b.x, f[5] = b.y, f[6] = b.z, f[7] = b.w;
}
__device__ __forceinline__ void q_st8(__nv_bfloat16* p, const float* f) {
    uint4 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
    *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ void q_st8(float* p, const float* f) {
    reinterpret_cast<float4*>(p)[0] = make_float4(f[0], f[1], f[2], f[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(f[4], f[5], f[6], f[7]);
}

// The reference quantises with s' = grad_scale(s, g) = (s - s*g).detach() + s*g (quant_noise_utils.py:51-54), whose
// VALUE is fl(fl(s - fl(s*g)) + fl(s*g)) — not always s. Codes are decided with exactly that number.
__device__ __forceinline__ float effective_step(float s, float g) {
    const float y = __fmul_rn(s, g);
    return __fadd_rn(__fsub_rn(s, y), y);
}

// Tail of a reduction kernel: the block's partial is published, and when `total` is given the LAST CTA to arrive folds
// all partials in fixed order (all 256 threads: strided partial sums, shuffle tree, warps in order) and writes
// *total = sum * scale — no second launch for a [blocks] -> [1] reduction and no torch multiply behind it.
// `counter` is a zeroed uint32 that the tail returns to zero.
__device__ __forceinline__ void q_publish_and_fold(float block_sum, float* __restrict__ part, float* __restrict__ total,
                                                   unsigned int* counter, float scale, float* wsum) {
    if (threadIdx.x == 0) part[blockIdx.x] = block_sum;
    if (!total) return;
    __shared__ unsigned int s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int t = atomicAdd(counter, 1u);
        s_last = (t == gridDim.x - 1u) ? 1u : 0u;
        if (s_last) *counter = 0u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    float v = 0.f;
    for (unsigned int k = threadIdx.x; k < gridDim.x; k += blockDim.x) v += __ldcg(part + k);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int w = 0; w < 8; ++w) t += wsum[w];
        *total = t * scale;
    }
}

// ---- max |x| over a [rows, c] tensor (pitch ld) into *slot as float bits (non-negative floats order like uints)
template <typename T>
__global__ void absmax_kernel(const T* __restrict__ x, long long ld, long long rows, int c, unsigned int* slot) {
    pdl_enter();
    const int nchunk = c >> 3;
    const long long total = rows * nchunk;
    float m = 0.f;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / nchunk;
        const int ch = static_cast<int>(i - r * nchunk);
        float f[8];
        q_ld8(x + r * ld + ch * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) m = fmaxf(m, fabsf(f[j]));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(slot, __float_as_uint(m));  // integer max: order independent, exact
}

// init_step_size (layers_utils_lsq.py:31-34 via data_quant(isint=True), quant_noise_utils.py:8-28): while the
// parameter still holds its construction value 1.0, step = 1 / (1 / max|x| * Qp); an all-zero tensor leaves it at 1.
__global__ void lsq_init_step_kernel(float* step, const unsigned int* slot, float qp) {
    pdl_enter();
    if (*step != 1.0f) return;
    const float range = __uint_as_float(*slot);
    if (range == 0.f) return;
    const float scale = __fmul_rn(__fdiv_rn(1.0f, range), qp);
    *step = __fdiv_rn(1.0f, scale);
}

// mode 0: y = round(clamp(x / s', -Qp, Qp)) * s'          (fake quantisation, isint=False)
// mode 1: y = round(clamp(x / s', -Qp, Qp)) * s' / s'     (isint=True: the code, up to the reference's own rounding
//                                                          of q * s' / s'.detach(), quant_noise_utils.py:117-118)
template <typename TI, typename TO>
__global__ void lsq_quant_kernel(const TI* __restrict__ x, long long ld_x, TO* __restrict__ y, long long ld_y,
                                 long long rows, int c, const float* __restrict__ step, float qp, float g, int mode) {
    pdl_enter();
    const float s = effective_step(*step, g);
    const int nchunk = c >> 3;
    const long long total = rows * nchunk;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / nchunk;
        const int ch = static_cast<int>(i - r * nchunk);
        float f[8];
        q_ld8(x + r * ld_x + ch * 8, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float q = rintf(fminf(fmaxf(__fdiv_rn(f[j], s), -qp), qp));
            const float v = __fmul_rn(q, s);
            f[j] = mode == 0 ? v : __fdiv_rn(v, s);
        }
        q_st8(y + r * ld_y + ch * 8, f);
    }
}

// Backward of data_quant_lsq: with u = x / s, mask = (-Qp <= u <= Qp), q = round(clamp(u)):
//   mode 0: dx = dy * mask;      d s' = sum dy * (q - u * mask)
//   mode 1: dx = dy * mask / s;  d s' = sum dy * (q - u * mask) / s      (out = q * s' / s'.detach())
// part[block] receives the block's share of d s' (the caller folds the partials and applies the LSQ grad scale
// 1 / sqrt(Qp * numel), quant_noise_utils.py:57-58).
template <typename TI>
__global__ void __launch_bounds__(256)
    lsq_quant_bwd_kernel(const __nv_bfloat16* __restrict__ dy, long long ld_dy, const TI* __restrict__ x, long long ld_x,
                         __nv_bfloat16* __restrict__ dx, long long ld_dx, long long rows, int c,
                         const float* __restrict__ step, float qp, float g, int mode, float* __restrict__ part,
                         float* __restrict__ sum_out, unsigned int* counter) {
    pdl_enter();
    __shared__ float wsum[8];
    const float s = effective_step(*step, g);
    const float inv_s = __fdiv_rn(1.0f, s);
    const int nchunk = c >> 3;
    const long long total = rows * nchunk;
    float acc = 0.f;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / nchunk;
        const int ch = static_cast<int>(i - r * nchunk);
        float f[8], g[8];
        q_ld8(x + r * ld_x + ch * 8, f);
        q_ld8(dy + r * ld_dy + ch * 8, g);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float u = __fdiv_rn(f[j], s);
            const bool in = (u >= -qp) && (u <= qp);
            const float q = rintf(fminf(fmaxf(u, -qp), qp));
            acc += g[j] * (q - (in ? u : 0.f));
            g[j] = in ? (mode == 0 ? g[j] : g[j] * inv_s) : 0.f;
        }
        q_st8(dx + r * ld_dx + ch * 8, g);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = acc;
    __syncthreads();
    float t = 0.f;
    if (threadIdx.x == 0) {
        for (int w = 0; w < 8; ++w) t += wsum[w];
        t = mode == 0 ? t : t * inv_s;
    }
    q_publish_and_fold(t, part, sum_out, counter, g, wsum);   // total = d(step) with the LSQ gradient scale applied
}

// bit_split_tensor (layers_utils_adda.py:161-185): a (nearly) integer code v -> `slices` signed digits of slice_bit
// bits, stacked along rows ([slices * rows, c], slice k in rows [k * rows, (k + 1) * rows)):
//   digit_k = sign(v) * floor(|v| / 2^lsb) - trunc(v / 2^msb) * 2^slice_bit,  lsb = k * slice_bit,
//   msb = min(lsb + slice_bit, x_bit - 1)
// evaluated in fp32 like the reference (its codes are q * s' / s', which can sit one ulp below the integer q — the
// floor then yields q - 1; reproduced, not corrected).
// Divisions by 2^lsb / 2^msb are written as multiplications by the exact reciprocal power of two: the same fp32
// result (codes are < 2^9, nowhere near the subnormal range) without four IEEE division sequences per element and
// slice, which made this kernel instruction-bound (14.6 us per launch in the ADDA training leg).
__device__ __forceinline__ float split_digit(float v, int lsb, int msb, int slice_bit) {
    const float a = __fmul_rn(v, __int_as_float((127 - lsb) << 23));
    const float shift = floorf(fabsf(a)) * (a > 0.f ? 1.f : (a < 0.f ? -1.f : 0.f));
    const float t = __fmul_rn(v, __int_as_float((127 - msb) << 23));
    const float resid = (t >= 0.f ? floorf(t) : ceilf(t)) * static_cast<float>(1 << slice_bit);
    return shift - resid;
}

// quantise = 0: x holds the codes; quantise = 1: x holds activations, codes are computed here (LSQ isint=True) and
// never written out.
template <typename TI>
__global__ void bit_split_kernel(const TI* __restrict__ q, long long ld_q, __nv_bfloat16* __restrict__ out,
                                 long long ld_o, long long rows, int c, int slices, int slice_bit, int x_bit,
                                 int quantise, const float* __restrict__ step, float qp, float g) {
    pdl_enter();
    const float s = quantise ? effective_step(*step, g) : 1.f;
    const int nchunk = c >> 3;
    const long long total = rows * nchunk;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / nchunk;
        const int ch = static_cast<int>(i - r * nchunk);
        float f[8];
        q_ld8(q + r * ld_q + ch * 8, f);
        if (quantise) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float code = rintf(fminf(fmaxf(__fdiv_rn(f[j], s), -qp), qp));
                f[j] = __fdiv_rn(__fmul_rn(code, s), s);
            }
        }
        for (int k = 0; k < slices; ++k) {
            const int lsb = k * slice_bit;
            const int msb = min(lsb + slice_bit, x_bit - 1);
            float d[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) d[j] = split_digit(f[j], lsb, msb, slice_bit);
            q_st8(out + (static_cast<long long>(k) * rows + r) * ld_o + ch * 8, d);
        }
    }
}

// The ADC of one mapped weight block (layers_qn_lsq_adda_cim.py:464-480 + bit_concat_tensor): for every slice k the
// analogue partial sum P_k = digits_k @ W_block arrives as fp32 [slices * rows, ncol];
//   code_k = round(clamp(alpha * P_k * (1 + gain_noise[col]) + R * offset_noise[col], -R - 1, R))
//   out[r, col0 + col] += sum_k 2^(k * slice_bit) * code_k
__global__ void adc_combine_kernel(const float* __restrict__ p, long long ld_p, float* __restrict__ out, long long ld_o,
                                   long long rows, int ncol, int col0, int slices, int slice_bit,
                                   const float* __restrict__ alpha_p, float R, const float* __restrict__ gain_noise,
                                   const float* __restrict__ offset_noise, int accumulate) {
    pdl_enter();
    const float alpha = *alpha_p;
    const long long total = rows * ncol;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / ncol;
        const int col = static_cast<int>(i - r * ncol);
        float sum = 0.f;
        for (int k = 0; k < slices; ++k) {
            float v = __fmul_rn(alpha, p[(static_cast<long long>(k) * rows + r) * ld_p + col]);
            if (gain_noise) {
                v = __fmul_rn(v, __fadd_rn(1.0f, gain_noise[col0 + col]));
                v = __fadd_rn(v, __fmul_rn(R, offset_noise[col0 + col]));
            }
            v = rintf(fminf(fmaxf(v, -R - 1.0f), R));
            sum += v * static_cast<float>(1 << (k * slice_bit));   // integers: exact in fp32
        }
        float* o = out + r * ld_o + col0 + col;
        *o = accumulate ? *o + sum : sum;
    }
}

// Backward of the ADC of one mapped block. Forward: a_k = alpha * P_k (+ noise), code_k = round(clamp(a_k, -R-1, R)),
// acc += sum_k 2^(k*slice_bit) * code_k. With g = d acc (per output element) the straight-through gradients are
//   d P_k = alpha * 2^(k*slice_bit) * g * [a_k in range]          (bf16 [slices * rows, ld_d], GEMM operand)
//   d alpha += sum 2^(k*slice_bit) * g * P_k * [a_k in range]     (fixed-order partial per CTA)
__global__ void __launch_bounds__(256)
    adc_bwd_kernel(const float* __restrict__ p, long long ld_p, const float* __restrict__ dacc, long long ld_g,
                   __nv_bfloat16* __restrict__ dp, long long ld_d, long long rows, int ncol, int col0, int slices,
                   int slice_bit, const float* __restrict__ alpha_p, float R, const float* __restrict__ gain_noise,
                   const float* __restrict__ offset_noise, float* __restrict__ part, float* __restrict__ sum_out,
                   unsigned int* counter) {
    pdl_enter();
    __shared__ float wsum[8];
    const float alpha = *alpha_p;
    const long long total = rows * ncol;
    float acc = 0.f;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / ncol;
        const int col = static_cast<int>(i - r * ncol);
        const float g = dacc[r * ld_g + col0 + col];
        for (int k = 0; k < slices; ++k) {
            const float pv = p[(static_cast<long long>(k) * rows + r) * ld_p + col];
            float v = __fmul_rn(alpha, pv);
            if (gain_noise) {
                v = __fmul_rn(v, __fadd_rn(1.0f, gain_noise[col0 + col]));
                v = __fadd_rn(v, __fmul_rn(R, offset_noise[col0 + col]));
            }
            const bool in = (v >= -R - 1.0f) && (v <= R);
            const float gk = in ? g * static_cast<float>(1 << (k * slice_bit)) : 0.f;
            acc += gk * pv;
            dp[(static_cast<long long>(k) * rows + r) * ld_d + col] = __float2bfloat16_rn(alpha * gk);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = acc;
    __syncthreads();
    float t = 0.f;
    if (threadIdx.x == 0)
        for (int w = 0; w < 8; ++w) t += wsum[w];
    q_publish_and_fold(t, part, sum_out, counter, 1.f, wsum);
}

// Backward of bit_split_tensor: d code = sum_k d digit_k / (slices * 2^(k*slice_bit)) (layers_utils_adda.py:178-183:
// every slice passes the gradient of shift / slices, shift = floor_pass(code / 2^lsb)).
__global__ void bit_split_bwd_kernel(const __nv_bfloat16* __restrict__ dd, long long ld_d, __nv_bfloat16* __restrict__ dq,
                                     long long ld_q, long long rows, int c, int slices, int slice_bit) {
    pdl_enter();
    const int nchunk = c >> 3;
    const long long total = rows * nchunk;
    const float inv = 1.f / static_cast<float>(slices);
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / nchunk;
        const int ch = static_cast<int>(i - r * nchunk);
        float s[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) s[j] = 0.f;
        for (int k = 0; k < slices; ++k) {
            float f[8];
            q_ld8(dd + (static_cast<long long>(k) * rows + r) * ld_d + ch * 8, f);
            const float w = inv / static_cast<float>(1 << (k * slice_bit));
#pragma unroll
            for (int j = 0; j < 8; ++j) s[j] += f[j] * w;
        }
        q_st8(dq + r * ld_q + ch * 8, s);
    }
}

static int q_grid(long long items, int per_block, int cap_mult = 16) {
    long long b = (items + per_block - 1) / per_block;
    const long long cap = static_cast<long long>(cap_mult) * sm_count();
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return static_cast<int>(b);
}

// CTAs per SM of the backward kernels that end in a last-CTA fold (one atomic per CTA on one counter, then a pass over
// the partials): B200SD_QBWD_CAP overrides (tools/quant_bwd_sweep.py).
static int qbwd_cap() {
    static const int cap = [] {
        const char* e = getenv("B200SD_QBWD_CAP");
        const int v = e ? atoi(e) : 0;
        return v > 0 && v <= 16 ? v : 4;
    }();
    return cap;
}

}  // namespace b200

using namespace b200;

#define Q_SHAPE_OK(c, ld) ((c) > 0 && (c) % 8 == 0 && (ld) % 8 == 0 && (ld) >= (c))

extern "C" int b200sd_lsq_init_step(const void* x, int32_t x_f32, int64_t ld_x, int64_t rows, int32_t c, float* step,
                                    float qp, uint32_t* scratch, void* stream) {
    B200_REQUIRE(x && step && scratch && rows >= 0 && qp >= 1.f, "lsq_init_step: bad arguments");
    B200_REQUIRE(Q_SHAPE_OK(c, ld_x), "lsq_init_step: channels and pitch must be multiples of 8");
    cudaStream_t st = as_stream(stream);
    cudaMemsetAsync(scratch, 0, sizeof(uint32_t), st);
    if (rows > 0) {
        const int grid = q_grid(rows * (c / 8), 256, 4);
        if (x_f32)
            launch_k(absmax_kernel<float>, grid, 256, 0, st, static_cast<const float*>(x), ld_x, rows, c, scratch);
        else
            launch_k(absmax_kernel<__nv_bfloat16>, grid, 256, 0, st, static_cast<const __nv_bfloat16*>(x), ld_x, rows, c, scratch);
        B200_CHECK_LAUNCH("lsq_init_step(absmax)");
    }
    launch_k(lsq_init_step_kernel, 1, 1, 0, st, step, scratch, qp);
    B200_CHECK_LAUNCH("lsq_init_step");
    return B200SD_OK;
}

extern "C" int b200sd_lsq_quant(const void* x, int32_t x_f32, int64_t ld_x, void* y, int32_t y_f32, int64_t ld_y,
                                int64_t rows, int32_t c, const float* step, float qp, float g, int32_t mode,
                                void* stream) {
    B200_REQUIRE(x && y && step && rows >= 0 && qp >= 1.f && (mode == 0 || mode == 1), "lsq_quant: bad arguments");
    B200_REQUIRE(Q_SHAPE_OK(c, ld_x) && Q_SHAPE_OK(c, ld_y), "lsq_quant: channels and pitches must be multiples of 8");
    if (rows == 0) return B200SD_OK;
    const int grid = q_grid(rows * (c / 8), 256);
    cudaStream_t st = as_stream(stream);
    if (x_f32 && y_f32)
        launch_k(lsq_quant_kernel<float, float>, grid, 256, 0, st, static_cast<const float*>(x), ld_x, static_cast<float*>(y), ld_y, rows, c, step, qp, g, mode);
    else if (x_f32)
        launch_k(lsq_quant_kernel<float, __nv_bfloat16>, grid, 256, 0, st, static_cast<const float*>(x), ld_x, static_cast<__nv_bfloat16*>(y), ld_y, rows, c, step, qp, g, mode);
    else if (y_f32)
        launch_k(lsq_quant_kernel<__nv_bfloat16, float>, grid, 256, 0, st, static_cast<const __nv_bfloat16*>(x), ld_x, static_cast<float*>(y), ld_y, rows, c, step, qp, g, mode);
    else
        launch_k(lsq_quant_kernel<__nv_bfloat16, __nv_bfloat16>, grid, 256, 0, st, static_cast<const __nv_bfloat16*>(x), ld_x, static_cast<__nv_bfloat16*>(y), ld_y, rows, c, step, qp, g, mode);
    B200_CHECK_LAUNCH("lsq_quant");
    return B200SD_OK;
}

extern "C" int b200sd_lsq_quant_bwd_blocks(int64_t rows, int32_t c) { return q_grid(rows * (c / 8), 256, qbwd_cap()); }

extern "C" int b200sd_lsq_quant_bwd(const void* dy, int64_t ld_dy, const void* x, int32_t x_f32, int64_t ld_x, void* dx,
                                    int64_t ld_dx, int64_t rows, int32_t c, const float* step, float qp, float g,
                                    int32_t mode, float* part, float* total, uint32_t* counter, void* stream) {
    B200_REQUIRE(dy && x && dx && step && part && rows > 0 && qp >= 1.f && (mode == 0 || mode == 1),
                 "lsq_quant_bwd: bad arguments");
    B200_REQUIRE((total == nullptr) == (counter == nullptr), "lsq_quant_bwd: total and counter come together");
    B200_REQUIRE(Q_SHAPE_OK(c, ld_dy) && Q_SHAPE_OK(c, ld_x) && Q_SHAPE_OK(c, ld_dx), "lsq_quant_bwd: channels / pitches");
    const int grid = q_grid(rows * (c / 8), 256, qbwd_cap());
    cudaStream_t st = as_stream(stream);
    if (x_f32)
        launch_k(lsq_quant_bwd_kernel<float>, grid, 256, 0, st, static_cast<const __nv_bfloat16*>(dy), ld_dy, static_cast<const float*>(x), ld_x,
                                                          static_cast<__nv_bfloat16*>(dx), ld_dx, rows, c, step, qp, g, mode, part, total, counter);
    else
        launch_k(lsq_quant_bwd_kernel<__nv_bfloat16>, grid, 256, 0, st, static_cast<const __nv_bfloat16*>(dy), ld_dy, static_cast<const __nv_bfloat16*>(x), ld_x,
                                                                  static_cast<__nv_bfloat16*>(dx), ld_dx, rows, c, step, qp, g, mode, part, total, counter);
    B200_CHECK_LAUNCH("lsq_quant_bwd");
    return B200SD_OK;
}

extern "C" int b200sd_bit_split(const void* x, int32_t x_f32, int64_t ld_x, void* digits, int64_t ld_o, int64_t rows,
                                int32_t c, int32_t slices, int32_t slice_bit, int32_t x_bit, const float* step, float qp,
                                float g, void* stream) {
    B200_REQUIRE(x && digits && rows >= 0 && slices >= 1 && slice_bit >= 1 && x_bit >= 2, "bit_split: bad arguments");
    B200_REQUIRE(Q_SHAPE_OK(c, ld_x) && Q_SHAPE_OK(c, ld_o), "bit_split: channels and pitches must be multiples of 8");
    B200_REQUIRE(x_bit <= 9 && slice_bit <= 8, "bit_split: digits wider than 8 bits are not exact in bf16");
    B200_REQUIRE(!step || qp >= 1.f, "bit_split: bad quantiser range");
    if (rows == 0) return B200SD_OK;
    const int grid = q_grid(rows * (c / 8), 256);
    cudaStream_t st = as_stream(stream);
    if (x_f32)
        launch_k(bit_split_kernel<float>, grid, 256, 0, st, static_cast<const float*>(x), ld_x, static_cast<__nv_bfloat16*>(digits), ld_o,
                                                      rows, c, slices, slice_bit, x_bit, step != nullptr, step, qp, g);
    else
        launch_k(bit_split_kernel<__nv_bfloat16>, grid, 256, 0, st, static_cast<const __nv_bfloat16*>(x), ld_x, static_cast<__nv_bfloat16*>(digits),
                                                              ld_o, rows, c, slices, slice_bit, x_bit, step != nullptr, step, qp, g);
    B200_CHECK_LAUNCH("bit_split");
    return B200SD_OK;
}

extern "C" int b200sd_adc_combine(const float* partial, int64_t ld_p, float* out, int64_t ld_o, int64_t rows,
                                  int32_t ncol, int32_t col0, int32_t slices, int32_t slice_bit, const float* adc_scale,
                                  float adc_range, const float* gain_noise, const float* offset_noise,
                                  int32_t accumulate, void* stream) {
    B200_REQUIRE(partial && out && adc_scale && rows >= 0 && ncol > 0 && col0 >= 0 && slices >= 1 && slice_bit >= 1,
                 "adc_combine: bad arguments");
    B200_REQUIRE((gain_noise == nullptr) == (offset_noise == nullptr), "adc_combine: gain and offset noise come together");
    B200_REQUIRE(slices * slice_bit <= 24, "adc_combine: digit weights beyond 2^24 are not exact in fp32");
    if (rows == 0) return B200SD_OK;
    launch_k(adc_combine_kernel, q_grid(rows * ncol, 256), 256, 0, as_stream(stream), 
        partial, ld_p, out, ld_o, rows, ncol, col0, slices, slice_bit, adc_scale, adc_range, gain_noise, offset_noise,
        accumulate);
    B200_CHECK_LAUNCH("adc_combine");
    return B200SD_OK;
}

extern "C" int b200sd_adc_bwd_blocks(int64_t rows, int32_t ncol) { return q_grid(rows * ncol, 256, 4); }

extern "C" int b200sd_adc_bwd(const float* partial, int64_t ld_p, const float* dacc, int64_t ld_g, void* dpartial,
                              int64_t ld_d, int64_t rows, int32_t ncol, int32_t col0, int32_t slices, int32_t slice_bit,
                              const float* adc_scale, float adc_range, const float* gain_noise, const float* offset_noise,
                              float* part, float* total, uint32_t* counter, void* stream) {
    B200_REQUIRE(partial && dacc && dpartial && adc_scale && part && rows > 0 && ncol > 0 && col0 >= 0 && slices >= 1 &&
                     slice_bit >= 1, "adc_bwd: bad arguments");
    B200_REQUIRE((total == nullptr) == (counter == nullptr), "adc_bwd: total and counter come together");
    B200_REQUIRE((gain_noise == nullptr) == (offset_noise == nullptr), "adc_bwd: gain and offset noise come together");
    B200_REQUIRE(ld_d >= ncol, "adc_bwd: gradient pitch too small");
    launch_k(adc_bwd_kernel, q_grid(rows * ncol, 256, 4), 256, 0, as_stream(stream), 
        partial, ld_p, dacc, ld_g, static_cast<__nv_bfloat16*>(dpartial), ld_d, rows, ncol, col0, slices, slice_bit,
        adc_scale, adc_range, gain_noise, offset_noise, part, total, counter);
    B200_CHECK_LAUNCH("adc_bwd");
    return B200SD_OK;
}

extern "C" int b200sd_bit_split_bwd(const void* ddigits, int64_t ld_d, void* dcodes, int64_t ld_q, int64_t rows, int32_t c,
                                    int32_t slices, int32_t slice_bit, void* stream) {
    B200_REQUIRE(ddigits && dcodes && rows >= 0 && slices >= 1 && slice_bit >= 1, "bit_split_bwd: bad arguments");
    B200_REQUIRE(Q_SHAPE_OK(c, ld_d) && Q_SHAPE_OK(c, ld_q), "bit_split_bwd: channels and pitches must be multiples of 8");
    if (rows == 0) return B200SD_OK;
    launch_k(bit_split_bwd_kernel, q_grid(rows * (c / 8), 256), 256, 0, as_stream(stream), 
        static_cast<const __nv_bfloat16*>(ddigits), ld_d, static_cast<__nv_bfloat16*>(dcodes), ld_q, rows, c, slices,
        slice_bit);
    B200_CHECK_LAUNCH("bit_split_bwd");
    return B200SD_OK;
}

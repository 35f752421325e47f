// Output side of a compute-in-memory layer (reference cim_layers/layers_qn_lsq_adda_cim.py:482-500): the integer
// accumulation leaves the ADC as acc[M, N]; then
//     y   = acc / w_scale / in_scale / adc_scale + bias        (three true divisions in that order, then the add)
//     out = data_quant_lsq(y, step_out)                        (quant_noise_utils.py:101-120, isint = False)
// As torch ops that is 5 elementwise kernels forward and ~25 backward per layer (strided-broadcast divisions by 0-dim
// tensors, reciprocals, negations, the sums behind d(adc_scale)); here it is ONE launch each way.
//   forward   reads acc (fp32), writes out (bf16 quantised / bf16 plain / fp32 y for the lazy step-size initialisation).
//   backward  reads g = d out (bf16) and acc, recomputes y; writes d acc (fp32, the ADC backward's operand) and the
//             masked gradient d y (bf16; its column sums are d bias); per-CTA partials of d(step_out) and d(adc_scale)
//             are folded in fixed order by the last CTA.
// w_scale = 1 / step_weight and in_scale = 1 / step_input are detached in the reference: no gradient flows through them.
// Also here: get_adc_scale (layers_utils_adda.py:40-49) as one single-thread launch.
#include "host.h"
#include <cuda_bf16.h>

namespace b200 {

constexpr int AO_THREADS = 256;
constexpr int AO_MAX_CTAS = 4 * 148;

// scratch: [0] uint arrival counter, [1] pad, [2 ..) two partials per CTA
struct AoScratch {
    unsigned int arrive, pad;
    float part[2 * AO_MAX_CTAS];
};

__device__ __forceinline__ float ao_effective_step(float s, float g) {
    const float y = __fmul_rn(s, g);
    return __fadd_rn(__fsub_rn(s, y), y);
}

struct AoScales {
    float ws, is, alpha;
};
__device__ __forceinline__ AoScales ao_scales(const float* step_w, const float* step_in, const float* alpha) {
    AoScales k;
    k.ws = step_w ? __fdiv_rn(1.0f, *step_w) : 1.0f;   // (1.0 / step).detach()
    k.is = __fdiv_rn(1.0f, *step_in);
    k.alpha = *alpha;
    return k;
}
__device__ __forceinline__ float ao_rescale(float a, const AoScales& k) {
    return __fdiv_rn(__fdiv_rn(__fdiv_rn(a, k.ws), k.is), k.alpha);
}

__device__ __forceinline__ void ao_ld8(const float* p, float* f) {
    const float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
    f[0] = a.x, f[1] = a.y, f[2] = a.z, f[3] = a.w, f[4] = b.x, f[5] = b.y, f[6] = b.z, f[7] = b.w;
}
__device__ __forceinline__ void ao_ld8(const __nv_bfloat16* p, float* f) {
    const uint4 u = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float2 v = __bfloat1622float2(h[j]);
        f[2 * j] = v.x, f[2 * j + 1] = v.y;
    }
}
__device__ __forceinline__ void ao_st8(float* p, const float* f) {
    reinterpret_cast<float4*>(p)[0] = make_float4(f[0], f[1], f[2], f[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(f[4], f[5], f[6], f[7]);
}
__device__ __forceinline__ void ao_st8(__nv_bfloat16* p, const float* f) {
    uint4 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
    *reinterpret_cast<uint4*>(p) = u;
}

// mode 0: out = fp32 y;  mode 1: out = bf16 round(clamp(y / s', -Qp, Qp)) * s';  mode 2: out = bf16 y
template <typename TO>
__global__ void __launch_bounds__(AO_THREADS)
    adda_out_kernel(const float* __restrict__ acc, long long ld_a, const float* __restrict__ bias,
                    const float* __restrict__ step_w, const float* __restrict__ step_in, const float* __restrict__ alpha,
                    const float* __restrict__ step_out, float qp, float g, TO* __restrict__ out, long long ld_o,
                    long long rows, int c, int mode) {
    pdl_enter();
    const AoScales k = ao_scales(step_w, step_in, alpha);
    const float s = mode == 1 ? ao_effective_step(*step_out, g) : 1.f;
    const int nchunk = c >> 3;
    const long long total = rows * nchunk;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / nchunk;
        const int ch = static_cast<int>(i - r * nchunk);
        float f[8], b[8];
        ao_ld8(acc + r * ld_a + ch * 8, f);
        if (bias) ao_ld8(bias + ch * 8, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float y = ao_rescale(f[j], k);
            if (bias) y = __fadd_rn(y, b[j]);
            if (mode == 1) y = __fmul_rn(rintf(fminf(fmaxf(__fdiv_rn(y, s), -qp), qp)), s);
            f[j] = y;
        }
        ao_st8(out + r * ld_o + ch * 8, f);
    }
}

__device__ __forceinline__ bool ao_last_cta(unsigned int* counter, unsigned int total) {
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

__device__ __forceinline__ float ao_block_sum(float v, float* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = red[0];
#pragma unroll
    for (int w = 1; w < AO_THREADS / 32; ++w) r += red[w];
    return r;
}

// quant != 0: the LSQ output quantiser sits behind y (mask + step gradient); res[0] = d(step_out) * g, res[1] = d(alpha)
__global__ void __launch_bounds__(AO_THREADS)
    adda_out_bwd_kernel(const __nv_bfloat16* __restrict__ gout, long long ld_g, const float* __restrict__ acc,
                        long long ld_a, const float* __restrict__ bias, const float* __restrict__ step_w,
                        const float* __restrict__ step_in, const float* __restrict__ alpha,
                        const float* __restrict__ step_out, float qp, float g, int quant, float* __restrict__ dacc,
                        __nv_bfloat16* __restrict__ dy, long long rows, int c, AoScratch* __restrict__ sc,
                        float* __restrict__ res) {
    pdl_enter();
    __shared__ float red[AO_THREADS / 32];
    const AoScales k = ao_scales(step_w, step_in, alpha);
    const float s = quant ? ao_effective_step(*step_out, g) : 1.f;
    const float inv_alpha = 1.f / k.alpha, inv_all = 1.f / (k.alpha * k.is * k.ws);
    const int nchunk = c >> 3;
    const long long total = rows * nchunk;
    float a_s = 0.f, a_al = 0.f;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / nchunk;
        const int ch = static_cast<int>(i - r * nchunk);
        float f[8], b[8], gg[8];
        ao_ld8(acc + r * ld_a + ch * 8, f);
        ao_ld8(gout + r * ld_g + ch * 8, gg);
        if (bias) ao_ld8(bias + ch * 8, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float y2 = ao_rescale(f[j], k);
            if (quant) {
                const float y = bias ? __fadd_rn(y2, b[j]) : y2;
                const float u = __fdiv_rn(y, s);
                const bool in = (u >= -qp) && (u <= qp);
                const float q = rintf(fminf(fmaxf(u, -qp), qp));
                a_s += gg[j] * (q - (in ? u : 0.f));
                gg[j] = in ? gg[j] : 0.f;
            }
            // y2 = (acc / ws / is) / alpha:  d alpha = -g * y2 / alpha,  d acc = g / (alpha * is * ws). The forward
            // value (and with it the clamp mask) is recomputed with the reference's divisions; the two gradient
            // factors use one reciprocal each, hoisted out of the loop (8 -> 4 IEEE divisions per element).
            a_al -= gg[j] * (y2 * inv_alpha);
            f[j] = gg[j] * inv_all;
        }
        ao_st8(dacc + r * static_cast<long long>(c) + ch * 8, f);
        if (dy) ao_st8(dy + r * static_cast<long long>(c) + ch * 8, gg);
    }
    a_s = ao_block_sum(a_s, red);
    a_al = ao_block_sum(a_al, red);
    if (threadIdx.x == 0) sc->part[2 * blockIdx.x] = a_s, sc->part[2 * blockIdx.x + 1] = a_al;
    if (ao_last_cta(&sc->arrive, gridDim.x)) {
        float v0 = 0.f, v1 = 0.f;
        for (unsigned int t = threadIdx.x; t < gridDim.x; t += AO_THREADS)
            v0 += __ldcg(sc->part + 2 * t), v1 += __ldcg(sc->part + 2 * t + 1);
        v0 = ao_block_sum(v0, red);
        v1 = ao_block_sum(v1, red);
        if (threadIdx.x == 0) res[0] = v0 * g, res[1] = v1;
    }
}

// get_adc_scale: clamp (straight-through: the VALUE is fl(fl(gc - g) + g)), then the rounded gain ('gain' mode) or the
// rounded range 1 / gain (other modes), times the scale of gain 1. out[0] = adc_scale, out[1] = d adc_scale / d gain.
__global__ void adc_scale_kernel(const float* __restrict__ gain, float gmin, float gmax, int gain_mode, float scale1,
                                 float* __restrict__ out) {
    pdl_enter();
    const float g0 = *gain;
    const float gc = fminf(fmaxf(g0, gmin), gmax);
    const float g = __fadd_rn(__fsub_rn(gc, g0), g0);
    float v, d;
    if (gain_mode) {
        v = __fadd_rn(__fsub_rn(rintf(g), g), g);
        d = 1.f;
    } else {
        const float r = __fdiv_rn(1.0f, g);
        const float rr = __fadd_rn(__fsub_rn(rintf(r), r), r);
        v = __fdiv_rn(1.0f, rr);
        d = 1.f / (rr * rr * g * g);   // d(1/rr)/d rr * d r/d g = (-1/rr^2) * (-1/g^2)
    }
    out[0] = __fmul_rn(v, scale1);
    out[1] = d * scale1;
}

static int ao_grid(long long items) {
    long long b = (items + AO_THREADS - 1) / AO_THREADS;
    const long long cap = static_cast<long long>(4) * sm_count();
    if (b > cap) b = cap;
    if (b > AO_MAX_CTAS) b = AO_MAX_CTAS;
    if (b < 1) b = 1;
    return static_cast<int>(b);
}

}  // namespace b200

using namespace b200;

#define AO_SHAPE_OK(c, ld) ((c) > 0 && (c) % 8 == 0 && (ld) % 8 == 0 && (ld) >= (c))

extern "C" int b200sd_adda_out_scratch_floats(void) { return static_cast<int>(sizeof(AoScratch) / sizeof(float)); }

extern "C" int b200sd_adda_out(const float* acc, int64_t ld_a, const float* bias, const float* step_w,
                               const float* step_in, const float* adc_scale, const float* step_out, float qp,
                               float grad_scale, void* out, int64_t ld_o, int64_t rows, int32_t c, int32_t mode,
                               void* stream) {
    B200_REQUIRE(acc && step_in && adc_scale && out && rows >= 0 && mode >= 0 && mode <= 2, "adda_out: bad arguments");
    B200_REQUIRE(mode != 1 || (step_out && qp >= 1.f), "adda_out: the output quantiser needs its step size and range");
    B200_REQUIRE(AO_SHAPE_OK(c, ld_a) && AO_SHAPE_OK(c, ld_o), "adda_out: columns and pitches must be multiples of 8");
    if (rows == 0) return B200SD_OK;
    const int grid = ao_grid(rows * (c / 8));
    cudaStream_t st = as_stream(stream);
    if (mode == 0)
        launch_k(adda_out_kernel<float>, grid, AO_THREADS, 0, st, acc, ld_a, bias, step_w, step_in, adc_scale, step_out, qp,
                 grad_scale, static_cast<float*>(out), ld_o, rows, c, mode);
    else
        launch_k(adda_out_kernel<__nv_bfloat16>, grid, AO_THREADS, 0, st, acc, ld_a, bias, step_w, step_in, adc_scale,
                 step_out, qp, grad_scale, static_cast<__nv_bfloat16*>(out), ld_o, rows, c, mode);
    B200_CHECK_LAUNCH("adda_out");
    return B200SD_OK;
}

extern "C" int b200sd_adda_out_bwd(const void* gout, int64_t ld_g, const float* acc, int64_t ld_a, const float* bias,
                                   const float* step_w, const float* step_in, const float* adc_scale,
                                   const float* step_out, float qp, float grad_scale, float* dacc, void* dy,
                                   int64_t rows, int32_t c, float* scratch, float* res, void* stream) {
    B200_REQUIRE(gout && acc && step_in && adc_scale && dacc && scratch && res && rows > 0, "adda_out_bwd: bad arguments");
    B200_REQUIRE(!step_out || qp >= 1.f, "adda_out_bwd: bad quantiser range");
    B200_REQUIRE(AO_SHAPE_OK(c, ld_a) && AO_SHAPE_OK(c, ld_g), "adda_out_bwd: columns and pitches must be multiples of 8");
    launch_k(adda_out_bwd_kernel, ao_grid(rows * (c / 8)), AO_THREADS, 0, as_stream(stream),
             static_cast<const __nv_bfloat16*>(gout), ld_g, acc, ld_a, bias, step_w, step_in, adc_scale, step_out, qp,
             grad_scale, step_out != nullptr ? 1 : 0, dacc, static_cast<__nv_bfloat16*>(dy), rows, c,
             reinterpret_cast<AoScratch*>(scratch), res);
    B200_CHECK_LAUNCH("adda_out_bwd");
    return B200SD_OK;
}

extern "C" int b200sd_adc_scale(const float* gain, float gain_min, float gain_max, int32_t gain_mode, float scale1,
                                float* out, void* stream) {
    B200_REQUIRE(gain && out && gain_min > 0.f && gain_max >= gain_min, "adc_scale: bad arguments");
    launch_k(adc_scale_kernel, 1, 1, 0, as_stream(stream), gain, gain_min, gain_max, gain_mode, scale1, out);
    B200_CHECK_LAUNCH("adc_scale");
    return B200SD_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// Single-block layers (the whole weight fits one array: every Linear of the DiT-9L except the MLP's second): ADC of the
// slices, their recombination, the division chain, the bias and the output quantiser in ONE pass over the fp32 block
// product — the integer accumulation `acc` and its gradient are never written to memory. Same arithmetic, in the same
// order, as adc_combine_kernel followed by adda_out_kernel (csrc/quant.cu, above); the backward is adda_out_bwd_kernel
// followed by adc_bwd_kernel.
namespace b200 {

__device__ __forceinline__ float at_adc(float p, float alpha, float R, bool noisy, float gn, float on) {
    float v = __fmul_rn(alpha, p);
    if (noisy) {
        v = __fmul_rn(v, __fadd_rn(1.0f, gn));
        v = __fadd_rn(v, __fmul_rn(R, on));
    }
    return v;
}

// part fp32 [slices * rows, ld_p]; out bf16 [rows, c]; one thread = 4 consecutive columns of one row
__global__ void __launch_bounds__(AO_THREADS)
    adda_tail_kernel(const float* __restrict__ part, long long ld_p, const float* __restrict__ bias,
                     const float* __restrict__ step_w, const float* __restrict__ step_in, const float* __restrict__ alpha_p,
                     const float* __restrict__ step_out, float qp, float g, float R, int slices, int slice_bit,
                     const float* __restrict__ gain_noise, const float* __restrict__ offset_noise,
                     __nv_bfloat16* __restrict__ out, long long rows, int c) {
    pdl_enter();
    const AoScales k = ao_scales(step_w, step_in, alpha_p);
    const float s = step_out ? ao_effective_step(*step_out, g) : 1.f;
    const bool noisy = gain_noise != nullptr;
    const int nq = c >> 2;
    const long long total = rows * nq;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / nq;
        const int col = static_cast<int>(i - r * nq) * 4;
        float sum[4] = {0.f, 0.f, 0.f, 0.f};
        for (int sl = 0; sl < slices; ++sl) {
            const float4 p = *reinterpret_cast<const float4*>(part + (static_cast<long long>(sl) * rows + r) * ld_p + col);
            const float pv[4] = {p.x, p.y, p.z, p.w};
            const float wgt = static_cast<float>(1 << (sl * slice_bit));
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float v = at_adc(pv[j], k.alpha, R, noisy, noisy ? gain_noise[col + j] : 0.f,
                                       noisy ? offset_noise[col + j] : 0.f);
                sum[j] += rintf(fminf(fmaxf(v, -R - 1.0f), R)) * wgt;   // integers: exact in fp32
            }
        }
        __nv_bfloat162 o[2];
        float y[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float v = ao_rescale(sum[j], k);
            if (bias) v = __fadd_rn(v, bias[col + j]);
            if (step_out) v = __fmul_rn(rintf(fminf(fmaxf(__fdiv_rn(v, s), -qp), qp)), s);
            y[j] = v;
        }
        o[0] = __floats2bfloat162_rn(y[0], y[1]), o[1] = __floats2bfloat162_rn(y[2], y[3]);
        *reinterpret_cast<uint2*>(out + r * static_cast<long long>(c) + col) = *reinterpret_cast<const uint2*>(o);
    }
}

// res[0] = d(step_out) * g, res[1] = d(adc_scale) (through the division AND through the ADC's scaled input)
__global__ void __launch_bounds__(AO_THREADS)
    adda_tail_bwd_kernel(const __nv_bfloat16* __restrict__ gout, long long ld_g, const float* __restrict__ part,
                         long long ld_p, const float* __restrict__ bias, const float* __restrict__ step_w,
                         const float* __restrict__ step_in, const float* __restrict__ alpha_p,
                         const float* __restrict__ step_out, float qp, float g, float R, int slices, int slice_bit,
                         const float* __restrict__ gain_noise, const float* __restrict__ offset_noise,
                         __nv_bfloat16* __restrict__ dp, long long ld_d, __nv_bfloat16* __restrict__ dy, long long rows,
                         int c, AoScratch* __restrict__ sc, float* __restrict__ res) {
    pdl_enter();
    __shared__ float red[AO_THREADS / 32];
    const AoScales k = ao_scales(step_w, step_in, alpha_p);
    const float s = step_out ? ao_effective_step(*step_out, g) : 1.f;
    const float inv_alpha = 1.f / k.alpha, inv_all = 1.f / (k.alpha * k.is * k.ws);
    const bool noisy = gain_noise != nullptr;
    const int nq = c >> 2;
    const long long total = rows * nq;
    float a_s = 0.f, a_al = 0.f;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / nq;
        const int col = static_cast<int>(i - r * nq) * 4;
        // (the slices are walked twice — the second read of the 16 bytes hits L1 — rather than kept in per-thread arrays
        // indexed by a run-time slice count, which would live in local memory)
        float sum[4] = {0.f, 0.f, 0.f, 0.f};
        for (int sl = 0; sl < slices; ++sl) {
            const float4 p = *reinterpret_cast<const float4*>(part + (static_cast<long long>(sl) * rows + r) * ld_p + col);
            const float pv[4] = {p.x, p.y, p.z, p.w};
            const float wgt = static_cast<float>(1 << (sl * slice_bit));
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float v = at_adc(pv[j], k.alpha, R, noisy, noisy ? gain_noise[col + j] : 0.f,
                                       noisy ? offset_noise[col + j] : 0.f);
                sum[j] += rintf(fminf(fmaxf(v, -R - 1.0f), R)) * wgt;
            }
        }
        const uint2 gu = *reinterpret_cast<const uint2*>(gout + r * ld_g + col);
        const __nv_bfloat162* gh = reinterpret_cast<const __nv_bfloat162*>(&gu);
        const float2 g01 = __bfloat1622float2(gh[0]), g23 = __bfloat1622float2(gh[1]);
        float gg[4] = {g01.x, g01.y, g23.x, g23.y};
        float dacc[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float y2 = ao_rescale(sum[j], k);
            if (step_out) {
                const float y = bias ? __fadd_rn(y2, bias[col + j]) : y2;
                const float u = __fdiv_rn(y, s);
                const bool in = (u >= -qp) && (u <= qp);
                const float q = rintf(fminf(fmaxf(u, -qp), qp));
                a_s += gg[j] * (q - (in ? u : 0.f));
                gg[j] = in ? gg[j] : 0.f;
            }
            a_al -= gg[j] * (y2 * inv_alpha);
            dacc[j] = gg[j] * inv_all;
        }
        if (dy) {
            __nv_bfloat162 o[2] = {__floats2bfloat162_rn(gg[0], gg[1]), __floats2bfloat162_rn(gg[2], gg[3])};
            *reinterpret_cast<uint2*>(dy + r * static_cast<long long>(c) + col) = *reinterpret_cast<const uint2*>(o);
        }
        for (int sl = 0; sl < slices; ++sl) {
            const float4 p = *reinterpret_cast<const float4*>(part + (static_cast<long long>(sl) * rows + r) * ld_p + col);
            const float pv[4] = {p.x, p.y, p.z, p.w};
            const float wgt = static_cast<float>(1 << (sl * slice_bit));
            float d[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float v = at_adc(pv[j], k.alpha, R, noisy, noisy ? gain_noise[col + j] : 0.f,
                                       noisy ? offset_noise[col + j] : 0.f);
                const bool in = (v >= -R - 1.0f) && (v <= R);
                const float gk = in ? dacc[j] * wgt : 0.f;
                a_al += gk * pv[j];
                d[j] = k.alpha * gk;
            }
            __nv_bfloat162 o[2] = {__floats2bfloat162_rn(d[0], d[1]), __floats2bfloat162_rn(d[2], d[3])};
            *reinterpret_cast<uint2*>(dp + (static_cast<long long>(sl) * rows + r) * ld_d + col) = *reinterpret_cast<const uint2*>(o);
        }
    }
    a_s = ao_block_sum(a_s, red);
    a_al = ao_block_sum(a_al, red);
    if (threadIdx.x == 0) sc->part[2 * blockIdx.x] = a_s, sc->part[2 * blockIdx.x + 1] = a_al;
    if (ao_last_cta(&sc->arrive, gridDim.x)) {
        float v0 = 0.f, v1 = 0.f;
        for (unsigned int t = threadIdx.x; t < gridDim.x; t += AO_THREADS)
            v0 += __ldcg(sc->part + 2 * t), v1 += __ldcg(sc->part + 2 * t + 1);
        v0 = ao_block_sum(v0, red);
        v1 = ao_block_sum(v1, red);
        if (threadIdx.x == 0) res[0] = v0 * g, res[1] = v1;
    }
}

}  // namespace b200

extern "C" int b200sd_adda_tail(const float* part, int64_t ld_p, const float* bias, const float* step_w,
                                const float* step_in, const float* adc_scale, const float* step_out, float qp,
                                float grad_scale, float adc_range, int32_t slices, int32_t slice_bit,
                                const float* gain_noise, const float* offset_noise, void* out, int64_t rows, int32_t c,
                                void* stream) {
    B200_REQUIRE(part && step_in && adc_scale && out && rows >= 0 && slices >= 1 && slices <= 8 && slice_bit >= 1,
                 "adda_tail: bad arguments");
    B200_REQUIRE(!step_out || qp >= 1.f, "adda_tail: bad quantiser range");
    B200_REQUIRE((gain_noise == nullptr) == (offset_noise == nullptr), "adda_tail: gain and offset noise come together");
    B200_REQUIRE(slices * slice_bit <= 24, "adda_tail: digit weights beyond 2^24 are not exact in fp32");
    B200_REQUIRE(c > 0 && c % 4 == 0 && ld_p % 4 == 0 && ld_p >= c && (reinterpret_cast<uintptr_t>(part) & 15) == 0,
                 "adda_tail: columns / pitch must be multiples of 4 (16-byte loads)");
    if (rows == 0) return B200SD_OK;
    launch_k(adda_tail_kernel, ao_grid(rows * (c / 4)), AO_THREADS, 0, as_stream(stream), part, ld_p, bias, step_w, step_in,
             adc_scale, step_out, qp, grad_scale, adc_range, slices, slice_bit, gain_noise, offset_noise,
             static_cast<__nv_bfloat16*>(out), rows, c);
    B200_CHECK_LAUNCH("adda_tail");
    return B200SD_OK;
}

extern "C" int b200sd_adda_tail_bwd(const void* gout, int64_t ld_g, const float* part, int64_t ld_p, const float* bias,
                                    const float* step_w, const float* step_in, const float* adc_scale,
                                    const float* step_out, float qp, float grad_scale, float adc_range, int32_t slices,
                                    int32_t slice_bit, const float* gain_noise, const float* offset_noise,
                                    void* dpartial, int64_t ld_d, void* dy, int64_t rows, int32_t c, float* scratch,
                                    float* res, void* stream) {
    B200_REQUIRE(gout && part && step_in && adc_scale && dpartial && scratch && res && rows > 0 && slices >= 1 &&
                     slices <= 8 && slice_bit >= 1, "adda_tail_bwd: bad arguments");
    B200_REQUIRE(!step_out || qp >= 1.f, "adda_tail_bwd: bad quantiser range");
    B200_REQUIRE((gain_noise == nullptr) == (offset_noise == nullptr), "adda_tail_bwd: gain and offset noise come together");
    B200_REQUIRE(c > 0 && c % 4 == 0 && ld_p % 4 == 0 && ld_p >= c && ld_g % 4 == 0 && ld_g >= c && ld_d % 4 == 0 &&
                     ld_d >= c && (reinterpret_cast<uintptr_t>(part) & 15) == 0,
                 "adda_tail_bwd: columns / pitches must be multiples of 4");
    launch_k(adda_tail_bwd_kernel, ao_grid(rows * (c / 4)), AO_THREADS, 0, as_stream(stream),
             static_cast<const __nv_bfloat16*>(gout), ld_g, part, ld_p, bias, step_w, step_in, adc_scale, step_out, qp,
             grad_scale, adc_range, slices, slice_bit, gain_noise, offset_noise, static_cast<__nv_bfloat16*>(dpartial),
             ld_d, static_cast<__nv_bfloat16*>(dy), rows, c, reinterpret_cast<AoScratch*>(scratch), res);
    B200_CHECK_LAUNCH("adda_tail_bwd");
    return B200SD_OK;
}

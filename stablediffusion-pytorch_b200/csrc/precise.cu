// Precise mode (DESIGN §4 "error budget"): the residual stream and every stored GEMM result stay fp32, and every
// tensor-core operand is SPLIT into two bf16 values, hi = bf16(v) and lo = bf16(v - hi) (16 mantissa bits together), so
// that a product runs on the same tcgen05 kind::f16 GEMM as A_hi*[W_hi|W_lo] + A_lo*W_hi with fp32 accumulation.
// These are the two producers of split A operands: GroupNorm (+activation) of an fp32 tensor, and a plain split.
// Not a fast path: HBM-bound helpers written for accuracy (fp32 two-pass statistics), sized for parity runs.
#include "host.h"
#include <cuda_bf16.h>

namespace b200 {

__device__ __forceinline__ float p_act(float v, int act) {
    if (act == B200SD_ACT_SILU) return v / (1.f + expf(-v));   // IEEE division, libm-grade exp: accuracy over speed
    if (act == B200SD_ACT_RELU) return fmaxf(v, 0.f);
    return v;
}

__device__ __forceinline__ void put_split(__nv_bfloat16* row, int hi_col, int lo_col, float v) {
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    row[hi_col] = hi;
    if (lo_col >= 0) row[lo_col] = __float2bfloat16_rn(v - __bfloat162float(hi));
}

__device__ __forceinline__ float block_sum(float v, float* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();   // red may still be read by the previous call
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float s = 0.f;
    for (int i = 0; i < nw; ++i) s += red[i];   // fixed order: every thread gets the same value
    return s;
}

// CTA = (group, image). Three passes over the group's hw x cpg fp32 slab (L2-resident at parity sizes).
__global__ void __launch_bounds__(256)
    groupnorm_split_kernel(const float* __restrict__ x, long long ld_x, __nv_bfloat16* __restrict__ y, long long ld_y,
                           int lo_off, const float* __restrict__ gamma, const float* __restrict__ beta, long long hw,
                           int c, int cpg, float eps, int act) {
    pdl_enter();
    __shared__ float red[8];
    const int g = blockIdx.x, img = blockIdx.y;
    const float* xs = x + static_cast<long long>(img) * hw * ld_x + g * cpg;
    __nv_bfloat16* ys = y + static_cast<long long>(img) * hw * ld_y;
    const long long total = hw * cpg;
    float s = 0.f;
    for (long long i = threadIdx.x; i < total; i += blockDim.x) s += xs[(i / cpg) * ld_x + (i % cpg)];
    const float mean = block_sum(s, red) / static_cast<float>(total);
    float q = 0.f;
    for (long long i = threadIdx.x; i < total; i += blockDim.x) {
        const float d = xs[(i / cpg) * ld_x + (i % cpg)] - mean;
        q = fmaf(d, d, q);
    }
    const float rstd = 1.f / sqrtf(block_sum(q, red) / static_cast<float>(total) + eps);
    for (long long i = threadIdx.x; i < total; i += blockDim.x) {
        const long long r = i / cpg;
        const int j = static_cast<int>(i % cpg), ch = g * cpg + j;
        const float v = (xs[r * ld_x + j] - mean) * rstd * (gamma ? gamma[ch] : 1.f) + (beta ? beta[ch] : 0.f);
        put_split(ys + r * ld_y, ch, lo_off >= 0 ? lo_off + ch : -1, p_act(v, act));
    }
}

__global__ void split_f32_kernel(const float* __restrict__ src, long long ld_s, __nv_bfloat16* __restrict__ dst,
                                 long long ld_d, int hi_off, int lo_off, long long rows, int c, int act) {
    pdl_enter();
    const long long total = rows * c;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / c;
        const int j = static_cast<int>(i - r * c);
        put_split(dst + r * ld_d, hi_off + j, lo_off >= 0 ? lo_off + j : -1, p_act(src[r * ld_s + j], act));
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200sd_groupnorm_split(const float* x, int64_t ld_x, void* y, int64_t ld_y, int32_t lo_off,
                                      const float* gamma, const float* beta, int64_t n, int64_t hw, int c, int groups,
                                      float eps, int act, void* stream) {
    B200_REQUIRE(x && y, "groupnorm_split: null pointer");
    B200_REQUIRE(groups > 0 && c > 0 && c % groups == 0, "groupnorm_split: %d channels not divisible by %d groups", c,
                 groups);
    B200_REQUIRE(ld_x >= c && ld_y >= c && (lo_off < 0 || (lo_off >= c && ld_y >= lo_off + c)),
                 "groupnorm_split: output pitch %lld cannot hold hi [0,%d) and lo [%d,%d)", (long long)ld_y, c, lo_off,
                 lo_off + c);
    B200_REQUIRE(n <= 65535 && groups <= 65535, "groupnorm_split: grid limit");
    if (n == 0 || hw == 0) return B200SD_OK;
    launch_k(groupnorm_split_kernel, dim3((unsigned)groups, (unsigned)n), 256, 0, as_stream(stream), x, ld_x,
             static_cast<__nv_bfloat16*>(y), ld_y, lo_off, gamma, beta, hw, c, c / groups, eps, act);
    B200_CHECK_LAUNCH("groupnorm_split");
    return B200SD_OK;
}

extern "C" int b200sd_split_f32(const float* src, int64_t ld_s, void* dst, int64_t ld_d, int32_t hi_off, int32_t lo_off,
                                int64_t rows, int32_t c, int act, void* stream) {
    B200_REQUIRE(src && dst, "split_f32: null pointer");
    B200_REQUIRE(rows >= 0 && c > 0 && ld_s >= c && hi_off >= 0 && ld_d >= hi_off + c &&
                     (lo_off < 0 || (ld_d >= lo_off + c && (lo_off >= hi_off + c || lo_off + c <= hi_off))),
                 "split_f32: bad geometry (c %d, hi at %d, lo at %d, pitch %lld)", c, hi_off, lo_off, (long long)ld_d);
    if (rows == 0) return B200SD_OK;
    long long blocks = (rows * c + 255) / 256;
    if (blocks > 8ll * sm_count()) blocks = 8ll * sm_count();
    launch_k(split_f32_kernel, (int)blocks, 256, 0, as_stream(stream), src, ld_s, static_cast<__nv_bfloat16*>(dst), ld_d,
             hi_off, lo_off, rows, c, act);
    B200_CHECK_LAUNCH("split_f32");
    return B200SD_OK;
}

// Glue of the Enhance wrappers (reference cim_qn_train/layers_enhance.py):
//   * GroupedLinear / grouped 1x1 Conv2d (:324-365, :368-410, :439-535) run as ONE dense tcgen05 GEMM over the
//     block-diagonal form of their (groups, out_group, in_group) parameter. The dense bf16 matrix (and its transpose for
//     the input gradient) lives in a persistent zero-initialised buffer per module; each forward rewrites only the
//     diagonal blocks (blockdiag_pack), the weight gradient is read back out of the dense wgrad (blockdiag_gather).
//     Before: zero fill of the fp32 dense matrix + index_put + cast + two packing copies per map and step.
//   * merge_feature_map (:596-608): out = orig * (1 - s) + enh * s with s = sigmoid(blend_factor) (or blend_factor), and
//     the std / mean of both operands the progressive trainer logs (progressive_qn_train.py:990-996) — one launch
//     forward (blend + four moments, folded by the last CTA in double), one backward (both input gradients and
//     d blend_factor).
#include "host.h"
#include <cuda_bf16.h>

namespace b200 {

constexpr int EN_THREADS = 256;
constexpr int EN_MAX_CTAS = 4 * 148;

__global__ void blockdiag_pack_kernel(const float* __restrict__ w, int g, int og, int ig, __nv_bfloat16* __restrict__ dense,
                                      long long ld_d, __nv_bfloat16* __restrict__ dense_t, long long ld_t) {
    pdl_enter();
    const long long total = static_cast<long long>(g) * og * ig;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int c = static_cast<int>(i % ig);
        const long long t = i / ig;
        const int r = static_cast<int>(t % og), gi = static_cast<int>(t / og);
        const __nv_bfloat16 v = __float2bfloat16_rn(w[i]);
        const long long row = static_cast<long long>(gi) * og + r, col = static_cast<long long>(gi) * ig + c;
        dense[row * ld_d + col] = v;
        if (dense_t) dense_t[col * ld_t + row] = v;
    }
}

__global__ void blockdiag_gather_kernel(const float* __restrict__ dense, long long ld, int g, int og, int ig,
                                        float* __restrict__ dw) {
    pdl_enter();
    const long long total = static_cast<long long>(g) * og * ig;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int c = static_cast<int>(i % ig);
        const long long t = i / ig;
        const int r = static_cast<int>(t % og), gi = static_cast<int>(t / og);
        dw[i] = dense[(static_cast<long long>(gi) * og + r) * ld + static_cast<long long>(gi) * ig + c];
    }
}

__device__ __forceinline__ void en_ld8(const __nv_bfloat16* p, float* f) {
    const uint4 u = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float2 v = __bfloat1622float2(h[j]);
        f[2 * j] = v.x, f[2 * j + 1] = v.y;
    }
}
__device__ __forceinline__ void en_st8(__nv_bfloat16* p, const float* f) {
    uint4 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
    *reinterpret_cast<uint4*>(p) = u;
}

__device__ __forceinline__ float en_factor(const float* blend, int sigmoid) {
    const float b = *blend;
    return sigmoid ? 1.f / (1.f + expf(-b)) : b;
}

__device__ __forceinline__ bool en_last_cta(unsigned int* counter, unsigned int total) {
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

template <typename T>
__device__ __forceinline__ T en_block_sum(T v, T* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    T r = red[0];
#pragma unroll
    for (int w = 1; w < EN_THREADS / 32; ++w) r += red[w];
    return r;
}

// out = orig * (1 - s) + enh * s (fp32, one rounding to bf16), optional ReLU; stats[0..3] = std(orig), mean(orig),
// std(enh), mean(enh) over the rows x c elements (unbiased std, like Tensor.std()).
__global__ void __launch_bounds__(EN_THREADS)
    blend_kernel(const __nv_bfloat16* __restrict__ orig, long long ld_a, const __nv_bfloat16* __restrict__ enh,
                 long long ld_b, const float* __restrict__ blend, int sigmoid, int relu, __nv_bfloat16* __restrict__ out,
                 long long ld_o, long long rows, int c, float* __restrict__ part, unsigned int* counter,
                 float* __restrict__ stats) {
    pdl_enter();
    __shared__ float red[EN_THREADS / 32];
    __shared__ double dred[EN_THREADS / 32];
    const float s = en_factor(blend, sigmoid), s1 = 1.f - s;
    const int nchunk = c >> 3;
    const long long total = rows * nchunk;
    float sa = 0.f, qa = 0.f, sb = 0.f, qb = 0.f;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / nchunk;
        const int ch = static_cast<int>(i - r * nchunk);
        float a[8], b[8];
        en_ld8(orig + r * ld_a + ch * 8, a);
        en_ld8(enh + r * ld_b + ch * 8, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            sa += a[j], qa += a[j] * a[j], sb += b[j], qb += b[j] * b[j];
            const float v = __fadd_rn(__fmul_rn(a[j], s1), __fmul_rn(b[j], s));
            a[j] = relu ? fmaxf(v, 0.f) : v;
        }
        en_st8(out + r * ld_o + ch * 8, a);
    }
    if (!stats) return;
    sa = en_block_sum(sa, red), qa = en_block_sum(qa, red), sb = en_block_sum(sb, red), qb = en_block_sum(qb, red);
    if (threadIdx.x == 0) {
        float* p = part + 4 * blockIdx.x;
        p[0] = sa, p[1] = qa, p[2] = sb, p[3] = qb;
    }
    if (en_last_cta(counter, gridDim.x)) {
        double v[4] = {0.0, 0.0, 0.0, 0.0};
        for (unsigned int t = threadIdx.x; t < gridDim.x; t += EN_THREADS)
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] += static_cast<double>(__ldcg(part + 4 * t + k));
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = en_block_sum(v[k], dred);
        if (threadIdx.x == 0) {
            const double n = static_cast<double>(rows) * c;
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                const double mean = v[2 * k] / n;
                const double var = n > 1.0 ? fmax(v[2 * k + 1] - n * mean * mean, 0.0) / (n - 1.0) : nan("");
                stats[2 * k] = static_cast<float>(sqrt(var));
                stats[2 * k + 1] = static_cast<float>(mean);
            }
        }
    }
}

// d orig = g (1 - s), d enh = g s, d blend_factor = sum g (enh - orig) * (sigmoid ? s (1 - s) : 1); `out` (the forward
// result) masks g where a fused ReLU clipped.
__global__ void __launch_bounds__(EN_THREADS)
    blend_bwd_kernel(const __nv_bfloat16* __restrict__ g, long long ld_g, const __nv_bfloat16* __restrict__ orig,
                     long long ld_a, const __nv_bfloat16* __restrict__ enh, long long ld_b,
                     const __nv_bfloat16* __restrict__ out, long long ld_o, const float* __restrict__ blend, int sigmoid,
                     __nv_bfloat16* __restrict__ d_orig, __nv_bfloat16* __restrict__ d_enh, long long rows, int c,
                     float* __restrict__ part, unsigned int* counter, float* __restrict__ d_blend) {
    pdl_enter();
    __shared__ float red[EN_THREADS / 32];
    const float s = en_factor(blend, sigmoid), s1 = 1.f - s;
    const int nchunk = c >> 3;
    const long long total = rows * nchunk;
    float acc = 0.f;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / nchunk;
        const int ch = static_cast<int>(i - r * nchunk);
        float gg[8], a[8], b[8], o[8];
        en_ld8(g + r * ld_g + ch * 8, gg);
        en_ld8(orig + r * ld_a + ch * 8, a);
        en_ld8(enh + r * ld_b + ch * 8, b);
        if (out) en_ld8(out + r * ld_o + ch * 8, o);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float gv = (out && !(o[j] > 0.f)) ? 0.f : gg[j];
            acc += gv * (b[j] - a[j]);
            a[j] = gv * s1, b[j] = gv * s;
        }
        en_st8(d_orig + r * static_cast<long long>(c) + ch * 8, a);
        en_st8(d_enh + r * static_cast<long long>(c) + ch * 8, b);
    }
    acc = en_block_sum(acc, red);
    if (threadIdx.x == 0) part[blockIdx.x] = acc;
    if (en_last_cta(counter, gridDim.x)) {
        float v = 0.f;
        for (unsigned int t = threadIdx.x; t < gridDim.x; t += EN_THREADS) v += __ldcg(part + t);
        v = en_block_sum(v, red);
        if (threadIdx.x == 0) *d_blend = sigmoid ? v * s * s1 : v;
    }
}

static int en_grid(long long items) {
    long long b = (items + EN_THREADS - 1) / EN_THREADS;
    const long long cap = static_cast<long long>(4) * sm_count();
    if (b > cap) b = cap;
    if (b > EN_MAX_CTAS) b = EN_MAX_CTAS;
    if (b < 1) b = 1;
    return static_cast<int>(b);
}

}  // namespace b200

using namespace b200;

extern "C" int b200sd_blockdiag_pack(const float* w, int32_t groups, int32_t out_group, int32_t in_group, void* dense,
                                     int64_t ld_d, void* dense_t, int64_t ld_t, void* stream) {
    B200_REQUIRE(w && dense && groups >= 1 && out_group >= 1 && in_group >= 1, "blockdiag_pack: bad arguments");
    B200_REQUIRE(ld_d >= static_cast<int64_t>(groups) * in_group, "blockdiag_pack: dense pitch too small");
    B200_REQUIRE(!dense_t || ld_t >= static_cast<int64_t>(groups) * out_group, "blockdiag_pack: transposed pitch too small");
    launch_k(blockdiag_pack_kernel, en_grid(static_cast<long long>(groups) * out_group * in_group), EN_THREADS, 0,
             as_stream(stream), w, groups, out_group, in_group, static_cast<__nv_bfloat16*>(dense), ld_d,
             static_cast<__nv_bfloat16*>(dense_t), ld_t);
    B200_CHECK_LAUNCH("blockdiag_pack");
    return B200SD_OK;
}

extern "C" int b200sd_blockdiag_gather(const float* dense, int64_t ld, int32_t groups, int32_t out_group,
                                       int32_t in_group, float* dw, void* stream) {
    B200_REQUIRE(dense && dw && groups >= 1 && out_group >= 1 && in_group >= 1, "blockdiag_gather: bad arguments");
    B200_REQUIRE(ld >= static_cast<int64_t>(groups) * in_group, "blockdiag_gather: pitch too small");
    launch_k(blockdiag_gather_kernel, en_grid(static_cast<long long>(groups) * out_group * in_group), EN_THREADS, 0,
             as_stream(stream), dense, ld, groups, out_group, in_group, dw);
    B200_CHECK_LAUNCH("blockdiag_gather");
    return B200SD_OK;
}

#define EN_SHAPE_OK(c, ld) ((c) > 0 && (c) % 8 == 0 && (ld) % 8 == 0 && (ld) >= (c))

extern "C" int b200sd_blend_part_floats(void) { return 4 * EN_MAX_CTAS; }

extern "C" int b200sd_blend(const void* orig, int64_t ld_a, const void* enh, int64_t ld_b, const float* blend,
                            int32_t sigmoid, int32_t relu, void* out, int64_t ld_o, int64_t rows, int32_t c, float* part,
                            uint32_t* counter, float* stats, void* stream) {
    B200_REQUIRE(orig && enh && blend && out && rows > 0, "blend: bad arguments");
    B200_REQUIRE(!stats || (part && counter), "blend: statistics need the partial buffer and a counter");
    B200_REQUIRE(EN_SHAPE_OK(c, ld_a) && EN_SHAPE_OK(c, ld_b) && EN_SHAPE_OK(c, ld_o), "blend: columns / pitches (multiples of 8)");
    launch_k(blend_kernel, en_grid(rows * (c / 8)), EN_THREADS, 0, as_stream(stream),
             static_cast<const __nv_bfloat16*>(orig), ld_a, static_cast<const __nv_bfloat16*>(enh), ld_b, blend, sigmoid,
             relu, static_cast<__nv_bfloat16*>(out), ld_o, rows, c, part, counter, stats);
    B200_CHECK_LAUNCH("blend");
    return B200SD_OK;
}

extern "C" int b200sd_blend_bwd(const void* g, int64_t ld_g, const void* orig, int64_t ld_a, const void* enh,
                                int64_t ld_b, const void* out, int64_t ld_o, const float* blend, int32_t sigmoid,
                                void* d_orig, void* d_enh, int64_t rows, int32_t c, float* part, uint32_t* counter,
                                float* d_blend, void* stream) {
    B200_REQUIRE(g && orig && enh && blend && d_orig && d_enh && part && counter && d_blend && rows > 0,
                 "blend_bwd: bad arguments");
    B200_REQUIRE(EN_SHAPE_OK(c, ld_g) && EN_SHAPE_OK(c, ld_a) && EN_SHAPE_OK(c, ld_b) && (!out || EN_SHAPE_OK(c, ld_o)),
                 "blend_bwd: columns / pitches (multiples of 8)");
    launch_k(blend_bwd_kernel, en_grid(rows * (c / 8)), EN_THREADS, 0, as_stream(stream),
             static_cast<const __nv_bfloat16*>(g), ld_g, static_cast<const __nv_bfloat16*>(orig), ld_a,
             static_cast<const __nv_bfloat16*>(enh), ld_b, static_cast<const __nv_bfloat16*>(out), ld_o, blend, sigmoid,
             static_cast<__nv_bfloat16*>(d_orig), static_cast<__nv_bfloat16*>(d_enh), rows, c, part, counter, d_blend);
    B200_CHECK_LAUNCH("blend_bwd");
    return B200SD_OK;
}

// VQVAE.quantize (models/vqvae.py:93-128): nearest codebook vector per latent pixel, the straight-through output
// x + (e - x), and the mean squared error that both quantisation losses equal in value. One thread per pixel, the
// codebook streamed through shared memory in tiles; distances are sums of squared differences in fp32 (the reference's
// cdist goes through ||x||^2 + ||e||^2 - 2 x.e; the argmin is the same except between numerically tied codes).
#include "host.h"

namespace b200 {

constexpr int VQ_MAXC = 16, VQ_TILE = 512, VQ_THREADS = 256;   // 32 KB of codes per tile

template <int CT>   // compile-time channel count (4 = every config of the reference), 0 = run-time c
__global__ void __launch_bounds__(VQ_THREADS)
    vq_nearest_kernel(const float* __restrict__ x, long long ld_x, const float* __restrict__ cb, int n_codes, int c,
                      long long rows, long long hw, long long* __restrict__ idx, float* __restrict__ zq_nchw,
                      float* __restrict__ partial, float* __restrict__ loss, unsigned int* counter) {
    pdl_enter();
    __shared__ __align__(16) float tile[VQ_TILE * VQ_MAXC];
    __shared__ float red[VQ_THREADS / 32];
    __shared__ unsigned int s_last;
    const long long m = static_cast<long long>(blockIdx.x) * VQ_THREADS + threadIdx.x;
    const bool live = m < rows;
    float xv[VQ_MAXC];
#pragma unroll
    for (int j = 0; j < VQ_MAXC; ++j) xv[j] = (live && j < c) ? x[m * ld_x + j] : 0.f;
    float best = 3.4e38f;
    int best_i = 0;
    for (int k0 = 0; k0 < n_codes; k0 += VQ_TILE) {
        const int nk = min(VQ_TILE, n_codes - k0);
        __syncthreads();
        for (int i = threadIdx.x; i < nk * c; i += VQ_THREADS) tile[i] = __ldg(cb + static_cast<long long>(k0) * c + i);
        __syncthreads();
        for (int k = 0; k < nk; ++k) {
            float d = 0.f;
            if (CT == 4) {
                const float4 e = reinterpret_cast<const float4*>(tile)[k];     // same address in every lane: broadcast
                const float t0 = xv[0] - e.x, t1 = xv[1] - e.y, t2 = xv[2] - e.z, t3 = xv[3] - e.w;
                d = fmaf(t3, t3, fmaf(t2, t2, fmaf(t1, t1, t0 * t0)));
            } else {
#pragma unroll
                for (int j = 0; j < VQ_MAXC; ++j)
                    if (j < c) {
                        const float t = xv[j] - tile[k * c + j];
                        d = fmaf(t, t, d);
                    }
            }
            if (d < best) best = d, best_i = k0 + k;      // strict: the first of equal minima, like torch.argmin
        }
    }
    float se = 0.f;
    if (live) {
        idx[m] = best_i;
        const long long b = m / hw, p = m - b * hw;
        for (int j = 0; j < c; ++j) {
            const float e = __ldg(cb + static_cast<long long>(best_i) * c + j);
            const float diff = __fsub_rn(e, xv[j]);
            zq_nchw[(b * c + j) * hw + p] = __fadd_rn(xv[j], diff);      // x + (quant_out - x).detach()
            se += diff * diff;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = se;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f;
        for (int i = 0; i < VQ_THREADS / 32; ++i) s += red[i];
        partial[blockIdx.x] = s;
        __threadfence();
        const unsigned int t = atomicAdd(counter, 1u);
        s_last = (t == gridDim.x - 1u) ? 1u : 0u;
        if (s_last) *counter = 0u;
    }
    __syncthreads();
    if (s_last && threadIdx.x == 0) {       // a few hundred partials, summed in order
        __threadfence();
        float s = 0.f;
        for (unsigned int i = 0; i < gridDim.x; ++i) s += __ldcg(partial + i);
        loss[0] = s / (static_cast<float>(rows) * static_cast<float>(c));
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200sd_vq_nearest(const float* x, int64_t ld_x, const float* codebook, int32_t n_codes, int32_t c,
                                 int64_t n, int64_t hw, int64_t* idx, float* zq_nchw, float* partial, float* loss,
                                 uint32_t* counter, void* stream) {
    B200_REQUIRE(x && codebook && idx && zq_nchw && partial && loss && counter, "vq_nearest: null pointer");
    B200_REQUIRE(c >= 1 && c <= VQ_MAXC && n_codes >= 1 && ld_x >= c, "vq_nearest: 1 <= channels <= %d required", VQ_MAXC);
    const long long rows = n * hw;
    if (rows == 0) return B200SD_OK;
    const long long blocks = (rows + VQ_THREADS - 1) / VQ_THREADS;
    B200_REQUIRE(blocks <= 0x7fffffffLL, "vq_nearest: too many rows");
    if (c == 4)
        launch_k(vq_nearest_kernel<4>, static_cast<unsigned>(blocks), VQ_THREADS, 0, as_stream(stream), x,
                 static_cast<long long>(ld_x), codebook, n_codes, c, rows, static_cast<long long>(hw),
                 reinterpret_cast<long long*>(idx), zq_nchw, partial, loss, counter);
    else
        launch_k(vq_nearest_kernel<0>, static_cast<unsigned>(blocks), VQ_THREADS, 0, as_stream(stream), x,
                 static_cast<long long>(ld_x), codebook, n_codes, c, rows, static_cast<long long>(hw),
                 reinterpret_cast<long long*>(idx), zq_nchw, partial, loss, counter);
    B200_CHECK_LAUNCH("vq_nearest");
    return B200SD_OK;
}

extern "C" int b200sd_vq_blocks(int64_t rows) { return static_cast<int>((rows + VQ_THREADS - 1) / VQ_THREADS); }

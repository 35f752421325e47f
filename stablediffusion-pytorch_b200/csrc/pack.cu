// Weight-side data movement of a training step, table driven:
//
//  pack_weights   fp32 parameters (reference layouts: Conv2d [Cout, Cin, kh, kw], ConvTranspose2d [Cin, Cout, 4, 4],
//                 Linear [N, K]) -> the bf16 K-major matrices igemm / wgrad read ([rows, taps * Cin_pad], Cin padded to
//                 64 per tap with zeros), for EVERY layer of the model in ONE launch. A training step re-packs all
//                 weights after each optimizer step; done leaf by leaf with torch ops that is ~5 small kernels per
//                 layer and direction (cast, permute, flip, pad, concat) — about 700 launches per step for the
//                 conditioned UNet. Forward, transposed (dgrad) and flipped-tap (conv dgrad) forms are all the same
//                 job type with different strides.
//  wgrad_scatter  split-K partials of the weight-gradient GEMM ([splits, n_out, k_total] packed like the forward
//                 weights) -> the gradient in the parameter's own layout, summed in fixed split order, written
//                 straight into the flat gradient buffer the optimizer / all-reduce reads.
#include "host.h"
#include <cuda_bf16.h>

namespace b200 {

constexpr int PK_ROWS = 32, PK_CH = 32, PK_MAX_TAPS = 16;

struct PackJob {               // = b200sd_pack_job (include/b200sd.h), mirrored by ctypes (runtime.PackJob)
    const float* src;
    __nv_bfloat16* dst;
    long long s_r, s_c;        // source strides (elements) of the packed row index / channel index
    long long dst_ld;          // destination row pitch (elements); dst already points at the first column of the job
    int rows, cin, cp, taps;
    int tap_off[PK_MAX_TAPS];  // source offset of tap t
    int tile0;                 // first tile of this job in the launch
    int tiles_c;               // channel tiles per row tile (cp / 32)
};
static_assert(sizeof(PackJob) == sizeof(b200sd_pack_job), "PackJob must match the C-ABI struct");

// tile = 32 packed rows x 32 channels x all taps
__global__ void __launch_bounds__(256) pack_weights_kernel(const PackJob* __restrict__ jobs, int n_jobs) {
    pdl_enter();
    __shared__ __nv_bfloat16 tile[PK_ROWS][PK_MAX_TAPS * PK_CH + 2];
    // binary search: last job with tile0 <= blockIdx.x
    int lo = 0, hi = n_jobs - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (jobs[mid].tile0 <= static_cast<int>(blockIdx.x)) lo = mid; else hi = mid - 1;
    }
    const PackJob j = jobs[lo];
    const int local = blockIdx.x - j.tile0;
    const int r0 = (local / j.tiles_c) * PK_ROWS, c0 = (local % j.tiles_c) * PK_CH;
    const int tid = threadIdx.x;
    // lanes run along whichever of (row, channel) is closer in the source
    const bool lanes_c = j.s_c <= j.s_r;
    for (int i = tid; i < PK_ROWS * PK_CH; i += 256) {
        const int a = i & 31, b = i >> 5;
        const int r = lanes_c ? b : a, c = lanes_c ? a : b;
        const int gr = r0 + r, gc = c0 + c;
        const bool ok = gr < j.rows && gc < j.cin;
        const float* s = j.src + gr * j.s_r + gc * j.s_c;
        for (int t = 0; t < j.taps; ++t)
            tile[r][t * PK_CH + c] = __float2bfloat16_rn(ok ? __ldg(s + j.tap_off[t]) : 0.f);
    }
    __syncthreads();
    for (int i = tid; i < PK_ROWS * j.taps * (PK_CH / 2); i += 256) {
        const int c2 = i % (PK_CH / 2);
        const int t = (i / (PK_CH / 2)) % j.taps;
        const int r = i / ((PK_CH / 2) * j.taps);
        if (r0 + r < j.rows) {
            const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&tile[r][t * PK_CH + 2 * c2]);
            *reinterpret_cast<__nv_bfloat162*>(j.dst + (r0 + r) * j.dst_ld + t * j.cp + c0 + 2 * c2) = v;
        }
    }
}

struct ScatterArgs {
    const float* partial;      // [splits][n_out][k_total]
    float* out;
    long long s_co, s_ci;      // destination strides (elements) of the output-row index / channel index
    int splits, n_out, k_total, col0, cin, cp, taps, accumulate;
    int tap_off[PK_MAX_TAPS];  // destination offset of tap t
};

// One CTA: one output row x SC_CH channels x all taps. Reads run along the packed K axis (channels contiguous per
// tap), writes along the parameter's own innermost axis; the tile turns around in shared memory.
constexpr int SC_CH = 128;

__global__ void __launch_bounds__(256) wgrad_scatter_kernel(const ScatterArgs a) {
    pdl_enter();
    __shared__ float tile[SC_CH * (PK_MAX_TAPS + 1)];
    const int ts = a.taps | 1;   // odd row stride: conflict-free turn-around
    const int co = blockIdx.y, c0 = blockIdx.x * SC_CH;
    const int nch = min(SC_CH, a.cin - c0);
    const long long split_stride = static_cast<long long>(a.n_out) * a.k_total;
    const float* row = a.partial + static_cast<long long>(co) * a.k_total + a.col0 + c0;
    const int per = nch * a.taps;
    for (int e = threadIdx.x; e < per; e += 256) {
        const int t = e / nch, ci = e - t * nch;
        const float* p = row + t * a.cp + ci;
        float s = 0.f;
        for (int k = 0; k < a.splits; ++k) s += __ldcs(p + k * split_stride);
        tile[ci * ts + t] = s;
    }
    __syncthreads();
    float* orow = a.out + co * a.s_co;
    for (int e = threadIdx.x; e < per; e += 256) {
        const int ci = e / a.taps, t = e - ci * a.taps;
        float* o = orow + (c0 + ci) * a.s_ci + a.tap_off[t];
        const float s = tile[ci * ts + t];
        *o = a.accumulate ? *o + s : s;
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200sd_pack_job_size(void) { return static_cast<int>(sizeof(PackJob)); }

extern "C" int b200sd_pack_weights(const void* jobs_dev, int32_t n_jobs, int32_t n_tiles, void* stream) {
    B200_REQUIRE(jobs_dev && n_jobs > 0 && n_tiles > 0, "pack_weights: bad arguments");
    launch_k(pack_weights_kernel, n_tiles, 256, 0, as_stream(stream), static_cast<const PackJob*>(jobs_dev), n_jobs);
    B200_CHECK_LAUNCH("pack_weights");
    return B200SD_OK;
}

extern "C" int b200sd_wgrad_scatter(const float* partial, int32_t splits, int32_t n_out, int32_t k_total, int32_t col0,
                                    int32_t cin, int32_t cp, int32_t taps, const int32_t* tap_off, float* out,
                                    int64_t s_co, int64_t s_ci, int32_t accumulate, void* stream) {
    B200_REQUIRE(partial && out && tap_off && splits >= 1 && n_out >= 1 && cin >= 1 && taps >= 1 && taps <= PK_MAX_TAPS,
                 "wgrad_scatter: bad arguments");
    B200_REQUIRE(col0 >= 0 && cp >= cin && col0 + taps * cp <= k_total, "wgrad_scatter: column range outside the partials");
    B200_REQUIRE(n_out <= 65535, "wgrad_scatter: too many output rows");
    ScatterArgs a;
    a.partial = partial, a.out = out, a.s_co = s_co, a.s_ci = s_ci;
    a.splits = splits, a.n_out = n_out, a.k_total = k_total, a.col0 = col0, a.cin = cin, a.cp = cp, a.taps = taps;
    a.accumulate = accumulate;
    for (int t = 0; t < PK_MAX_TAPS; ++t) a.tap_off[t] = t < taps ? tap_off[t] : 0;
    launch_k(wgrad_scatter_kernel, dim3((cin + SC_CH - 1) / SC_CH, n_out), 256, 0, as_stream(stream), a);
    B200_CHECK_LAUNCH("wgrad_scatter");
    return B200SD_OK;
}

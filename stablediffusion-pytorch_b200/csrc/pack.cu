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

// tile = 32 packed rows x 32 channels x all taps. TAPS > 0: tap count known at compile time — the four (row, channel)
// positions of a thread and their taps unroll into up to 64 independent loads in flight; the first version walked them
// in two nested runtime loops, a chain of dependent L2 / DRAM round trips that held the whole-model pack at 0.96 ms per
// training step, ~5x its memory time. TAPS == 0: any tap count, the original loops.
template <int TAPS>
__device__ __forceinline__ void pack_tile_load(const PackJob& j, int r0, int c0, __nv_bfloat16 (*tile)[PK_MAX_TAPS * PK_CH + 2]) {
    const int tid = threadIdx.x;
    const bool lanes_c = j.s_c <= j.s_r;   // lanes run along whichever of (row, channel) is closer in the source
    constexpr int NT = TAPS > 0 ? TAPS : 1;
    if (TAPS > 0) {
        float v[4][NT];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = tid + k * 256;
            const int a = i & 31, b = i >> 5;
            const int r = lanes_c ? b : a, c = lanes_c ? a : b;
            const int gr = r0 + r, gc = c0 + c;
            const bool ok = gr < j.rows && gc < j.cin;
            const float* s = j.src + gr * j.s_r + gc * j.s_c;
#pragma unroll
            for (int t = 0; t < NT; ++t) v[k][t] = ok ? __ldg(s + j.tap_off[t]) : 0.f;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = tid + k * 256;
            const int a = i & 31, b = i >> 5;
            const int r = lanes_c ? b : a, c = lanes_c ? a : b;
#pragma unroll
            for (int t = 0; t < NT; ++t) tile[r][t * PK_CH + c] = __float2bfloat16_rn(v[k][t]);
        }
        return;
    }
    for (int i = tid; i < PK_ROWS * PK_CH; i += 256) {
        const int a = i & 31, b = i >> 5;
        const int r = lanes_c ? b : a, c = lanes_c ? a : b;
        const int gr = r0 + r, gc = c0 + c;
        const bool ok = gr < j.rows && gc < j.cin;
        const float* s = j.src + gr * j.s_r + gc * j.s_c;
        for (int t = 0; t < j.taps; ++t)
            tile[r][t * PK_CH + c] = __float2bfloat16_rn(ok ? __ldg(s + j.tap_off[t]) : 0.f);
    }
}

__global__ void __launch_bounds__(256) pack_weights_kernel(const PackJob* __restrict__ jobs, int n_jobs) {
    pdl_enter();
    __shared__ __nv_bfloat16 tile[PK_ROWS][PK_MAX_TAPS * PK_CH + 2];
    // binary search: last job with tile0 <= blockIdx.x
    int lo = 0, hi = n_jobs - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (jobs[mid].tile0 <= static_cast<int>(blockIdx.x)) lo = mid; else hi = mid - 1;
    }
    // The job is staged in SHARED memory: a per-thread copy whose tap_off[] is indexed at run time lives in local
    // memory — 128 B of stack stores per thread, 4.5 GB of L2 write traffic per whole-model pack (ncu), 10x the
    // packed bytes.
    __shared__ PackJob sj;
    {
        const int* src = reinterpret_cast<const int*>(jobs + lo);
        int* dst = reinterpret_cast<int*>(&sj);
        for (int i = threadIdx.x; i < static_cast<int>(sizeof(PackJob) / sizeof(int)); i += 256) dst[i] = src[i];
    }
    __syncthreads();
    const PackJob& j = sj;
    const int local = blockIdx.x - j.tile0;
    const int r0 = (local / j.tiles_c) * PK_ROWS, c0 = (local % j.tiles_c) * PK_CH;
    const int tid = threadIdx.x;
    switch (j.taps) {
        case 1: pack_tile_load<1>(j, r0, c0, tile); break;
        case 4: pack_tile_load<4>(j, r0, c0, tile); break;
        case 9: pack_tile_load<9>(j, r0, c0, tile); break;
        case 16: pack_tile_load<16>(j, r0, c0, tile); break;
        default: pack_tile_load<0>(j, r0, c0, tile); break;
    }
    __syncthreads();
    // 16-byte stores: 8 channels of one (row, tap) per thread (rows are 64-channel aligned in the destination)
    const int items = PK_ROWS * j.taps * (PK_CH / 8);
    const bool vec_ok = (j.dst_ld % 8 == 0) && (j.cp % 8 == 0) && ((reinterpret_cast<uintptr_t>(j.dst) & 15) == 0);
    if (vec_ok) {
        for (int i = tid; i < items; i += 256) {
            const int c8 = i % (PK_CH / 8);
            const int t = (i / (PK_CH / 8)) % j.taps;
            const int r = i / ((PK_CH / 8) * j.taps);
            if (r0 + r < j.rows) {
                const __nv_bfloat16* sp = &tile[r][t * PK_CH + 8 * c8];
                uint4 u;
                u.x = *reinterpret_cast<const unsigned int*>(sp), u.y = *reinterpret_cast<const unsigned int*>(sp + 2);
                u.z = *reinterpret_cast<const unsigned int*>(sp + 4), u.w = *reinterpret_cast<const unsigned int*>(sp + 6);
                *reinterpret_cast<uint4*>(j.dst + (r0 + r) * j.dst_ld + t * j.cp + c0 + 8 * c8) = u;
            }
        }
        return;
    }
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
// Every thread issues ALL its loads before the first add (float4 along the channels, the element loop fully unrolled):
// the first version walked its ~5 elements one after the other, a chain of dependent L2 round trips that made the
// kernel latency-bound at 4-5x its memory time (2.45 ms of a 16.7 ms training step, profiles/r2e_*).
constexpr int SC_CH = 128;
constexpr int SC_V4 = (SC_CH / 4) * PK_MAX_TAPS;          // float4 items per CTA at most
constexpr int SC_IT = (SC_V4 + 255) / 256;                // per thread
constexpr int SC_OUT_IT = (SC_CH * PK_MAX_TAPS + 255) / 256;

// (__grid_constant__: tap_off[] is indexed with a runtime tap — without it nvcc copies the whole argument struct into
// every thread's LOCAL memory to take its address: 128 B of stack stores per thread, 690 MB of L2 write traffic for a
// launch whose output is 7 MB — ncu, profiles/r2f_pack_scatter.md.)
__global__ void __launch_bounds__(256) wgrad_scatter_kernel(const __grid_constant__ ScatterArgs a) {
    pdl_enter();
    __shared__ float tile[SC_CH * (PK_MAX_TAPS + 1)];
    const int ts = a.taps | 1;   // odd row stride: conflict-free turn-around
    const int co = blockIdx.y, c0 = blockIdx.x * SC_CH;
    const int nch = min(SC_CH, a.cin - c0);
    const int nq = (nch + 3) >> 2;                       // float4 groups per tap (the padded K columns exist: cp >= cin)
    const long long split_stride = static_cast<long long>(a.n_out) * a.k_total;
    const float* row = a.partial + static_cast<long long>(co) * a.k_total + a.col0 + c0;
    const int per4 = nq * a.taps;
    float4 acc[SC_IT];
#pragma unroll
    for (int i = 0; i < SC_IT; ++i) {
        acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        const int e = threadIdx.x + i * 256;
        if (e < per4) {
            const int t = e / nq, q = e - t * nq;
            const float4* p = reinterpret_cast<const float4*>(row + t * a.cp + 4 * q);
            if (a.splits <= 4) {                         // the common depths: every split's load in flight at once
                float4 v[4];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    v[k] = k < a.splits ? __ldcs(p + k * (split_stride >> 2)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int k = 0; k < 4; ++k)              // fixed order 0, 1, 2, 3 (adding 0.f for absent splits is exact)
                    acc[i].x += v[k].x, acc[i].y += v[k].y, acc[i].z += v[k].z, acc[i].w += v[k].w;
            } else {
                for (int k = 0; k < a.splits; ++k) {
                    const float4 v = __ldcs(p + k * (split_stride >> 2));
                    acc[i].x += v.x, acc[i].y += v.y, acc[i].z += v.z, acc[i].w += v.w;
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < SC_IT; ++i) {
        const int e = threadIdx.x + i * 256;
        if (e < per4) {
            const int t = e / nq, q = e - t * nq;
            float* d = tile + (4 * q) * ts + t;
            d[0] = acc[i].x, d[ts] = acc[i].y, d[2 * ts] = acc[i].z, d[3 * ts] = acc[i].w;
        }
    }
    __syncthreads();
    float* orow = a.out + co * a.s_co;
    const int per = nch * a.taps;
    float old[SC_OUT_IT];
    if (a.accumulate) {
#pragma unroll
        for (int i = 0; i < SC_OUT_IT; ++i) {
            const int e = threadIdx.x + i * 256;
            old[i] = 0.f;
            if (e < per) {
                const int ci = e / a.taps, t = e - ci * a.taps;
                old[i] = orow[(c0 + ci) * a.s_ci + a.tap_off[t]];
            }
        }
    }
#pragma unroll
    for (int i = 0; i < SC_OUT_IT; ++i) {
        const int e = threadIdx.x + i * 256;
        if (e < per) {
            const int ci = e / a.taps, t = e - ci * a.taps;
            const float sum = tile[ci * ts + t];
            orow[(c0 + ci) * a.s_ci + a.tap_off[t]] = a.accumulate ? old[i] + sum : sum;
        }
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
    B200_REQUIRE(k_total % 4 == 0 && col0 % 4 == 0 && cp % 4 == 0 && (reinterpret_cast<uintptr_t>(partial) & 15) == 0,
                 "wgrad_scatter: the partials are read as float4 (k_total, col0, cp multiples of 4, 16-byte aligned)");
    ScatterArgs a;
    a.partial = partial, a.out = out, a.s_co = s_co, a.s_ci = s_ci;
    a.splits = splits, a.n_out = n_out, a.k_total = k_total, a.col0 = col0, a.cin = cin, a.cp = cp, a.taps = taps;
    a.accumulate = accumulate;
    for (int t = 0; t < PK_MAX_TAPS; ++t) a.tap_off[t] = t < taps ? tap_off[t] : 0;
    launch_k(wgrad_scatter_kernel, dim3((cin + SC_CH - 1) / SC_CH, n_out), 256, 0, as_stream(stream), a);
    B200_CHECK_LAUNCH("wgrad_scatter");
    return B200SD_OK;
}

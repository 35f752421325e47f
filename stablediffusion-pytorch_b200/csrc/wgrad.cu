// Weight-gradient GEMM on tcgen05 / TMEM / TMA (backward of every nn.Conv2d / nn.ConvTranspose2d / nn.Linear leaf of
// the denoiser: the `loss.backward()` of train_ddpm_cond_celebhq_multi_gpu.py:362 through models/blocks.py:45-109).
//
//   dWp[co, k(seg, c)] = sum over output pixels p:  dY[p, co] * X_seg[p shifted by (dy_seg, dx_seg), c0_seg + c]
//
// i.e. the gradient in the SAME packed [n_out, k_total] layout the forward kernel (igemm.cu) reads its weights in; the
// host unpacks it to the parameter's layout. The reduction dimension is the pixel axis, which is the slow axis of both
// NHWC operands, so both MMA operands are **MN-major**: a TMA box (64 channels x 64 pixels) lands as 64 pixel rows of
// 128 B (128B-swizzled) and is consumed by tcgen05.mma through an MN-major shared-memory descriptor — no transposed
// copy of dY or X is ever made. The conv tap is again a TMA coordinate offset with out-of-bounds zero fill.
//
// Work item = (128 output channels) x (BN packed-K columns = BN/64 consecutive 64-column blocks, each with its own
// source/tap) x (one of S contiguous ranges of 64-pixel chunks: split-K, because a weight matrix has far fewer tiles
// than the GPU has SMs). Each work item stores its fp32 accumulator to partial[split][co][col] with plain vector stores;
// wgrad_reduce sums the S partials in fixed order (deterministic, no atomics, no zero-initialisation).
//   warp 0: TMA producer   warp 1: MMA issuer (two TMEM accumulators)   warps 2..5: epilogue (tcgen05.ld -> st.global.v4)
#include "host.h"
#include "ptx.cuh"
#include <cudaTypedefs.h>
#include <mutex>

namespace b200 {

constexpr int WG_THREADS = 192;
constexpr int WG_PIX = 64;  // pixels per K chunk
constexpr int WG_MAX_BLK = 192;  // packed K <= 12288 columns

struct WgBlk {  // one 64-column block of the packed K axis
    short src, dy, dx, c0;
};

struct alignas(64) WgradArgs {
    CUtensorMap x_map[4];
    CUtensorMap dy_map;
    int n_out, k_blocks;       // 64-column blocks in the packed K axis
    int ow, oh, on;
    int bw, bh, bn;            // pixel box, bw*bh*bn == 64
    int chunks_x, chunks_y, chunks;
    int m_blocks, n_groups, splits, items;
    float* partial;            // [splits][n_out][k_blocks*64]
    WgBlk blk[WG_MAX_BLK];     // block table, by value (kernel parameter space)
};

// MN-major operand, 128-byte swizzle: 8 k-rows of 128 B per swizzle atom (SBO = 1024 B between 8-row groups),
// 64-element MN blocks LBO bytes apart (cute/atom/mma_traits_sm100.hpp, make_umma_desc<Major::MN>, LayoutType::B128).
__device__ __forceinline__ uint64_t umma_desc_mn128(uint32_t saddr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((saddr >> 4) & 0x3FFF);
    d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= static_cast<uint64_t>(1024 >> 4) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(2) << 61;
    return d;
}

template <int BN, int STAGES>
struct WgSmem {
    static constexpr int A_BYTES = 2 * WG_PIX * 128;          // two 64-channel boxes of dY
    static constexpr int B_BYTES = (BN / 64) * WG_PIX * 128;  // BN/64 boxes of X
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int BAR_OFF = STAGES * STAGE_BYTES;
    static constexpr int TOTAL = BAR_OFF + (2 * STAGES + 4) * 8 + 16 + 1024;
};

struct WgItem {
    int m_blk, n_grp, c_begin, c_end, split;
};

__device__ __forceinline__ WgItem wg_decode(const WgradArgs& a, int item) {
    WgItem w;
    w.n_grp = item % a.n_groups;
    int r = item / a.n_groups;
    w.m_blk = r % a.m_blocks;
    w.split = r / a.m_blocks;
    w.c_begin = static_cast<int>(static_cast<long long>(a.chunks) * w.split / a.splits);
    w.c_end = static_cast<int>(static_cast<long long>(a.chunks) * (w.split + 1) / a.splits);
    return w;
}

template <int BN, int STAGES>
__global__ void __launch_bounds__(WG_THREADS, 1) wgrad_kernel(const __grid_constant__ WgradArgs args) {
    using L = WgSmem<BN, STAGES>;
    constexpr int NB = BN / 64;
    constexpr uint32_t TMEM_COLS = 2 * BN;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tfull_bar = empty_bar + STAGES;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    pdl_trigger();
    if (warp == 0 && lane == 0) {
        for (int i = 0; i < 4; ++i) tma_prefetch_desc(&args.x_map[i]);
        tma_prefetch_desc(&args.dy_map);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tfull_bar[b], 1);
            mbar_init(&tempty_bar[b], 128);
        }
        fence_barrier_init();
    }
    if (warp == 2) {
        tmem_alloc(tmem_slot, TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();

    if (warp == 0) {
        if (lane == 0) {
            int it = 0;
            for (int item = blockIdx.x; item < args.items; item += gridDim.x) {
                const WgItem w = wg_decode(args, item);
                WgBlk blk[NB];
                int nvalid = 0;
#pragma unroll
                for (int j = 0; j < NB; ++j) {
                    const int b = w.n_grp * NB + j;
                    if (b < args.k_blocks) blk[j] = args.blk[b], nvalid = j + 1;
                }
                for (int c = w.c_begin; c < w.c_end; ++c, ++it) {
                    const int cx = c % args.chunks_x;
                    const int r = c / args.chunks_x;
                    const int cy = r % args.chunks_y, cn = r / args.chunks_y;
                    const int x0 = cx * args.bw, y0 = cy * args.bh, n0 = cn * args.bn;
                    const int stage = it % STAGES;
                    mbar_wait(&empty_bar[stage], ((it / STAGES) & 1) ^ 1);
                    uint8_t* sa = smem + stage * L::STAGE_BYTES;
                    uint8_t* sb = sa + L::A_BYTES;
                    mbar_expect_tx(&full_bar[stage], L::A_BYTES + nvalid * WG_PIX * 128);
                    tma_load_4d(sa, &args.dy_map, &full_bar[stage], w.m_blk * 128, x0, y0, n0);
                    tma_load_4d(sa + WG_PIX * 128, &args.dy_map, &full_bar[stage], w.m_blk * 128 + 64, x0, y0, n0);
#pragma unroll
                    for (int j = 0; j < NB; ++j)
                        if (j < nvalid)
                            tma_load_4d(sb + j * WG_PIX * 128, &args.x_map[blk[j].src], &full_bar[stage], blk[j].c0,
                                        x0 + blk[j].dx, y0 + blk[j].dy, n0);
                }
            }
        }
    } else if (warp == 1) {
        // both operands MN-major: instruction-descriptor bits 15 (A) and 16 (B)
        constexpr uint32_t idesc = umma_idesc_bf16(128, BN) | (1u << 15) | (1u << 16);
        int it = 0, lt = 0;
        for (int item = blockIdx.x; item < args.items; item += gridDim.x, ++lt) {
            const WgItem w = wg_decode(args, item);
            const int buf = lt & 1;
            mbar_wait(&tempty_bar[buf], ((lt >> 1) & 1) ^ 1);
            tc_fence_after();
            const uint32_t tacc = tmem_base + buf * BN;
            for (int c = w.c_begin; c < w.c_end; ++c, ++it) {
                const int stage = it % STAGES;
                mbar_wait(&full_bar[stage], (it / STAGES) & 1);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t sa = smem_u32(smem + stage * L::STAGE_BYTES);
                    const uint64_t da = umma_desc_mn128(sa, WG_PIX * 128);
                    const uint64_t db = umma_desc_mn128(sa + L::A_BYTES, WG_PIX * 128);
#pragma unroll
                    for (int k = 0; k < WG_PIX / 16; ++k)  // 16 pixel rows = 2048 B per K step
                        umma_bf16(tacc, da + (2048 >> 4) * k, db + (2048 >> 4) * k, idesc, (c != w.c_begin) || k != 0);
                    umma_commit(&empty_bar[stage]);
                }
                __syncwarp();
            }
            if (lane == 0) umma_commit(&tfull_bar[buf]);
            __syncwarp();
        }
    } else {
        const int q = warp & 3;
        const int r = q * 32 + lane;
        const long long k_total = static_cast<long long>(args.k_blocks) * 64;
        int lt = 0;
        for (int item = blockIdx.x; item < args.items; item += gridDim.x, ++lt) {
            const WgItem w = wg_decode(args, item);
            const int buf = lt & 1;
            const int co = w.m_blk * 128 + r;
            float* dst = args.partial + (static_cast<long long>(w.split) * args.n_out + co) * k_total +
                         static_cast<long long>(w.n_grp) * BN;
            mbar_wait(&tfull_bar[buf], (lt >> 1) & 1);
            tc_fence_after();
            const bool empty = w.c_begin == w.c_end;  // more splits than chunks: this range holds nothing
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                if (w.n_grp * BN + c * 32 >= k_total) break;
                uint32_t acc[32];
                tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * BN + c * 32, acc);
                tmem_ld_wait();
                if (co < args.n_out) {
#pragma unroll
                    for (int g = 0; g < 8; ++g) {
                        float4 v = make_float4(__uint_as_float(acc[4 * g]), __uint_as_float(acc[4 * g + 1]),
                                               __uint_as_float(acc[4 * g + 2]), __uint_as_float(acc[4 * g + 3]));
                        if (empty) v = make_float4(0.f, 0.f, 0.f, 0.f);
                        *reinterpret_cast<float4*>(dst + c * 32 + 4 * g) = v;
                    }
                }
            }
            tc_fence_before();
            mbar_arrive(&tempty_bar[buf]);
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

__global__ void wgrad_reduce_kernel(const float* __restrict__ partial, int splits, long long elems,
                                    float* __restrict__ out, int accumulate) {
    pdl_enter();
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < (elems >> 2); i += stride) {
        float4 s = reinterpret_cast<const float4*>(partial)[i];
        for (int k = 1; k < splits; ++k) {
            const float4 v = reinterpret_cast<const float4*>(partial + k * elems)[i];
            s.x += v.x, s.y += v.y, s.z += v.z, s.w += v.w;
        }
        if (accumulate) {
            const float4 o = reinterpret_cast<float4*>(out)[i];
            s.x += o.x, s.y += o.y, s.z += o.z, s.w += o.w;
        }
        reinterpret_cast<float4*>(out)[i] = s;
    }
}

static PFN_cuTensorMapEncodeTiled_v12000 wg_encode_fn() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    });
    return fn;
}

static int wg_encode(CUtensorMap* m, const b200sd_src& s, const uint32_t* box, const char* what) {
    auto enc = wg_encode_fn();
    if (!enc) {
        set_error("wgrad: cuTensorMapEncodeTiled driver entry point unavailable");
        return B200SD_ERR_CUDA;
    }
    if (!s.base || (reinterpret_cast<uintptr_t>(s.base) & 15)) {
        set_error("wgrad: %s base unaligned", what);
        return B200SD_ERR_ARG;
    }
    uint64_t dims[4] = {(uint64_t)s.dims[0], (uint64_t)s.dims[1], (uint64_t)s.dims[2], (uint64_t)s.dims[3]};
    uint64_t str[3];
    for (int d = 0; d < 3; ++d) {
        if (s.strides[d] <= 0 || s.strides[d] % 16) {
            set_error("wgrad: %s stride[%d]=%lld B is not a positive multiple of 16", what, d, (long long)s.strides[d]);
            return B200SD_ERR_ARG;
        }
        str[d] = (uint64_t)s.strides[d];
    }
    uint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(s.base), dims, str, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("wgrad: cuTensorMapEncodeTiled(%s) failed (%d)", what, (int)r);
        return B200SD_ERR_CUDA;
    }
    return B200SD_OK;
}

static void wg_box(int ow, int oh, int& bw, int& bh, int& bn) {
    bw = 1;
    while (bw < ow && bw < WG_PIX) bw <<= 1;
    bh = 1;
    while (bw * bh < WG_PIX && bh < oh) bh <<= 1;
    bn = WG_PIX / (bw * bh);
}

template <int BN, int STAGES>
static int launch_wgrad(const WgradArgs& a, int grid, cudaStream_t st) {
    using L = WgSmem<BN, STAGES>;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(wgrad_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             L::TOTAL);
        if (e != cudaSuccess) {
            set_error("wgrad: cannot set %d B dynamic smem: %s", L::TOTAL, cudaGetErrorString(e));
            return B200SD_ERR_CUDA;
        }
        configured = true;
    }
    launch_k(wgrad_kernel<BN, STAGES>, grid, WG_THREADS, L::TOTAL, st, a);
    B200_CHECK_LAUNCH("wgrad");
    return B200SD_OK;
}

}  // namespace b200

using namespace b200;

static int wg_bn(int k_blocks) { return k_blocks >= 4 ? 256 : 128; }

extern "C" int b200sd_wgrad_splits(int n_out, int k_total, int on, int oh, int ow) {
    int bw, bh, bn;
    wg_box(ow, oh, bw, bh, bn);
    const long long chunks = (long long)((ow + bw - 1) / bw) * ((oh + bh - 1) / bh) * ((on + bn - 1) / bn);
    const int k_blocks = (k_total + 63) / 64;
    const int BN = wg_bn(k_blocks);
    const long long base = (long long)((n_out + 127) / 128) * ((k_blocks * 64 + BN - 1) / BN);
    // Split-K depth by a small cost model (microseconds, fitted to the batch-16 UNet shapes on B200): each split adds a
    // full fp32 copy of the gradient to write (here) and to read back (wgrad_scatter / wgrad_reduce), which costs more
    // than the tensor-core time it saves once one wave of CTAs is busy:
    //   rounds(s) * (chunks/s * t_chunk + t_tile_store) + s * tiles * t_tile_load
    const double f = BN / 256.0;
    const double t_chunk = 0.4 * f, t_store = 6.5 * f, t_load = 0.055 * f;
    const int nsm = sm_count();
    long long smax = chunks < 64 ? chunks : 64;
    if (smax < 1) smax = 1;
    long long s = 1;
    double best = 1e30;
    for (long long c = 1; c <= smax; ++c) {
        const double rounds = (double)((base * c + nsm - 1) / nsm);
        const double cost = rounds * ((double)((chunks + c - 1) / c) * t_chunk + t_store) + (double)(c * base) * t_load;
        if (cost < best - 1e-9) best = cost, s = c;
    }
    return (int)s;
}

extern "C" int b200sd_wgrad(const b200sd_wgrad_params* p, void* stream) {
    B200_REQUIRE(p, "wgrad: null params");
    B200_REQUIRE(p->n_src >= 1 && p->n_src <= 4, "wgrad: n_src=%d not in [1,4]", p->n_src);
    B200_REQUIRE(p->n_seg >= 1 && p->n_seg <= 20, "wgrad: n_seg=%d not in [1,20]", p->n_seg);
    B200_REQUIRE(p->partial && p->splits >= 1, "wgrad: null partial buffer / splits < 1");
    B200_REQUIRE(p->n_out >= 1 && p->on >= 1 && p->oh >= 1 && p->ow >= 1, "wgrad: empty problem");
    B200_REQUIRE((reinterpret_cast<uintptr_t>(p->partial) & 15) == 0, "wgrad: partial buffer unaligned");
    WgradArgs a;
    memset(&a, 0, sizeof(a));
    int nb = 0;
    for (int s = 0; s < p->n_seg; ++s) {
        const b200sd_seg& sg = p->segs[s];
        B200_REQUIRE(sg.src >= 0 && sg.src < p->n_src && sg.nchunk >= 1 && sg.c0 >= 0, "wgrad: bad segment %d", s);
        for (int ch = 0; ch < sg.nchunk; ++ch) {
            B200_REQUIRE(nb < WG_MAX_BLK, "wgrad: packed K axis longer than %d columns", WG_MAX_BLK * 64);
            a.blk[nb++] = WgBlk{(short)sg.src, (short)sg.dy, (short)sg.dx, (short)(sg.c0 + ch * 64)};
        }
    }
    a.k_blocks = nb;
    a.n_out = p->n_out;
    a.ow = p->ow, a.oh = p->oh, a.on = p->on;
    wg_box(p->ow, p->oh, a.bw, a.bh, a.bn);
    a.chunks_x = (p->ow + a.bw - 1) / a.bw;
    a.chunks_y = (p->oh + a.bh - 1) / a.bh;
    a.chunks = a.chunks_x * a.chunks_y * ((p->on + a.bn - 1) / a.bn);
    const uint32_t box[4] = {64u, (uint32_t)a.bw, (uint32_t)a.bh, (uint32_t)a.bn};
    for (int i = 0; i < p->n_src; ++i) {
        int rc = wg_encode(&a.x_map[i], p->srcs[i], box, "source");
        if (rc) return rc;
    }
    for (int i = p->n_src; i < 4; ++i) a.x_map[i] = a.x_map[0];
    int rc = wg_encode(&a.dy_map, p->dy, box, "dY");
    if (rc) return rc;
    const int BN = wg_bn(nb);
    a.m_blocks = (p->n_out + 127) / 128;
    a.n_groups = (nb * 64 + BN - 1) / BN;
    a.splits = p->splits;
    const long long items = (long long)a.m_blocks * a.n_groups * a.splits;
    B200_REQUIRE(items < (1ll << 31), "wgrad: too many work items");
    a.items = (int)items;
    a.partial = p->partial;
    const int grid = a.items < sm_count() ? a.items : sm_count();
    cudaStream_t st = as_stream(stream);
    if (BN == 256) return launch_wgrad<256, 4>(a, grid, st);
    return launch_wgrad<128, 6>(a, grid, st);
}

extern "C" int b200sd_wgrad_reduce(const float* partial, int splits, int64_t elems, float* out, int accumulate,
                                   void* stream) {
    B200_REQUIRE(partial && out && splits >= 1 && elems >= 0 && elems % 4 == 0, "wgrad_reduce: bad arguments");
    B200_REQUIRE(((reinterpret_cast<uintptr_t>(partial) | reinterpret_cast<uintptr_t>(out)) & 15) == 0,
                 "wgrad_reduce: pointers unaligned");
    if (elems == 0) return B200SD_OK;
    long long blocks = (elems / 4 + 255) / 256;
    const long long cap = 8ll * sm_count();
    if (blocks > cap) blocks = cap;
    launch_k(wgrad_reduce_kernel, (int)blocks, 256, 0, as_stream(stream), partial, splits, elems, out, accumulate);
    B200_CHECK_LAUNCH("wgrad_reduce");
    return B200SD_OK;
}

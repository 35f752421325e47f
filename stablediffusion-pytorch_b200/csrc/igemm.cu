// Implicit-GEMM core on tcgen05 / TMEM / TMA (SURVEY K1-K4, K6, K9-K13 projections).
//
// Persistent kernel, each CTA walks output tiles of 128 (pixels) x BN (channels), BN in {256, 192, 128, 64, 32}. Three
// residency forms of the same code (template parameters MINB, CL2; b200sd_igemm picks per call):
//   one CTA per SM                     the default
//   two CTAs per SM (<128,2,2>,<64,2,2>)  K <= 768 with more than one wave of tiles: overlaps two epilogue chains
//   CTA pairs, cta_group::2 (CL2)      K >= 1024: two adjacent row tiles of one column tile as ONE 256 x BN MMA, each CTA
//                                      staging half of the weight tile, which buys a 4-5 stage ring (see CL2 below)
// Roles inside a CTA:
//   warp 0      : TMA producer — per 64-channel K chunk one 4-D box load of the activation tile (the conv tap is a
//                 coordinate offset; image borders are TMA out-of-bounds zero fill = conv padding) and one 2-D box
//                 load of the weight tile, both landing 128B-swizzled, K-major in a STAGES-deep smem ring
//   warp 1      : MMA issuer — lane 0 issues 4 x tcgen05.mma (M=128, N=BN, K=16) per chunk into one of TWO TMEM
//                 accumulators, releases ring stages with tcgen05.commit and hands the accumulator to the epilogue
//   warps 2..9  : epilogue — tcgen05.ld the accumulator (one TMEM lane = one output pixel per thread), bias /
//                 time-embedding row bias / activation / gate / residual, bf16 convert into a swizzled smem staging
//                 tile, optional per-channel sum / sum-of-squares of the tile (the next GroupNorm's statistics), then
//                 ONE thread issues TMA tensor stores. The residual tile is TMA-loaded into the same staging buffer
//                 ahead of time. While the epilogue drains accumulator A the MMA warp fills accumulator B.
// The epilogue's global traffic is therefore all bulk-async (TMA): a handful of threads keep tens of KB in flight,
// which per-thread 16-byte loads/stores from 128 threads cannot (r1a profile: 5 % tensor-pipe on K=384 GEMMs).
// Outputs TMA cannot address (fp32, row pitch not a multiple of 16 B, fewer than 64 columns) take a direct path.
#include "host.h"
#include <cstdio>
#include "ptx.cuh"
#include <cudaTypedefs.h>
#include <mutex>
#include <stdlib.h>

namespace b200 {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int IGEMM_THREADS = 320;  // TMA warp, MMA warp, 8 epilogue warps
constexpr int MAX_SEG = 40;
constexpr int EPI_THREADS = 256;
constexpr int PANEL_BYTES = BM * 128;  // 128 rows x 64 bf16

struct SegDev {
    int src, dy, dx, c0, nchunk;
};

struct alignas(64) IgemmArgs {
    CUtensorMap a_map[4];
    CUtensorMap b_map;
    CUtensorMap b_half_map;  // box of BN/2 rows: each CTA of a pair stages one half of the weight tile (CL2 kernels)
    CUtensorMap out_map;
    CUtensorMap res_map;
    SegDev seg[MAX_SEG];
    int n_seg, total_chunks;
    int n_out;
    int on, oh, ow;  // output pixel grid
    int bw, bh, bn;  // tile box (bw*bh*bn == 128)
    int tiles_x, tiles_y, tiles_img, tiles_col, num_tiles;
    int m_tiles, num_ptiles;  // row tiles; (row-tile pair, column tile) work items of the CL2 kernels
    int tma_out;     // 1: staged TMA-store epilogue, 0: direct global stores
    int tma_res;     // 1: residual arrives by TMA into the staging buffer
    void* out;
    int out_f32;
    long long out_sn, out_sy, out_sx;
    const float* bias;
    const __nv_bfloat16* rowbias;
    long long rowbias_ld;
    const __nv_bfloat16* gate;
    long long gate_ld;
    const __nv_bfloat16* residual;
    long long res_sn, res_sy, res_sx;
    int act;
    int res_f32, rowbias_f32;   // precise mode: fp32 residual / row bias (direct-store epilogue only)
    int vec_ok;    // every optional epilogue operand (and a directly stored output) is 16-byte addressable
    float* stats;  // [on][stats_p][stats_ld][2] partial sums, or null
    int stats_p, stats_slot0, stats_ld;
    // split-K (non-pair kernels): every tile is S work items, each reducing a contiguous range of the K chunks; the
    // items store their fp32 accumulators to `sk_ws` [tile][split][128][BN]; the LAST item of a tile to arrive (tile
    // counter) sums the S partials in split order and runs the epilogue. 1 = off.
    int splits;
    float* sk_ws;
    unsigned int* sk_count;
};

struct Tile {
    int col_t, x0, y0, n0, slot;
};

__device__ __forceinline__ Tile decode_tile(const IgemmArgs& a, int tile) {
    // column tiles vary fastest so CTAs that run together share activation tiles in L2
    Tile t;
    t.col_t = tile % a.tiles_col;
    int m = tile / a.tiles_col;
    const int tx = m % a.tiles_x;
    m /= a.tiles_x;
    const int ty = m % a.tiles_y;
    t.x0 = tx * a.bw;
    t.y0 = ty * a.bh;
    t.n0 = (m / a.tiles_y) * a.bn;
    t.slot = ty * a.tiles_x + tx;
    return t;
}

__device__ __forceinline__ float apply_act(float v, int act) {
    if (act == B200SD_ACT_SILU) return silu_mufu(v);
    if (act == B200SD_ACT_RELU) return fmaxf(v, 0.f);
    return v;
}

__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float2 t = __bfloat1622float2(h[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}
__device__ __forceinline__ uint4 pack8(const float (&v)[8]) {
    uint4 pk;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
    for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
    return pk;
}

// 8 consecutive bf16 / fp32 values starting at p[col], masked to nv valid entries (vector path when aligned)
__device__ __forceinline__ void load8_bf16(const __nv_bfloat16* p, int nv, float (&f)[8]) {
    if (nv == 8 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
        unpack8(__ldg(reinterpret_cast<const uint4*>(p)), f);
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = j < nv ? __bfloat162float(p[j]) : 0.f;
    }
}
__device__ __forceinline__ void load8_f32(const float* p, int nv, float (&f)[8]) {
    if (nv == 8 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
        f[0] = a.x, f[1] = a.y, f[2] = a.z, f[3] = a.w, f[4] = b.x, f[5] = b.y, f[6] = b.z, f[7] = b.w;
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = j < nv ? __ldg(p + j) : 0.f;
    }
}

template <int BN, int STAGES, int MINB = 1, bool CL2 = false>
struct IgemmSmem {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = (CL2 ? BN / 2 : BN) * BK * 2;  // a CTA of a pair stages half of the weight tile's rows
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGING_OFF = STAGES * STAGE_BYTES;
    static constexpr int STAGING_ONE = BN >= 64 ? (BN / 64) * PANEL_BYTES : 0;
    // staging tiles: two when shared memory allows (one CTA per SM: BN <= 128; two CTAs per SM: within 96 KB)
    static constexpr int NSTG =
        BN >= 64 && BN <= 128 && (MINB == 1 || STAGES * STAGE_BYTES + 2 * STAGING_ONE <= 96 * 1024) ? 2 : 1;
    static constexpr int STAGING_BYTES = NSTG * STAGING_ONE;
    static constexpr int BIAS_MAX = 2560;                       // floats of bias kept in shared memory
    static constexpr int BIAS_OFF = STAGING_OFF + STAGING_BYTES;
    static constexpr int BAR_OFF = BIAS_OFF + BIAS_MAX * 4;
    static constexpr int NBAR = 2 * STAGES + 6;  // full[S], empty[S], tmem_full[2], tmem_empty[2], res_full[2]
    static constexpr int RED_OFF = BAR_OFF + (NBAR * 8 + 16 + 15) / 16 * 16;  // 8 warps x 32 float4: statistics exchange
    static constexpr int TOTAL = RED_OFF + 4096 + 1024;      // +1024: manual alignment slack
};

// Epilogue arithmetic on 8 consecutive output columns of one row.
__device__ __forceinline__ void epilogue8(const IgemmArgs& args, float (&v)[8], int col, int nv, int pn,
                                          bool bias_from_global) {
    float f[8];
    if (args.bias && bias_from_global) {
        load8_f32(args.bias + col, nv, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] += f[j];
    }
    if (args.rowbias) {
        if (args.rowbias_f32) load8_f32(reinterpret_cast<const float*>(args.rowbias) + pn * args.rowbias_ld + col, nv, f);
        else load8_bf16(args.rowbias + pn * args.rowbias_ld + col, nv, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] += f[j];
    }
    if (args.act != B200SD_ACT_NONE) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = apply_act(v[j], args.act);
    }
    if (args.gate) {
        load8_bf16(args.gate + pn * args.gate_ld + col, nv, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] *= f[j];
    }
}

// MINB = 2: two CTAs per SM (<= 102 registers, <= 113 KB shared memory, 2 x 2 x BN <= 512 TMEM columns). Short-K
// GEMMs are bound by the serial latency of one tile's epilogue (TMEM load -> operands -> staging -> barrier -> TMA store
// -> statistics); a second resident CTA runs its epilogue chain in the gaps of the first.
//
// CL2: CTA pairs (thread-block clusters of two, tcgen05 cta_group::2) that compute two adjacent row tiles of the SAME
// column tile as one 256 x BN MMA. Each CTA stages its own 128 activation rows and HALF of the weight tile's rows; the
// leader (rank 0) issues the MMAs, which read both CTAs' shared memory and write each CTA's 128 accumulator rows into
// its own TMEM. Per K chunk a CTA now fills and reads 128 + BN/2 rows of shared memory instead of 128 + BN: the
// one-CTA 128 x 256 tile was bound by shared-memory bandwidth (TMA fill + MMA operand reads ~ 190 B/clk against 128),
// which is why it stopped at ~70 % of the tensor pipe. Barriers: every load of a stage completes bytes on the LEADER's
// full barrier; the leader's tcgen05.commit multicasts the stage release / accumulator-ready arrivals to both CTAs;
// both CTAs' epilogue threads arrive on the leader's accumulator-empty barrier.
template <int BN, int STAGES, int MINB, bool CL2 = false>
__global__ void __launch_bounds__(IGEMM_THREADS, MINB) igemm_kernel(const __grid_constant__ IgemmArgs args) {
    using L = IgemmSmem<BN, STAGES, MINB, CL2>;
    constexpr uint32_t TMEM_COLS = 2 * BN <= 32 ? 32 : 2 * BN <= 64 ? 64 : 2 * BN <= 128 ? 128 : 2 * BN <= 256 ? 256 : 512;  // two accumulators, rounded up to a power of two
    constexpr int PANELS = BN / 64;
    constexpr int NSTG = L::NSTG;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* staging = smem + L::STAGING_OFF;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tfull_bar = empty_bar + STAGES;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint64_t* res_bar = tempty_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_bar + 2);
    float* sbias = reinterpret_cast<float*>(smem + L::BIAS_OFF);
    float* red = reinterpret_cast<float*>(smem + L::RED_OFF);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // this CTA's work list: tile_of(i), i < my_tiles
    const int crank = CL2 ? static_cast<int>(cluster_ctarank()) : 0;
    const int unit = CL2 ? blockIdx.x >> 1 : blockIdx.x, nunits = CL2 ? gridDim.x >> 1 : gridDim.x;
    const int S = CL2 ? 1 : args.splits;   // split-K depth (work item = tile * S + split: a tile's items run side by side)
    const int nwork = CL2 ? args.num_ptiles : args.num_tiles * S;
    const int my_tiles = unit < nwork ? (nwork - unit + nunits - 1) / nunits : 0;
    auto split_of = [&](int i) -> int { return S == 1 ? 0 : (unit + i * nunits) % S; };
    auto tile_of = [&](int i) -> int {
        const int w = unit + i * nunits;
        if (!CL2) return S == 1 ? w : w / S;
        // (row-tile pair, column tile); a pair's second row tile may lie past the last one: every access of such a
        // tile is out of bounds for TMA (zero fill / clipped) and masked in the epilogue
        return (2 * (w / args.tiles_col) + crank) * args.tiles_col + w % args.tiles_col;
    };

    pdl_trigger();
    if (warp == 0 && lane == 0) {
        for (int i = 0; i < 4; ++i) tma_prefetch_desc(&args.a_map[i]);
        tma_prefetch_desc(CL2 ? &args.b_half_map : &args.b_map);
        if (args.tma_out) tma_prefetch_desc(&args.out_map);
        if (args.tma_res) tma_prefetch_desc(&args.res_map);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tfull_bar[b], 1);
            mbar_init(&tempty_bar[b], CL2 ? 2 * EPI_THREADS : EPI_THREADS);
        }
        mbar_init(&res_bar[0], 1);
        mbar_init(&res_bar[1], 1);
        fence_barrier_init();
    }
    if (warp == 2) {
        if (CL2) {
            tmem_alloc_pair(tmem_slot, TMEM_COLS);
            tmem_relinquish_pair();
        } else {
            tmem_alloc(tmem_slot, TMEM_COLS);
            tmem_relinquish();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (CL2) cluster_sync_all();  // both CTAs' barriers exist before any load or arrive that names the peer's
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();   // everything above overlapped the previous kernel's tail; from here on its results are read

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int it = 0;
            for (int lt = 0; lt < my_tiles; ++lt) {
                const Tile t = decode_tile(args, tile_of(lt));
                const int sp = split_of(lt);
                const int kb = static_cast<int>(static_cast<long long>(args.total_chunks) * sp / S);
                const int ke = static_cast<int>(static_cast<long long>(args.total_chunks) * (sp + 1) / S);
                int kc = 0;
                for (int s = 0; s < args.n_seg; ++s) {
                    const SegDev sg = args.seg[s];
                    for (int ch = 0; ch < sg.nchunk; ++ch, ++kc) {
                        if (kc < kb || kc >= ke) continue;   // another split's chunk
                        const int stage = it % STAGES;
                        const uint32_t phase = (it / STAGES) & 1;
                        mbar_wait(&empty_bar[stage], phase ^ 1);
                        uint8_t* sa = smem + stage * L::STAGE_BYTES;
                        if (CL2) {
                            // both CTAs' bytes complete on the leader's barrier; only the leader's MMA warp waits on it
                            const uint32_t lbar = mapa_rank(&full_bar[stage], 0);
                            if (crank == 0) mbar_expect_tx(&full_bar[stage], 2 * L::STAGE_BYTES);
                            tma_load_4d_pair(sa, &args.a_map[sg.src], lbar, sg.c0 + ch * BK, t.x0 + sg.dx, t.y0 + sg.dy,
                                             t.n0);
                            tma_load_2d_pair(sa + L::A_BYTES, &args.b_half_map, lbar, kc * BK,
                                             t.col_t * BN + crank * (BN / 2));
                        } else {
                            mbar_expect_tx(&full_bar[stage], L::STAGE_BYTES);
                            tma_load_4d(sa, &args.a_map[sg.src], &full_bar[stage], sg.c0 + ch * BK, t.x0 + sg.dx,
                                        t.y0 + sg.dy, t.n0);
                            tma_load_2d(sa + L::A_BYTES, &args.b_map, &full_bar[stage], kc * BK, t.col_t * BN);
                        }
                        ++it;
                    }
                }
            }
        }
    } else if (warp == 1 && (!CL2 || crank == 0)) {
        // ------------------------------------------------------------------ MMA issuer (the pair's leader only)
        constexpr uint32_t idesc = umma_idesc_bf16(CL2 ? 2 * BM : BM, BN);
        int it = 0;
        for (int lt = 0; lt < my_tiles; ++lt) {
            const int buf = lt & 1;
            mbar_wait(&tempty_bar[buf], ((lt >> 1) & 1) ^ 1);  // epilogue has drained this accumulator
            tc_fence_after();
            const uint32_t tacc = tmem_base + buf * BN;
            const int sp = split_of(lt);
            const int kb = static_cast<int>(static_cast<long long>(args.total_chunks) * sp / S);
            const int ke = static_cast<int>(static_cast<long long>(args.total_chunks) * (sp + 1) / S);
            for (int kc = kb; kc < ke; ++kc, ++it) {
                const int stage = it % STAGES;
                mbar_wait(&full_bar[stage], (it / STAGES) & 1);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t sa = smem_u32(smem + stage * L::STAGE_BYTES);
                    const uint64_t da = umma_desc_k128(sa);
                    const uint64_t db = umma_desc_k128(sa + L::A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        // +32 B per K=16 step inside the 128 B swizzle atom: +2 in the (>>4) start-address field
                        if (CL2) umma_bf16_pair(tacc, da + 2 * k, db + 2 * k, idesc, (kc | k) != 0);
                        else umma_bf16(tacc, da + 2 * k, db + 2 * k, idesc, ((kc - kb) | k) != 0);
                    }
                    if (CL2) umma_commit_pair(&empty_bar[stage], 3);
                    else umma_commit(&empty_bar[stage]);
                }
                __syncwarp();
            }
            if (lane == 0) {
                if (CL2) umma_commit_pair(&tfull_bar[buf], 3);
                else umma_commit(&tfull_bar[buf]);
            }
            __syncwarp();
        }
    } else if (warp >= 2) {
        // ------------------------------------------------------------------ epilogue (warps 2..9)
        const int e = warp - 2;  // 0..7
        const int q = warp & 3;  // TMEM lane quarter this warp may read
        const int cg = e >> 2;   // column group: this warp handles 32-column chunks cg, cg+2, ...
        const int r = q * 32 + lane;
        const int et = threadIdx.x - 64;
        const int rpi = args.bw * args.bh;  // tile rows per image
        const int lx = r % args.bw, ly = (r / args.bw) % args.bh, ln = r / rpi;
        const bool bias_sm = args.bias && args.n_out <= L::BIAS_MAX;
        if (bias_sm) {
            for (int i = et; i < args.n_out; i += EPI_THREADS) sbias[i] = __ldg(args.bias + i);
            named_bar_sync(1, EPI_THREADS);
        }
        auto load_residual = [&](const Tile& t, uint8_t* dst, uint64_t* bar) {
            int np = 0;
#pragma unroll
            for (int p = 0; p < PANELS; ++p)
                if (t.col_t * BN + p * 64 < args.n_out) ++np;
            mbar_expect_tx(bar, np * PANEL_BYTES);
#pragma unroll
            for (int p = 0; p < PANELS; ++p)
                if (t.col_t * BN + p * 64 < args.n_out)
                    tma_load_4d(dst + p * PANEL_BYTES, &args.res_map, bar, t.col_t * BN + p * 64, t.x0, t.y0, t.n0);
        };
        if (args.tma_res && et == 0 && my_tiles > 0)
            load_residual(decode_tile(args, tile_of(0)), staging, &res_bar[0]);

        const bool fast = args.vec_ok != 0;
        const bool has_bias = args.bias != nullptr, has_rowbias = args.rowbias != nullptr;
        const bool has_gate = args.gate != nullptr, has_res = args.residual != nullptr;
        const int act = args.act;
        for (int lt = 0; lt < my_tiles; ++lt) {
            const Tile t = decode_tile(args, tile_of(lt));
            const int buf = lt & 1;
            const int sb = NSTG == 2 ? (lt & 1) : 0;
            uint8_t* stg = staging + sb * L::STAGING_ONE;
            const int px = t.x0 + lx, py = t.y0 + ly, pn = t.n0 + ln;
            const bool row_ok = (px < args.ow) && (py < args.oh) && (pn < args.on);
            const int pnc = pn < args.on ? pn : args.on - 1;  // clamped image index for per-image operand reads
            const long long o_off = pn * args.out_sn + py * args.out_sy + px * args.out_sx;
            const long long r_off = pn * args.res_sn + py * args.res_sy + px * args.res_sx;

            mbar_wait(&tfull_bar[buf], (lt >> 1) & 1);
            tc_fence_after();
            // ---- split-K: park this item's accumulator, then either leave the tile to a later item or finish it
            bool discard = false;           // this item is not the tile's last: run the epilogue's motions, write nothing
            const float* sk_tile = nullptr; // last item: the tile's S partials [S][128][BN]
            if (!CL2 && S > 1) {
                const int tile_id = tile_of(lt);
                float* mine = args.sk_ws + (static_cast<long long>(tile_id) * S + split_of(lt)) * (BM * BN) + r * BN;
#pragma unroll 1
                for (int c = cg; c < BN / 32; c += 2) {
                    if (t.col_t * BN + c * 32 >= args.n_out) break;
                    uint32_t acc[32];
                    tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * BN + c * 32, acc);
                    tmem_ld_wait();
                    uint4* dst = reinterpret_cast<uint4*>(mine + c * 32);
#pragma unroll
                    for (int g = 0; g < 8; ++g) dst[g] = make_uint4(acc[4 * g], acc[4 * g + 1], acc[4 * g + 2], acc[4 * g + 3]);
                }
                tc_fence_before();
                mbar_arrive(&tempty_bar[buf]);   // the accumulator is free already: the MMA warp may start tile lt + 2
                __threadfence();
                named_bar_sync(3, EPI_THREADS);
                if (et == 0) {
                    const unsigned int seen = atomicAdd(args.sk_count + tile_id, 1u);
                    const bool last = seen == static_cast<unsigned int>(S - 1);
                    if (last) args.sk_count[tile_id] = 0u;   // ready for the next launch
                    red[0] = last ? 1.f : 0.f;
                }
                named_bar_sync(3, EPI_THREADS);
                discard = red[0] == 0.f;
                named_bar_sync(3, EPI_THREADS);    // red[0] is free again (the statistics path below uses `red`)
                if (!discard) {
                    __threadfence();
                    sk_tile = args.sk_ws + static_cast<long long>(tile_id) * S * (BM * BN) + r * BN;
                }
            }
            const bool row_live = row_ok && !discard;
            if (args.tma_res) mbar_wait(&res_bar[sb], NSTG == 2 ? (lt >> 1) & 1 : lt & 1);
#pragma unroll 1
            for (int c = cg; c < BN / 32; c += 2) {
                const int col0 = t.col_t * BN + c * 32;
                if (col0 >= args.n_out) break;  // warp-uniform
                uint32_t acc[32];
                if (!CL2 && S > 1) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) acc[j] = 0u;
                    if (sk_tile) {   // fixed order: split 0, 1, ..., S-1 (plain loads: written by other CTAs of this launch)
                        float sum[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) sum[j] = 0.f;
                        for (int sp = 0; sp < S; ++sp) {
                            const float4* src = reinterpret_cast<const float4*>(sk_tile + static_cast<long long>(sp) * (BM * BN) + c * 32);
#pragma unroll
                            for (int g = 0; g < 8; ++g) {
                                const float4 f = __ldcg(src + g);
                                sum[4 * g] += f.x, sum[4 * g + 1] += f.y, sum[4 * g + 2] += f.z, sum[4 * g + 3] += f.w;
                            }
                        }
#pragma unroll
                        for (int j = 0; j < 32; ++j) acc[j] = __float_as_uint(sum[j]);
                    }
                } else {
                    tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * BN + c * 32, acc);
                    tmem_ld_wait();
                }
                if (fast && col0 + 32 <= args.n_out) {
                    // full 32-column chunk with 16-byte addressable operands: straight-line code, one uniform branch
                    // per optional operand per chunk (the general path below pays them per 8 columns, with scalar
                    // fall-backs inlined each time — its instruction footprint stalls short-K tiles on fetch)
                    float v[4][8];
#pragma unroll
                    for (int g = 0; g < 4; ++g)
#pragma unroll
                        for (int j = 0; j < 8; ++j) v[g][j] = __uint_as_float(acc[g * 8 + j]);
                    if (has_bias) {
                        const uint32_t bs = smem_u32(sbias + col0);
                        const float4* bp = reinterpret_cast<const float4*>(args.bias + col0);
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            float4 b0, b1;
                            if (bias_sm) {
                                const uint4 u0 = lds128(bs + 32 * g), u1 = lds128(bs + 32 * g + 16);
                                b0 = make_float4(__uint_as_float(u0.x), __uint_as_float(u0.y), __uint_as_float(u0.z), __uint_as_float(u0.w));
                                b1 = make_float4(__uint_as_float(u1.x), __uint_as_float(u1.y), __uint_as_float(u1.z), __uint_as_float(u1.w));
                            } else {
                                b0 = __ldg(bp + 2 * g), b1 = __ldg(bp + 2 * g + 1);
                            }
                            v[g][0] += b0.x, v[g][1] += b0.y, v[g][2] += b0.z, v[g][3] += b0.w;
                            v[g][4] += b1.x, v[g][5] += b1.y, v[g][6] += b1.z, v[g][7] += b1.w;
                        }
                    }
                    if (has_rowbias && args.rowbias_f32) {
                        const float* rp = reinterpret_cast<const float*>(args.rowbias) + pnc * args.rowbias_ld + col0;
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            float f[8];
                            load8_f32(rp + 8 * g, 8, f);
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[g][j] += f[j];
                        }
                    } else if (has_rowbias) {
                        const uint4* rp = reinterpret_cast<const uint4*>(args.rowbias + pnc * args.rowbias_ld + col0);
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            float f[8];
                            unpack8(__ldg(rp + g), f);
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[g][j] += f[j];
                        }
                    }
                    if (act == B200SD_ACT_RELU) {
#pragma unroll
                        for (int g = 0; g < 4; ++g)
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[g][j] = fmaxf(v[g][j], 0.f);
                    } else if (act == B200SD_ACT_SILU) {
#pragma unroll
                        for (int g = 0; g < 4; ++g)
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[g][j] = silu_mufu(v[g][j]);
                    }
                    if (has_gate) {
                        const uint4* gp = reinterpret_cast<const uint4*>(args.gate + pnc * args.gate_ld + col0);
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            float f[8];
                            unpack8(__ldg(gp + g), f);
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[g][j] *= f[j];
                        }
                    }
                    if (has_res && args.res_f32 && row_live) {
                        // plain loads (not __ldg): with residual == out this launch accumulates in place
                        const float4* rp = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(args.residual) + r_off + col0);
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            const float4 a0 = rp[2 * g], a1 = rp[2 * g + 1];
                            v[g][0] += a0.x, v[g][1] += a0.y, v[g][2] += a0.z, v[g][3] += a0.w;
                            v[g][4] += a1.x, v[g][5] += a1.y, v[g][6] += a1.z, v[g][7] += a1.w;
                        }
                    } else if (has_res && !args.tma_res && row_live) {
                        const uint4* rp = reinterpret_cast<const uint4*>(args.residual + r_off + col0);
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            float f[8];
                            unpack8(__ldg(rp + g), f);
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[g][j] += f[j];
                        }
                    }
                    if (BN >= 64 && args.tma_out) {
                        const uint32_t row_s = smem_u32(stg) + r * 128;
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            const int jj = c * 4 + g;  // 16-byte chunk index within the tile row
                            const uint32_t sa = row_s + (jj >> 3) * PANEL_BYTES + (((jj & 7) ^ (r & 7)) << 4);
                            if (args.tma_res) {
                                float f[8];
                                unpack8(lds128(sa), f);
#pragma unroll
                                for (int j = 0; j < 8; ++j) v[g][j] += f[j];
                            }
                            sts128(sa, pack8(v[g]));
                        }
                    } else if (row_live) {
                        if (args.out_f32) {
                            float4* op = reinterpret_cast<float4*>(static_cast<float*>(args.out) + o_off + col0);
#pragma unroll
                            for (int g = 0; g < 4; ++g) {
                                op[2 * g] = make_float4(v[g][0], v[g][1], v[g][2], v[g][3]);
                                op[2 * g + 1] = make_float4(v[g][4], v[g][5], v[g][6], v[g][7]);
                            }
                        } else {
                            uint4* op = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(args.out) + o_off + col0);
#pragma unroll
                            for (int g = 0; g < 4; ++g) op[g] = pack8(v[g]);
                        }
                    }
                    continue;
                }
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const int col = col0 + g * 8;
                    if (col >= args.n_out) break;
                    const int nv = min(8, args.n_out - col);
                    float v[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(acc[g * 8 + j]);
                    if (bias_sm) {
                        if (nv == 8) {
                            const float4 b0 = *reinterpret_cast<const float4*>(sbias + col);
                            const float4 b1 = *reinterpret_cast<const float4*>(sbias + col + 4);
                            v[0] += b0.x, v[1] += b0.y, v[2] += b0.z, v[3] += b0.w;
                            v[4] += b1.x, v[5] += b1.y, v[6] += b1.z, v[7] += b1.w;
                        } else {
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                if (j < nv) v[j] += sbias[col + j];
                        }
                    }
                    epilogue8(args, v, col, nv, pnc, !bias_sm);
                    if (BN >= 64 && args.tma_out) {
                        const int jj = c * 4 + g;  // 16-byte chunk index within the tile row
                        const uint32_t sp = smem_u32(stg) + (jj >> 3) * PANEL_BYTES + r * 128 + (((jj & 7) ^ (r & 7)) << 4);
                        if (args.tma_res) {
                            float f[8];
                            unpack8(lds128(sp), f);
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[j] += f[j];
                        } else if (args.residual && row_live) {
                            float f[8];
                            load8_bf16(args.residual + r_off + col, nv, f);
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[j] += f[j];
                        }
                        sts128(sp, pack8(v));
                    } else if (row_live) {
                        if (args.residual && args.res_f32) {
                            const float* rp = reinterpret_cast<const float*>(args.residual) + r_off + col;
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                if (j < nv) v[j] += rp[j];
                        } else if (args.residual) {
                            float f[8];
                            load8_bf16(args.residual + r_off + col, nv, f);
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[j] += f[j];
                        }
                        if (args.out_f32) {
                            float* op = static_cast<float*>(args.out) + o_off + col;
                            if (nv == 8 && ((reinterpret_cast<uintptr_t>(op) & 15) == 0)) {
                                reinterpret_cast<float4*>(op)[0] = make_float4(v[0], v[1], v[2], v[3]);
                                reinterpret_cast<float4*>(op)[1] = make_float4(v[4], v[5], v[6], v[7]);
                            } else {
#pragma unroll
                                for (int j = 0; j < 8; ++j)
                                    if (j < nv) op[j] = v[j];
                            }
                        } else {
                            __nv_bfloat16* op = static_cast<__nv_bfloat16*>(args.out) + o_off + col;
                            if (nv == 8 && ((reinterpret_cast<uintptr_t>(op) & 15) == 0)) {
                                *reinterpret_cast<uint4*>(op) = pack8(v);
                            } else {
#pragma unroll
                                for (int j = 0; j < 8; ++j)
                                    if (j < nv) op[j] = __float2bfloat16_rn(v[j]);
                            }
                        }
                    }
                }
            }
            tc_fence_before();
            // accumulator free: the MMA warp may start tile lt+2 in it (CTA pair: the leader's barrier counts both CTAs;
            // split-K released it right after parking the partial)
            if (CL2) mbar_arrive_cluster(mapa_rank(&tempty_bar[buf], 0));
            else if (S == 1) mbar_arrive(&tempty_bar[buf]);

            if (BN >= 64 && args.tma_out) {
                fence_proxy_async();  // staging writes (generic proxy) -> visible to the TMA store (async proxy)
                const bool has_next = lt + 1 < my_tiles;
                if (NSTG == 2 && et == 0) {
                    tma_store_wait_read();  // the previous tile's store has finished reading the other buffer
                    if (args.tma_res && has_next)
                        load_residual(decode_tile(args, tile_of(lt + 1)), staging + (sb ^ 1) * L::STAGING_ONE,
                                      &res_bar[sb ^ 1]);
                }
                named_bar_sync(1, EPI_THREADS);
                if (et == 0) {
#pragma unroll
                    for (int p = 0; p < PANELS; ++p)
                        if (!discard && t.col_t * BN + p * 64 < args.n_out)
                            tma_store_4d(&args.out_map, stg + p * PANEL_BYTES, t.col_t * BN + p * 64, t.x0, t.y0, t.n0);
                    tma_store_commit();
                }
                if (discard) {
                    // nothing to count: another item owns this tile's output and statistics
                } else if (args.stats && rpi >= 8) {
                    // per-channel sum / sum of squares of the bf16-rounded tile, one partial per (image, tile slot),
                    // deterministic (no atomics). Warp e owns the BN/8 columns [e*W, e*W+W) for ALL 128 rows, so one
                    // image's partial is complete inside the warp: no shared-memory exchange and no barrier (the
                    // row-sliced version below spent 2 named barriers per 64-column panel and cost 8 us per 128x256
                    // tile). Lane = (row lane rl, column pair cp); in every 8-row group a lane reads rows
                    // rl*RG .. rl*RG+RG-1, which XOR the 128-byte swizzle to different 16-byte chunks for different
                    // rl: bank-conflict free 4-byte reads.
                    // (BN = 192: six warps take 32 columns each; BN = 32 never gets here)
                    constexpr int W = BN == 192 ? 32 : BN >= 64 ? BN / 8 : 8, CP = W / 2, RL = 32 / CP, RG = 8 / RL;
                    const int cp = lane % CP, rl = lane / CP;
                    const int colt = e * W + 2 * cp;  // column inside the tile
                    const int col = t.col_t * BN + colt;
                    // explicit ld.shared: through a generic pointer these reads took the L1TEX path (long scoreboard)
                    const uint32_t pb = smem_u32(stg) + (colt >> 6) * PANEL_BYTES + (colt & 7) * 2;
                    const int chunk = (colt & 63) >> 3;
                    const bool full = (t.x0 + args.bw <= args.ow) && (t.y0 + args.bh <= args.oh);
                    // Summation order is the same for every column tile width (so a sample's statistics, and with
                    // them its output, do not depend on the batch it travels in, which picks the tile): per residue
                    // class r = row % 8 the rows 8j + r are added in increasing j; the eight class sums are then
                    // combined as the balanced tree ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)) — in-lane for the RG classes a
                    // lane owns, by xor-shuffles across the row lanes for the rest.
                    const int seg_j = rpi >> 3;  // 8-row groups per image
                    for (int im = 0; im < args.bn && e * W < BN; ++im) {
                        float a0[RG], a1[RG], b0[RG], b1[RG];
#pragma unroll
                        for (int g = 0; g < RG; ++g) a0[g] = a1[g] = b0[g] = b1[g] = 0.f;
                        // branch-free body (staging rows outside the output are readable, just not counted):
                        // a per-lane `if` here cost a BSSY/BSYNC pair per row and made this pass 5 us per tile
                        if (full) {
#pragma unroll 4
                            for (int j = im * seg_j; j < (im + 1) * seg_j; ++j) {
#pragma unroll
                                for (int g = 0; g < RG; ++g) {
                                    const int row = rl * RG + g + 8 * j;
                                    const uint32_t u = lds32(pb + row * 128 + ((chunk ^ (row & 7)) << 4));
                                    const float fx = __uint_as_float(u << 16), fy = __uint_as_float(u & 0xffff0000u);
                                    a0[g] += fx, a1[g] += fy, b0[g] = fmaf(fx, fx, b0[g]), b1[g] = fmaf(fy, fy, b1[g]);
                                }
                            }
                        } else {
#pragma unroll 2
                            for (int j = im * seg_j; j < (im + 1) * seg_j; ++j) {
#pragma unroll
                                for (int g = 0; g < RG; ++g) {
                                    const int row = rl * RG + g + 8 * j;
                                    const bool ok = (t.x0 + row % args.bw < args.ow) && (t.y0 + (row / args.bw) % args.bh < args.oh);
                                    const uint32_t u = lds32(pb + row * 128 + ((chunk ^ (row & 7)) << 4));
                                    const float fx = ok ? __uint_as_float(u << 16) : 0.f;
                                    const float fy = ok ? __uint_as_float(u & 0xffff0000u) : 0.f;
                                    a0[g] += fx, a1[g] += fy, b0[g] = fmaf(fx, fx, b0[g]), b1[g] = fmaf(fy, fy, b1[g]);
                                }
                            }
                        }
#pragma unroll
                        for (int w = 1; w < RG; w <<= 1) {   // in-lane levels of the tree
#pragma unroll
                            for (int g = 0; g < RG; g += 2 * w)
                                a0[g] += a0[g + w], a1[g] += a1[g + w], b0[g] += b0[g + w], b1[g] += b1[g + w];
                        }
                        float s0 = a0[0], s1 = a1[0], q0 = b0[0], q1 = b1[0];
#pragma unroll
                        for (int off = CP; off < 32; off <<= 1) {   // remaining levels across the row lanes
                            s0 += __shfl_xor_sync(0xffffffffu, s0, off);
                            s1 += __shfl_xor_sync(0xffffffffu, s1, off);
                            q0 += __shfl_xor_sync(0xffffffffu, q0, off);
                            q1 += __shfl_xor_sync(0xffffffffu, q1, off);
                        }
                        const int img = t.n0 + im;
                        if (rl == 0 && col < args.n_out && img < args.on) {
                            float* dst = args.stats +
                                         ((static_cast<long long>(img) * args.stats_p + args.stats_slot0 + t.slot) *
                                              args.stats_ld + col) * 2;
                            *reinterpret_cast<float2*>(dst) = make_float2(s0, q0);
                            if (col + 1 < args.n_out) *reinterpret_cast<float2*>(dst + 2) = make_float2(s1, q1);
                        }
                    }
                } else if (args.stats) {
                    // images smaller than 8 pixels: row-sliced version. Warp e owns rows [16e, 16e+16), lane l owns
                    // columns 2l, 2l+1 of a 64-column panel: one conflict-free 4-byte smem read per row.
                    const bool full = (t.x0 + args.bw <= args.ow) && (t.y0 + args.bh <= args.oh);
                    const int rbase = e * 16;
#pragma unroll 1
                    for (int p = 0; p < PANELS; ++p) {
                        const int col = t.col_t * BN + p * 64 + 2 * lane;
                        const uint8_t* pb = stg + p * PANEL_BYTES + (lane & 3) * 4;
                        const int chunk = lane >> 2;
                        auto put = [&](int img_l, float s0, float s1, float q0, float q1) {
                            const int img = t.n0 + img_l;
                            if (col < args.n_out && img < args.on) {
                                float* dst = args.stats +
                                             ((static_cast<long long>(img) * args.stats_p + args.stats_slot0 + t.slot) *
                                                  args.stats_ld + col) * 2;
                                *reinterpret_cast<float2*>(dst) = make_float2(s0, q0);
                                if (col + 1 < args.n_out) *reinterpret_cast<float2*>(dst + 2) = make_float2(s1, q1);
                            }
                        };
                        const int seg = rpi < 16 ? rpi : 16;  // rows of one image inside this warp's 16 rows
                        float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll 4
                        for (int rr = 0; rr < 16; ++rr) {
                            const int row = rbase + rr;
                            bool ok = true;
                            if (!full) ok = (t.x0 + row % args.bw < args.ow) && (t.y0 + (row / args.bw) % args.bh < args.oh);
                            if (ok) {
                                const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(
                                    pb + row * 128 + ((chunk ^ (row & 7)) << 4)));
                                s0 += f.x, s1 += f.y, q0 += f.x * f.x, q1 += f.y * f.y;
                            }
                            if (seg < 16 && ((rr + 1) & (seg - 1)) == 0) {  // image boundary inside the warp's rows
                                put(row / rpi, s0, s1, q0, q1);
                                s0 = s1 = q0 = q1 = 0.f;
                            }
                        }
                        if (rpi >= 16) {
                            if (rpi == 16) {
                                put(e, s0, s1, q0, q1);
                            } else {  // 2, 4 or 8 warps share an image: combine through smem
                                float4* rd = reinterpret_cast<float4*>(red);
                                rd[e * 32 + lane] = make_float4(s0, s1, q0, q1);
                                named_bar_sync(2, EPI_THREADS);
                                const int wpi = rpi >> 4;  // warps per image
                                if ((e & (wpi - 1)) == 0) {
                                    for (int k = 1; k < wpi; ++k) {
                                        const float4 o = rd[(e + k) * 32 + lane];
                                        s0 += o.x, s1 += o.y, q0 += o.z, q1 += o.w;
                                    }
                                    put(rbase / rpi, s0, s1, q0, q1);
                                }
                                named_bar_sync(2, EPI_THREADS);
                            }
                        }
                    }
                }
                if (NSTG == 1) {
                    if (et == 0) tma_store_wait_read();
                    named_bar_sync(1, EPI_THREADS);  // staging may be overwritten
                    if (args.tma_res && et == 0 && has_next)
                        load_residual(decode_tile(args, tile_of(lt + 1)), staging, &res_bar[0]);
                } else if (args.stats) {
                    named_bar_sync(1, EPI_THREADS);  // all statistics reads done before this buffer is refilled
                }
            }
        }
        if (et == 0) tma_store_wait_read();  // smem must outlive the last store's reads
        tc_fence_before();
    }
    __syncthreads();
    if (CL2) cluster_sync_all();  // my shared memory, barriers and TMEM stay valid until the peer has finished too
    if (warp == 2) {
        tc_fence_after();
        if (CL2) tmem_dealloc_pair(tmem_base, TMEM_COLS);
        else tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------------ host side
static PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
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

static int encode_map(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_b,
                      const uint32_t* box) {
    auto enc = get_encode();
    if (!enc) {
        set_error("igemm: cuTensorMapEncodeTiled driver entry point unavailable (no CUDA driver?)");
        return B200SD_ERR_CUDA;
    }
    uint32_t estr[5] = {1, 1, 1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), dims, strides_b, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("igemm: cuTensorMapEncodeTiled failed (%d): base=%p rank=%d dims=[%llu,%llu,%llu,%llu] "
                  "strides=[%llu,%llu,%llu] box=[%u,%u,%u,%u]",
                  (int)r, base, rank, (unsigned long long)dims[0], (unsigned long long)dims[1],
                  (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0),
                  (unsigned long long)strides_b[0], (unsigned long long)(rank > 2 ? strides_b[1] : 0),
                  (unsigned long long)(rank > 3 ? strides_b[2] : 0), box[0], box[1], rank > 2 ? box[2] : 0,
                  rank > 3 ? box[3] : 0);
        return B200SD_ERR_CUDA;
    }
    return B200SD_OK;
}

// bf16, 128-byte swizzle, zero fill out of bounds: shared with csrc/attention_tc.cu
int encode_tma_map(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_b,
                   const uint32_t* box) {
    return encode_map(m, base, rank, dims, strides_b, box);
}

// (columns, ow, oh, on) view of a pixel-strided bf16 tensor for the epilogue's TMA store / residual load
static bool pixel_view_ok(const void* base, long long sn, long long sy, long long sx) {
    if (reinterpret_cast<uintptr_t>(base) & 15) return false;
    for (long long s : {sn, sy, sx})
        if (s < 0 || (s * 2) % 16 != 0) return false;
    return true;
}
static int encode_pixel_view(CUtensorMap* m, const void* base, int cols, int ow, int oh, int on, long long sn,
                             long long sy, long long sx, const uint32_t* box) {
    auto fix = [](long long s) { return (uint64_t)(s > 0 ? s * 2 : 16); };  // extent-1 dims may carry stride 0
    uint64_t dims[4] = {(uint64_t)cols, (uint64_t)ow, (uint64_t)oh, (uint64_t)on};
    uint64_t str[3] = {fix(sx), fix(sy), fix(sn)};
    return encode_map(m, base, 4, dims, str, box);
}

template <int BN, int STAGES, int MINB = 1, bool CL2 = false>
static int launch_igemm(const IgemmArgs& a, int grid, cudaStream_t st) {
    using L = IgemmSmem<BN, STAGES, MINB, CL2>;
    static_assert(L::TOTAL <= 227 * 1024, "shared memory budget");
    static_assert(MINB == 1 || L::TOTAL <= 113 * 1024, "two CTAs per SM need <= 113 KB each");
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(igemm_kernel<BN, STAGES, MINB, CL2>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL);
        if (e != cudaSuccess) {
            set_error("igemm: cannot set %d B dynamic smem: %s", L::TOTAL, cudaGetErrorString(e));
            return B200SD_ERR_CUDA;
        }
        configured = true;
    }
    if (CL2)
        launch_k_cluster(igemm_kernel<BN, STAGES, MINB, CL2>, grid, IGEMM_THREADS, L::TOTAL, st, 2, a);
    else
        launch_k(igemm_kernel<BN, STAGES, MINB, CL2>, grid, IGEMM_THREADS, L::TOTAL, st, a);
    B200_CHECK_LAUNCH("igemm");
    return B200SD_OK;
}

}  // namespace b200

using namespace b200;

static int g_last_splits = 1;   // split-K depth the last non-pair launch of this process used (tests, tools)

extern "C" int b200sd_igemm(const b200sd_igemm_params* p, void* stream) {
    B200_REQUIRE(p, "igemm: null params");
    B200_REQUIRE(p->n_src >= 1 && p->n_src <= 4, "igemm: n_src=%d not in [1,4]", p->n_src);
    B200_REQUIRE(p->n_seg >= 1 && p->n_seg <= MAX_SEG, "igemm: n_seg=%d not in [1,%d]", p->n_seg, MAX_SEG);
    B200_REQUIRE(p->weight && p->out, "igemm: null weight/out");
    B200_REQUIRE(p->n_out >= 1, "igemm: n_out=%d", p->n_out);
    B200_REQUIRE(p->on >= 1 && p->oh >= 1 && p->ow >= 1, "igemm: empty output grid %dx%dx%d", p->on, p->oh, p->ow);
    B200_REQUIRE(p->out_dtype == B200SD_DT_BF16 || p->out_dtype == B200SD_DT_F32, "igemm: bad out dtype");

    IgemmArgs a;
    memset(&a, 0, sizeof(a));
    // tile box over the output pixel grid: 128 pixels = bw x bh x bn
    int bw = p->ow >= BM ? BM : 1;
    while (bw < p->ow && bw < BM) bw <<= 1;  // next power of two >= ow, capped at 128
    int bh = 1;
    while (bw * bh < BM && bh < p->oh) bh <<= 1;
    int bn = BM / (bw * bh);
    a.bw = bw, a.bh = bh, a.bn = bn;
    a.tiles_x = (p->ow + bw - 1) / bw;
    a.tiles_y = (p->oh + bh - 1) / bh;
    a.tiles_img = (p->on + bn - 1) / bn;

    int total = 0;
    for (int s = 0; s < p->n_seg; ++s) {
        const b200sd_seg& sg = p->segs[s];
        B200_REQUIRE(sg.src >= 0 && sg.src < p->n_src, "igemm: segment %d names source %d", s, sg.src);
        B200_REQUIRE(sg.nchunk >= 1 && sg.c0 >= 0, "igemm: segment %d has nchunk=%d c0=%d", s, sg.nchunk, sg.c0);
        a.seg[s] = SegDev{sg.src, sg.dy, sg.dx, sg.c0, sg.nchunk};
        total += sg.nchunk;
    }
    a.n_seg = p->n_seg;
    a.total_chunks = total;

    const uint32_t abox[4] = {(uint32_t)BK, (uint32_t)bw, (uint32_t)bh, (uint32_t)bn};
    for (int i = 0; i < p->n_src; ++i) {
        const b200sd_src& s = p->srcs[i];
        B200_REQUIRE(s.base && (reinterpret_cast<uintptr_t>(s.base) & 15) == 0, "igemm: source %d base unaligned", i);
        for (int d = 0; d < 3; ++d)
            B200_REQUIRE(s.strides[d] > 0 && s.strides[d] % 16 == 0,
                         "igemm: source %d stride[%d]=%lld B is not a positive multiple of 16", i, d,
                         (long long)s.strides[d]);
        uint64_t dims[4] = {(uint64_t)s.dims[0], (uint64_t)s.dims[1], (uint64_t)s.dims[2], (uint64_t)s.dims[3]};
        uint64_t str[3] = {(uint64_t)s.strides[0], (uint64_t)s.strides[1], (uint64_t)s.strides[2]};
        int rc = encode_map(&a.a_map[i], s.base, 4, dims, str, abox);
        if (rc) return rc;
    }
    for (int i = p->n_src; i < 4; ++i) a.a_map[i] = a.a_map[0];

    // column tile: the persistent grid runs ceil(tiles / SMs) waves of tiles
    const int n = p->n_out;
    const long long m_tiles = (long long)a.tiles_x * a.tiles_y * a.tiles_img;
    const int sms = sm_count();
    int BNsel = 32;
    if (n > 32) {
        long long best = -1;
        for (int bnc : {256, 192, 128, 64}) {
            if (bnc > 64 && n <= bnc / 2) continue;
            // 192 only for long reductions whose width it divides (N = 384, 768 convs: measured +4..5 % over 256, whose
            // last tile is half empty); short-K GEMMs prefer the two-CTA 128 tile, wide ones the 256 tile
            if (bnc == 192 && (n % 192 != 0 || n > 768 || total < 24)) continue;
            const long long tiles = m_tiles * ((n + bnc - 1) / bnc);
            // per-tile time ~ max(mainloop, epilogue) + fixed: the mainloop is operand-feed bound (128 + BN rows of 128
            // bytes per K chunk), the epilogue is proportional to BN; constants fitted to tools/gemm_sweep.py on B200
            const long long main_c = (long long)total * (128 + bnc), epi_c = 10ll * bnc;
            long long cost = ((tiles + sms - 1) / sms) * ((main_c > epi_c ? main_c : epi_c) + 120);
            if (best < 0 || cost < best) best = cost, BNsel = bnc;
        }
    }
    if (const char* f = getenv("B200SD_FORCE_BN")) {  // tuning aid (tools/kernel_bench.py); not used by the product path
        const int v = atoi(f);
        if ((v == 256 || v == 192 || v == 128 || v == 64) && n > 32) BNsel = v;
    }
    // split-K depth (opt-in: the caller passes splitk_max > 1): only where at most half the SMs would have a tile; at
    // most 4 ranges of at least 16 K chunks each. Measured on B200 (tools/splitk_sweep.py, profiles/r2i_splitk.md): it
    // pays only for the 4x4-resolution layers of a batch-16 training step (16-24 tiles: -11 % .. -20 % at depth 3-4),
    // is neutral to harmful from 48 tiles up, and a deeper split loses to the partials' round trip.
    a.splits = 1;
    a.sk_ws = p->splitk_ws;
    a.sk_count = p->splitk_counters;
    {
        const long long tl = m_tiles * ((n + BNsel - 1) / BNsel);
        if (p->splitk_max > 1 && p->splitk_ws && p->splitk_counters && tl * 2 <= sms && total >= 32 &&
            tl <= p->splitk_counters_len && (reinterpret_cast<uintptr_t>(p->splitk_ws) & 15) == 0) {
            long long sd = sms / tl;
            if (sd > p->splitk_max) sd = p->splitk_max;
            if (sd > 4) sd = 4;
            if (sd > total / 16) sd = total / 16;
            if (sd >= 2 && tl * sd * BM * BNsel <= p->splitk_ws_floats) a.splits = (int)sd;
        }
    }
    if (const char* f = getenv("B200SD_FORCE_SPLITS")) {   // tuning aid (tools/splitk_sweep.py), with B200SD_FORCE_BN
        const long long v = atoi(f);
        const long long tl = m_tiles * ((n + BNsel - 1) / BNsel);
        if (p->splitk_ws && p->splitk_counters && v >= 1 && v <= total && tl <= p->splitk_counters_len &&
            tl * v * BM * BNsel <= p->splitk_ws_floats)
            a.splits = (int)v;
    }
    a.tiles_col = (n + BNsel - 1) / BNsel;
    const long long tiles_ll = m_tiles * a.tiles_col;
    B200_REQUIRE(tiles_ll < (1ll << 31), "igemm: too many tiles");
    a.num_tiles = (int)tiles_ll;
    a.m_tiles = (int)m_tiles;
    a.num_ptiles = (int)((m_tiles + 1) / 2) * a.tiles_col;

    {
        B200_REQUIRE((reinterpret_cast<uintptr_t>(p->weight) & 15) == 0, "igemm: weight unaligned");
        uint64_t dims[2] = {(uint64_t)total * BK, (uint64_t)n};
        uint64_t str[1] = {(uint64_t)total * BK * 2};
        uint32_t box[2] = {(uint32_t)BK, (uint32_t)BNsel};
        int rc = encode_map(&a.b_map, p->weight, 2, dims, str, box);
        if (rc) return rc;
        uint32_t hbox[2] = {(uint32_t)BK, (uint32_t)(BNsel / 2)};
        rc = encode_map(&a.b_half_map, p->weight, 2, dims, str, hbox);
        if (rc) return rc;
    }
    a.n_out = n;
    a.on = p->on, a.oh = p->oh, a.ow = p->ow;
    a.out = p->out;
    a.out_f32 = p->out_dtype == B200SD_DT_F32;
    a.out_sn = p->out_sn, a.out_sy = p->out_sy, a.out_sx = p->out_sx;
    a.bias = p->bias;
    a.rowbias = static_cast<const __nv_bfloat16*>(p->rowbias);
    a.rowbias_ld = p->rowbias_ld;
    a.gate = static_cast<const __nv_bfloat16*>(p->gate);
    a.gate_ld = p->gate_ld;
    a.residual = static_cast<const __nv_bfloat16*>(p->residual);
    a.res_sn = p->res_sn, a.res_sy = p->res_sy, a.res_sx = p->res_sx;
    a.act = p->act;
    a.res_f32 = p->residual && p->residual_dtype == B200SD_DT_F32;
    a.rowbias_f32 = p->rowbias && p->rowbias_dtype == B200SD_DT_F32;
    B200_REQUIRE(!a.res_f32 || a.out_f32, "igemm: an fp32 residual needs an fp32 output");

    {
        auto al = [](const void* ptr, long long bytes_a, long long bytes_b = 16, long long bytes_c = 16) {
            return (reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && bytes_a % 16 == 0 && bytes_b % 16 == 0 && bytes_c % 16 == 0;
        };
        const long long ob = a.out_f32 ? 4 : 2;
        const long long rb = a.res_f32 ? 4 : 2, wb = a.rowbias_f32 ? 4 : 2;
        a.vec_ok = (!p->bias || al(p->bias, 16)) && (!p->rowbias || al(p->rowbias, p->rowbias_ld * wb)) &&
                   (!p->gate || al(p->gate, p->gate_ld * 2)) &&
                   (!p->residual || al(p->residual, p->res_sn * rb, p->res_sy * rb, p->res_sx * rb)) &&
                   al(p->out, p->out_sn * ob, p->out_sy * ob, p->out_sx * ob);
    }
    // epilogue route: TMA tensor store from a staged bf16 tile when the output is addressable that way
    a.tma_out = BNsel >= 64 && !a.out_f32 && pixel_view_ok(p->out, p->out_sn, p->out_sy, p->out_sx);
    if (a.tma_out) {
        int rc = encode_pixel_view(&a.out_map, p->out, n, p->ow, p->oh, p->on, p->out_sn, p->out_sy, p->out_sx, abox);
        if (rc) return rc;
        // a zero stride over an extent > 1 is a broadcast (position embedding shared by all images): TMA cannot
        // express it, the epilogue's direct loads can
        const bool res_bcast = (p->res_sn == 0 && p->on > 1) || (p->res_sy == 0 && p->oh > 1) ||
                               (p->res_sx == 0 && p->ow > 1);
        a.tma_res = p->residual && !res_bcast && pixel_view_ok(p->residual, p->res_sn, p->res_sy, p->res_sx);
        if (a.tma_res) {
            rc = encode_pixel_view(&a.res_map, p->residual, n, p->ow, p->oh, p->on, p->res_sn, p->res_sy, p->res_sx,
                                   abox);
            if (rc) return rc;
        }
    }
    if (p->stats) {
        if (!a.tma_out) {
            set_error("igemm: fused statistics need the staged bf16 epilogue (bf16 output, 16-byte aligned pitch, "
                      "more than 32 output channels)");
            return B200SD_ERR_UNSUPPORTED;
        }
        const int per_img = bn > 1 ? 1 : a.tiles_x * a.tiles_y;
        B200_REQUIRE(p->stats_slot0 >= 0 && p->stats_slot0 + per_img <= p->stats_p && p->stats_ld >= n,
                     "igemm: statistics buffer has %d slots (need %d from slot %d) and pitch %d (need %d)", p->stats_p,
                     per_img, p->stats_slot0, p->stats_ld, n);
        a.stats = p->stats, a.stats_p = p->stats_p, a.stats_slot0 = p->stats_slot0, a.stats_ld = p->stats_ld;
    }

    cudaStream_t st = as_stream(stream);
    // short reductions with more than one wave of tiles: two co-resident CTAs per SM overlap their epilogue chains
    static const int dual_max = [] {
        const char* f = getenv("B200SD_DUAL_MAXCHUNKS");
        return f ? atoi(f) : 12;
    }();
    if (total <= dual_max && a.num_tiles > sms && (BNsel == 128 || BNsel == 64)) {
        const int grid2 = a.num_tiles < 2 * sms ? a.num_tiles : 2 * sms;
        return BNsel == 128 ? launch_igemm<128, 2, 2>(a, grid2, st) : launch_igemm<64, 2, 2>(a, grid2, st);
    }
    // long reductions: CTA pairs (cta_group::2), each CTA staging half of the weight tile
    static const int cl2_min = [] {
        const char* f = getenv("B200SD_CL2_MINCHUNKS");
        return f ? atoi(f) : 16;
    }();
    if (a.splits == 1 && total >= cl2_min && BNsel >= 128 && m_tiles >= 2) {
        const int ncl = a.num_ptiles < sms / 2 ? a.num_ptiles : sms / 2;
        switch (BNsel) {
            case 256: return launch_igemm<256, 4, 1, true>(a, 2 * ncl, st);
            case 192: return launch_igemm<192, 5, 1, true>(a, 2 * ncl, st);
            default: return launch_igemm<128, 5, 1, true>(a, 2 * ncl, st);
        }
    }
    g_last_splits = a.splits;
    static const bool dbg = getenv("B200SD_DEBUG_IGEMM") != nullptr;   // one line per launch: shape, tile choice, split depth
    if (dbg)
        fprintf(stderr, "igemm M=%lld N=%d chunks=%d BN=%d tiles=%d splits=%d splitk_max=%d\n",
                (long long)p->on * p->oh * p->ow, n, total, BNsel, a.num_tiles, a.splits, p->splitk_max);
    const long long items = (long long)a.num_tiles * a.splits;
    const int grid = items < sms ? (int)items : sms;
    switch (BNsel) {
        case 256: return launch_igemm<256, 3>(a, grid, st);
        case 192: return launch_igemm<192, 3>(a, grid, st);
        case 128: return launch_igemm<128, 4>(a, grid, st);
        case 64: return launch_igemm<64, 6>(a, grid, st);
        default: return launch_igemm<32, 6>(a, grid, st);
    }
}

extern "C" int b200sd_igemm_last_splits(void) { return g_last_splits; }

/* Number of per-image partial-statistics slots one b200sd_igemm call with this output grid writes. */
extern "C" int b200sd_igemm_stats_slots(int oh, int ow) {
    int bw = ow >= BM ? BM : 1;
    while (bw < ow && bw < BM) bw <<= 1;
    int bh = 1;
    while (bw * bh < BM && bh < oh) bh <<= 1;
    if (bw * bh < BM) return 1;
    return ((ow + bw - 1) / bw) * ((oh + bh - 1) / bh);
}

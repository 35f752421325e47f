// softmax(Q K^T * scale) V on tcgen05 / TMEM / TMA. One CTA = 128 queries of one
// (batch, head); the keys stream through in tiles of 64.
//
//   warp 0      one elected thread is both TMA producer and MMA issuer (a sequential controller: its waits are short and
//               in program order). Q once, then K_j / V_j, each a 4-D box (64, 1, rows, 1) of the (d, heads, L, batch)
//               view of the packed projection output: columns beyond the head dim and rows beyond L are TMA zero fill,
//               so a head's 8..64 channels land as a 128-byte-swizzled [rows x 64] bf16 panel — exactly the K-major
//               operand tile the implicit GEMM uses — with no padding in HBM.
//               S = Q K_j^T (tcgen05.mma kind::f16, M = 128, N = 64, K = d rounded up to 16) into TMEM columns [0, 64);
//               after the softmax has produced P_j: O_j = P_j V_j (M = 128, N = d rounded up to 16, K = 64; V is read
//               MN-major straight from its [keys x d] panel) into TMEM columns [64, 64 + N).
//   warps 1..4  softmax, one thread per query row (TMEM lane): ONE tcgen05.ld brings the row's 64 scores into registers,
//               online max / sum in the log2 domain, P_j as bf16 into a swizzled shared-memory A-operand panel; the
//               partial product O_j is pulled from TMEM one tile later and folded into the fp32 output row held in
//               registers (o = o * alpha_j + O_j), so TMEM never has to be rescaled.
// 128 TMEM columns and 48 KB of shared memory per CTA: three or four CTAs per SM. While one CTA's softmax warps are busy
// on the MUFU / FMA pipes the others' MMAs run, which is the overlap the mma.sync kernel (attention.cu) could not get —
// there one warp does QK^T, exp and PV back to back (r1k profile: tensor 48 %, MUFU 59 %, nothing saturated).
#include "host.h"
#include "ptx.cuh"
#include <cuda_bf16.h>
#include <string.h>

namespace b200 {

int encode_tma_map(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_b,
                   const uint32_t* box);   // igemm.cu

constexpr int ATC_THREADS = 160;
constexpr int ATC_KT = 64;                 // keys per tile
constexpr int ATC_QTILE = 128 * 128;       // [128 rows x 64 bf16] swizzled panel = 16 KB (Q, P)
constexpr int ATC_KTILE = ATC_KT * 128;    // [64 keys x 64 bf16] = 8 KB (K, V)

struct alignas(64) AttnTcArgs {
    CUtensorMap q_map, k_map, v_map;
    __nv_bfloat16* out;
    long long ld_o;
    float* lse;
    int heads, lq, lk, d, q_tiles, k_tiles;
    float scale_log2;
};

struct AtcSmem {
    static constexpr int Q_OFF = 0;
    // two stages each for K and V: K_{j+2} / V_{j+1} are requested while tile j is in its softmax, so the controller
    // thread never sits on a TMA latency between two MMA issues
    static constexpr int K_OFF = ATC_QTILE;
    static constexpr int V_OFF = K_OFF + 2 * ATC_KTILE;
    static constexpr int P_OFF = V_OFF + 2 * ATC_KTILE;
    static constexpr int BAR_OFF = P_OFF + ATC_QTILE;
    static constexpr int TOTAL = BAR_OFF + 13 * 8 + 16;
};

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

// MN-major operand, 128-byte swizzle (the V panel: rows = keys = the MMA's K dimension, 64 head-dim elements per
// 128-byte row) — same descriptor form as csrc/wgrad.cu.
__device__ __forceinline__ uint64_t atc_desc_mn128(uint32_t saddr) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((saddr >> 4) & 0x3FFF);
    d |= static_cast<uint64_t>((ATC_KTILE >> 4) & 0x3FFF) << 16;
    d |= static_cast<uint64_t>(1024 >> 4) << 32;
    d |= static_cast<uint64_t>(1) << 46;
    d |= static_cast<uint64_t>(2) << 61;
    return d;
}

template <int NOUT, int MINB>
__global__ void __launch_bounds__(ATC_THREADS, MINB) attention_tc_kernel(const __grid_constant__ AttnTcArgs args) {
    using L = AtcSmem;
    extern __shared__ uint8_t atc_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(atc_smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
    uint64_t *q_full = bars, *s_full = bars + 1, *p_full = bars + 2, *pv_full = bars + 3, *s_free = bars + 4,
             *k_full = bars + 5, *k_empty = bars + 7, *v_full = bars + 9, *v_empty = bars + 11;   // [2] each
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qt = blockIdx.x % args.q_tiles;
    const int bh = blockIdx.x / args.q_tiles;
    const int head = bh % args.heads, b = bh / args.heads;
    const int k_tiles = args.k_tiles;

    pdl_trigger();
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&args.q_map);
        tma_prefetch_desc(&args.k_map);
        tma_prefetch_desc(&args.v_map);
        for (int i = 0; i < 13; ++i) mbar_init(&bars[i], (i == 2 || i == 4) ? 128 : 1);   // p_full, s_free: every softmax thread
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 128);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t t_s = tmem_base, t_o = tmem_base + 64;
    pdl_wait();

    if (warp == 0) {
        // ------------------------------------------------------------------ controller: TMA loads + MMA issue
        if (lane == 0) {
            constexpr uint32_t idesc_s = umma_idesc_bf16(128, ATC_KT);
            constexpr uint32_t idesc_pv = umma_idesc_bf16(128, NOUT) | (1u << 16);   // B (V) is MN-major
            constexpr int KS = NOUT / 16;
            const uint64_t dq = umma_desc_k128(smem_u32(smem + L::Q_OFF));
            const uint64_t dk = umma_desc_k128(smem_u32(smem + L::K_OFF));
            const uint64_t dp = umma_desc_k128(smem_u32(smem + L::P_OFF));
            const uint64_t dv = atc_desc_mn128(smem_u32(smem + L::V_OFF));
            constexpr uint64_t STG = ATC_KTILE >> 4;   // descriptor start-address units between ring stages
            auto load_k = [&](int j) {
                mbar_expect_tx(&k_full[j & 1], ATC_KTILE);
                tma_load_4d(smem + L::K_OFF + (j & 1) * ATC_KTILE, &args.k_map, &k_full[j & 1], 0, head, j * ATC_KT, b);
            };
            auto load_v = [&](int j) {
                mbar_expect_tx(&v_full[j & 1], ATC_KTILE);
                tma_load_4d(smem + L::V_OFF + (j & 1) * ATC_KTILE, &args.v_map, &v_full[j & 1], 0, head, j * ATC_KT, b);
            };
            auto issue_s = [&](int j) {   // S_j = Q K_j^T, then K's stage goes back to the loader
                mbar_wait(&k_full[j & 1], (j >> 1) & 1);
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < KS; ++k) umma_bf16(t_s, dq + 2 * k, dk + (j & 1) * STG + 2 * k, idesc_s, k != 0);
                umma_commit(s_full);
                umma_commit(&k_empty[j & 1]);
            };
            mbar_expect_tx(q_full, ATC_QTILE);
            tma_load_4d(smem + L::Q_OFF, &args.q_map, q_full, 0, head, qt * 128, b);
            load_k(0);
            load_v(0);
            if (k_tiles > 1) load_k(1);
            mbar_wait(q_full, 0);
            issue_s(0);
            for (int j = 0; j < k_tiles; ++j) {
                if (j + 1 < k_tiles) {
                    // the softmax threads hold S_j in registers: S_{j+1} is computed while they exponentiate, so they
                    // never wait for the tensor pipe (first version: 30 % of all warp samples sat on s_full)
                    mbar_wait(s_free, j & 1);
                    issue_s(j + 1);
                    if (j >= 1) mbar_wait(&v_empty[(j + 1) & 1], ((j - 1) >> 1) & 1);   // P_{j-1} V_{j-1}: long done
                    load_v(j + 1);
                }
                if (j + 2 < k_tiles) {
                    mbar_wait(&k_empty[j & 1], (j >> 1) & 1);                           // S_j: long done
                    load_k(j + 2);
                }
                mbar_wait(p_full, j & 1);       // P_j is in shared memory and O_{j-1} has been pulled out of TMEM
                mbar_wait(&v_full[j & 1], (j >> 1) & 1);
                tc_fence_after();
                const int valid = min(ATC_KT, args.lk - j * ATC_KT);
                const int steps = (valid + 15) >> 4;
                for (int k = 0; k < steps; ++k)
                    umma_bf16(t_o, dp + 2 * k, dv + (j & 1) * STG + 128 * k, idesc_pv, k != 0);
                umma_commit(pv_full);
                umma_commit(&v_empty[j & 1]);
            }
        }
    } else {
        // ------------------------------------------------------------------ softmax: one thread per query row
        const int q4 = warp & 3;                 // TMEM lane quarter this warp may touch
        const int r = q4 * 32 + lane;
        const uint32_t lane_off = static_cast<uint32_t>(q4 * 32) << 16;
        const float sl2 = args.scale_log2;
        const uint32_t p_row = smem_u32(smem + L::P_OFF) + r * 128;
        const int sw = r & 7;
        float o[NOUT];
#pragma unroll
        for (int i = 0; i < NOUT; ++i) o[i] = 0.f;
        float m = -INFINITY, l = 0.f, alpha_prev = 0.f;
        auto fold = [&](float alpha) {   // o = o * alpha + (partial product in TMEM)
#pragma unroll
            for (int c = 0; c < NOUT / 16; ++c) {
                uint32_t pr[16];
                tmem_ld16(t_o + lane_off + c * 16, pr);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 16; ++i) o[c * 16 + i] = fmaf(o[c * 16 + i], alpha, __uint_as_float(pr[i]));
            }
        };
        for (int j = 0; j < k_tiles; ++j) {
            const int valid = min(ATC_KT, args.lk - j * ATC_KT);
            mbar_wait(s_full, j & 1);
            tc_fence_after();
            uint32_t sa[32], sb[32];
            tmem_ld32(t_s + lane_off, sa);
            tmem_ld32(t_s + lane_off + 32, sb);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(s_free);    // the scores are in registers: the tensor pipe may overwrite S
            if (valid < ATC_KT) {   // keys beyond lk: never the maximum, probability exactly 0 (ex2(-inf))
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    if (i >= valid) sa[i] = 0xff800000u;
                    if (32 + i >= valid) sb[i] = 0xff800000u;
                }
            }
            float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
                m0 = fmaxf(m0, __uint_as_float(sa[i])), m1 = fmaxf(m1, __uint_as_float(sa[i + 1]));
                m2 = fmaxf(m2, __uint_as_float(sb[i])), m3 = fmaxf(m3, __uint_as_float(sb[i + 1]));
            }
            const float m_new = fmaxf(m, fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)) * sl2);
            const float alpha = ex2_approx(m - m_new);
            if (j > 0) {   // P_{j-1} V_{j-1} has completed: its result is in TMEM and the P panel may be overwritten
                mbar_wait(pv_full, (j - 1) & 1);
                tc_fence_after();
                fold(alpha_prev);
            }
            alpha_prev = alpha;
            float r0 = 0.f, r1 = 0.f, r2 = 0.f, r3 = 0.f;
#pragma unroll
            for (int g = 0; g < 8; ++g) {   // 8 keys -> one 16-byte chunk of the row
                const uint32_t* sv = g < 4 ? &sa[g * 8] : &sb[(g - 4) * 8];
                float p[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) p[i] = ex2_approx(fmaf(__uint_as_float(sv[i]), sl2, -m_new));
                uint4 pk;
                pk.x = pack_bf16x2(p[0], p[1]);
                pk.y = pack_bf16x2(p[2], p[3]);
                pk.z = pack_bf16x2(p[4], p[5]);
                pk.w = pack_bf16x2(p[6], p[7]);
                sts128(p_row + ((g ^ sw) << 4), pk);
                r0 += p[0] + p[4], r1 += p[1] + p[5], r2 += p[2] + p[6], r3 += p[3] + p[7];
            }
            l = fmaf(l, alpha, (r0 + r1) + (r2 + r3));
            m = m_new;
            tc_fence_before();      // the TMEM reads above precede P_j V_j
            fence_proxy_async();    // P (generic-proxy stores) -> visible to the MMA's async-proxy reads
            mbar_arrive(p_full);
        }
        mbar_wait(pv_full, (k_tiles - 1) & 1);
        tc_fence_after();
        fold(alpha_prev);
        const int qi = qt * 128 + r;
        if (qi < args.lq) {
            const float inv = 1.f / l;
            __nv_bfloat16* op = args.out + (static_cast<long long>(b) * args.lq + qi) * args.ld_o + head * args.d;
#pragma unroll
            for (int c = 0; c < NOUT / 8; ++c) {
                if (c * 8 < args.d) {
                    uint4 pk;
                    pk.x = pack_bf16x2(o[c * 8 + 0] * inv, o[c * 8 + 1] * inv);
                    pk.y = pack_bf16x2(o[c * 8 + 2] * inv, o[c * 8 + 3] * inv);
                    pk.z = pack_bf16x2(o[c * 8 + 4] * inv, o[c * 8 + 5] * inv);
                    pk.w = pack_bf16x2(o[c * 8 + 6] * inv, o[c * 8 + 7] * inv);
                    *reinterpret_cast<uint4*>(op + c * 8) = pk;
                }
            }
            if (args.lse) args.lse[(static_cast<long long>(b) * args.heads + head) * args.lq + qi] = m + log2f(l);
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 128);
    }
}

static int atc_view(CUtensorMap* m, const void* base, int64_t ld, int d, int heads, int len, int64_t batch, int rows) {
    // (d, heads, L, batch) view of a [batch * L, ld] bf16 matrix whose head h occupies columns [h * d, (h + 1) * d)
    const uint64_t dims[4] = {(uint64_t)d, (uint64_t)heads, (uint64_t)len, (uint64_t)batch};
    const uint64_t str[3] = {(uint64_t)d * 2, (uint64_t)ld * 2, (uint64_t)len * (uint64_t)ld * 2};
    const uint32_t box[4] = {64, 1, (uint32_t)rows, 1};
    return encode_tma_map(m, base, 4, dims, str, box);
}

template <int NOUT, int MINB>
static int atc_launch(const AttnTcArgs& a, long long grid, cudaStream_t st) {
    constexpr int SMEM = AtcSmem::TOTAL + 1024;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(attention_tc_kernel<NOUT, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
        if (e != cudaSuccess) {
            set_error("attention_tc: cannot set %d B dynamic smem: %s", SMEM, cudaGetErrorString(e));
            return B200SD_ERR_CUDA;
        }
        configured = true;
    }
    launch_k(attention_tc_kernel<NOUT, MINB>, dim3((unsigned)grid), ATC_THREADS, SMEM, st, a);
    B200_CHECK_LAUNCH("attention_tc");
    return B200SD_OK;
}

// Returns 1 when the request is not one this kernel takes (the caller then uses the mma.sync kernels of attention.cu).
int attention_tc(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v, int64_t ld_v, void* out,
                 int64_t ld_o, int64_t batch, int heads, int lq, int lk, int d, float scale, float* lse,
                 cudaStream_t st) {
    if (!(scale > 0.f) || d % 8 || d < 8 || d > 64 || ld_o % 8 || (reinterpret_cast<uintptr_t>(out) & 15) ||
        ld_q % 8 || ld_k % 8 || ld_v % 8)
        return 1;
    const long long q_tiles = (lq + 127) / 128, grid = q_tiles * heads * batch;
    if (grid >= (1ll << 31)) return 1;
    AttnTcArgs a;
    memset(&a, 0, sizeof(a));
    int rc = atc_view(&a.q_map, q, ld_q, d, heads, lq, batch, 128);
    if (rc) return rc;
    if ((rc = atc_view(&a.k_map, k, ld_k, d, heads, lk, batch, ATC_KT))) return rc;
    if ((rc = atc_view(&a.v_map, v, ld_v, d, heads, lk, batch, ATC_KT))) return rc;
    a.out = static_cast<__nv_bfloat16*>(out);
    a.ld_o = ld_o;
    a.lse = lse;
    a.heads = heads, a.lq = lq, a.lk = lk, a.d = d;
    a.q_tiles = (int)q_tiles, a.k_tiles = (lk + ATC_KT - 1) / ATC_KT;
    a.scale_log2 = scale * 1.4426950408889634f;
    switch ((d + 15) / 16) {
        case 1: return atc_launch<16, 4>(a, grid, st);
        case 2: return atc_launch<32, 3>(a, grid, st);
        case 3: return atc_launch<48, 2>(a, grid, st);
        default: return atc_launch<64, 2>(a, grid, st);
    }
}

}  // namespace b200

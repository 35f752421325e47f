// Fused attention forward: softmax(Q K^T * scale) V per (batch, head), flash-style online softmax,
// never materialising the L x L score matrix the reference builds (models/attention.py:59-60,
// models/multihead_attention.py:58-61).
//
// Head dims on this path are 8..64 (16 heads over 128..768 channels; VQVAE 4 heads x 64; DiT 9 x 32). At these
// sizes the kernel is bound by the exp (MUFU) and fp32 softmax arithmetic, not by the MMA rate: QK^T costs 2*d
// flops per score against one exp per score, so the tensor pipe is <10 % busy whatever instruction feeds it.
// The MMAs are therefore warp-level mma.sync.m16n8k16 (bf16 in, fp32 accumulate) straight from registers, which
// keeps P in registers between the two GEMMs; tcgen05 would add a TMEM round trip per tile for no gain here.
//
// Two kernels: `attention_resident_kernel` (everything in the UNet / DiT: K and V of a (batch, head) staged once in
// shared memory, 1-16 warps x 16 queries, several query tiles per CTA; 16 warps x 2 CTAs per SM for L >= 512) and the
// streaming `attention_kernel` below for K/V that do not fit (VQVAE: L = 1024, d = 64): CTA = 4 warps = 64 queries of
// one (batch, head); KV streamed in 64-key tiles, double-buffered with cp.async.
#include "host.h"
#include <cuda_bf16.h>
#include <type_traits>
#include <stdlib.h>

namespace b200 {

constexpr int ATT_BQ = 64;
constexpr int ATT_BKV = 64;
constexpr int ATT_THREADS = 128;

__device__ __forceinline__ void cp_async16(void* dst, const void* src, int src_bytes) {
    uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(dst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const void* p) {
    uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(p));
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(a));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], const void* p) {
    uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(p));
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(a));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
}

// DP: head dim padded to a multiple of 16. Smem row pitch DP*2+16 B makes every ldmatrix conflict-free.
template <int DP>
__global__ void __launch_bounds__(ATT_THREADS)
    attention_kernel(const __nv_bfloat16* __restrict__ q, long long ld_q, const __nv_bfloat16* __restrict__ k,
                     long long ld_k, const __nv_bfloat16* __restrict__ v, long long ld_v,
                     __nv_bfloat16* __restrict__ out, long long ld_o, int lq, int lk, int d, float scale_log2) {
    pdl_enter();
    constexpr int PITCH = DP + 8;  // elements
    extern __shared__ __align__(16) uint8_t att_smem[];
    __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(att_smem);
    __nv_bfloat16* sK = sQ + ATT_BQ * PITCH;        // 2 stages
    __nv_bfloat16* sV = sK + 2 * ATT_BKV * PITCH;   // 2 stages

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int q0 = blockIdx.x * ATT_BQ;
    const int head = blockIdx.y;
    const long long b = blockIdx.z;
    const __nv_bfloat16* qg = q + (b * lq) * ld_q + head * d;
    const __nv_bfloat16* kg = k + (b * lk) * ld_k + head * d;
    const __nv_bfloat16* vg = v + (b * lk) * ld_v + head * d;
    const int cpr = d >> 3;  // valid 16-byte chunks per row

    // zero the padding columns [d, PITCH) of every buffer once; cp.async only ever writes [0, d)
    if (d < PITCH) {
        const int padw = PITCH - d;
        for (int i = tid; i < (ATT_BQ + 4 * ATT_BKV) * padw; i += ATT_THREADS) {
            const int row = i / padw, c = d + i % padw;
            sQ[row * PITCH + c] = __float2bfloat16(0.f);
        }
    }
    auto load_rows = [&](__nv_bfloat16* dst, const __nv_bfloat16* src, long long ld, int row0, int nrows_valid) {
        for (int i = tid; i < 64 * cpr; i += ATT_THREADS) {
            const int r = i / cpr, c = i - r * cpr;
            const bool ok = (row0 + r) < nrows_valid;
            const __nv_bfloat16* s = src + (long long)(ok ? row0 + r : 0) * ld + c * 8;
            cp_async16(dst + r * PITCH + c * 8, s, ok ? 16 : 0);
        }
    };
    load_rows(sQ, qg, ld_q, q0, lq);
    load_rows(sK, kg, ld_k, 0, lk);
    load_rows(sV, vg, ld_v, 0, lk);
    cp_async_commit();

    const int ntiles = (lk + ATT_BKV - 1) / ATT_BKV;
    float o[DP / 8][4];
#pragma unroll
    for (int i = 0; i < DP / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    uint32_t qf[DP / 16][4];

    for (int t = 0; t < ntiles; ++t) {
        const int st = t & 1;
        if (t + 1 < ntiles) {
            load_rows(sK + (st ^ 1) * ATT_BKV * PITCH, kg, ld_k, (t + 1) * ATT_BKV, lk);
            load_rows(sV + (st ^ 1) * ATT_BKV * PITCH, vg, ld_v, (t + 1) * ATT_BKV, lk);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        if (t == 0) {
#pragma unroll
            for (int kb = 0; kb < DP / 16; ++kb)
                ldsm_x4(qf[kb], sQ + (warp * 16 + (lane & 7) + 8 * ((lane >> 3) & 1)) * PITCH + kb * 16 + 8 * (lane >> 4));
        }
        const __nv_bfloat16* tK = sK + st * ATT_BKV * PITCH;
        const __nv_bfloat16* tV = sV + st * ATT_BKV * PITCH;

        // S = Q K^T : 16 x 64 per warp
        float s[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
        for (int kb = 0; kb < DP / 16; ++kb) {
#pragma unroll
            for (int np = 0; np < 4; ++np) {
                uint32_t bf[4];
                ldsm_x4(bf, tK + (np * 16 + (lane & 7) + 8 * (lane >> 4)) * PITCH + kb * 16 + 8 * ((lane >> 3) & 1));
                mma16816(s[2 * np], qf[kb], bf[0], bf[1]);
                mma16816(s[2 * np + 1], qf[kb], bf[2], bf[3]);
            }
        }
        // mask keys beyond lk, running max
        const int kbase = t * ATT_BKV + (lane & 3) * 2;
        float mx0 = m0, mx1 = m1;
#pragma unroll
        for (int nb = 0; nb < 8; ++nb) {
            const int key = kbase + nb * 8;
            if (key >= lk) s[nb][0] = s[nb][2] = -INFINITY;
            if (key + 1 >= lk) s[nb][1] = s[nb][3] = -INFINITY;
            mx0 = fmaxf(mx0, fmaxf(s[nb][0], s[nb][1]));
            mx1 = fmaxf(mx1, fmaxf(s[nb][2], s[nb][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float a0 = exp2f((m0 - mx0) * scale_log2), a1 = exp2f((m1 - mx1) * scale_log2);
        m0 = mx0, m1 = mx1;
        const float ms0 = mx0 * scale_log2, ms1 = mx1 * scale_log2;
        float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
        for (int nb = 0; nb < 8; ++nb) {
            s[nb][0] = exp2f(s[nb][0] * scale_log2 - ms0);
            s[nb][1] = exp2f(s[nb][1] * scale_log2 - ms0);
            s[nb][2] = exp2f(s[nb][2] * scale_log2 - ms1);
            s[nb][3] = exp2f(s[nb][3] * scale_log2 - ms1);
            rs0 += s[nb][0] + s[nb][1];
            rs1 += s[nb][2] + s[nb][3];
        }
        l0 = l0 * a0 + rs0;
        l1 = l1 * a1 + rs1;
#pragma unroll
        for (int i = 0; i < DP / 8; ++i) {
            o[i][0] *= a0, o[i][1] *= a0;
            o[i][2] *= a1, o[i][3] *= a1;
        }
        // O += P V
#pragma unroll
        for (int kb = 0; kb < 4; ++kb) {
            uint32_t pf[4];
            pf[0] = pack_bf16(s[2 * kb][0], s[2 * kb][1]);
            pf[1] = pack_bf16(s[2 * kb][2], s[2 * kb][3]);
            pf[2] = pack_bf16(s[2 * kb + 1][0], s[2 * kb + 1][1]);
            pf[3] = pack_bf16(s[2 * kb + 1][2], s[2 * kb + 1][3]);
#pragma unroll
            for (int dp = 0; dp < DP / 16; ++dp) {
                uint32_t bf[4];
                ldsm_x4_t(bf, tV + (kb * 16 + (lane & 7) + 8 * ((lane >> 3) & 1)) * PITCH + dp * 16 + 8 * (lane >> 4));
                mma16816(o[2 * dp], pf, bf[0], bf[1]);
                mma16816(o[2 * dp + 1], pf, bf[2], bf[3]);
            }
        }
        __syncthreads();  // everyone done with stage st before it is refilled at iteration t+1
    }
    // finalise: row sums across the quad, normalise, store
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.f / l0, i1 = 1.f / l1;
    const int r0 = q0 + warp * 16 + (lane >> 2), r1 = r0 + 8;
    __nv_bfloat16* og = out + (b * lq) * ld_o + head * d;
#pragma unroll
    for (int nb = 0; nb < DP / 8; ++nb) {
        const int col = nb * 8 + (lane & 3) * 2;
        if (col < d) {
            if (r0 < lq)
                *reinterpret_cast<__nv_bfloat162*>(og + (long long)r0 * ld_o + col) =
                    __floats2bfloat162_rn(o[nb][0] * i0, o[nb][1] * i0);
            if (r1 < lq)
                *reinterpret_cast<__nv_bfloat162*>(og + (long long)r1 * ld_o + col) =
                    __floats2bfloat162_rn(o[nb][2] * i1, o[nb][3] * i1);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Resident-KV variant (every attention inside the UNet / DiT: lk <= 1024 with d <= 32, or lk <= 256 with d <= 64).
// With head dims this small the kernel is bound by the softmax arithmetic (one exp + ~6 fp32 ops per score against
// 4*d tensor flops), so the structure removes everything else from the inner loop:
//   * K and V of the whole (batch, head) are staged ONCE in shared memory; after one barrier each warp streams
//     through all keys on its own — no per-tile barriers, no per-tile global loads
//   * the head dim is a template parameter in units of 8 and is NOT padded to 16: the odd 8 columns use
//     mma.m16n8k8 (d = 8 -> one k8 MMA, d = 24 -> k16 + k8), rows are 16-byte chunks at an odd 16-byte pitch so
//     every ldmatrix is bank-conflict free without padding columns to zero
//   * exp2 via ex2.approx.ftz with the softmax scale folded into one FFMA; key masking only in the last tile, and
//     the last tile only touches the 8-key groups that exist (lk = 77 costs 64 + 16 keys, not 128)
template <int D8>
struct AttRes {
    static constexpr int P16 = (D8 & 1) ? D8 : D8 + 1;  // row pitch in 16-byte units (odd)
    static constexpr int PITCH = P16 * 8;               // elements
};

__device__ __forceinline__ void ldsm_x2(uint32_t (&r)[2], const void* p) {
    uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(p));
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(a));
}
__device__ __forceinline__ void ldsm_x2_t(uint32_t (&r)[2], const void* p) {
    uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(p));
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(a));
}
__device__ __forceinline__ void mma1688(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ uint32_t ldg32(const __nv_bfloat16* p) { return *reinterpret_cast<const uint32_t*>(p); }

// NT / MINB: launch bounds. <512, 2> (64 registers) puts 32 warps on an SM for the L >= 512 sites, where the kernel
// was latency-bound at 16 warps (ncu r1j: issue 44 %, MUFU 48 %, tensor 39 %, none saturated).
// Each CTA stages K/V once and then walks `qiter` query tiles of (warps * 16) rows; the Q fragments come straight
// from global memory in mma A-operand layout (no shared-memory staging, no barrier per query tile).
template <int D8, int NT, int MINB>
__global__ void __launch_bounds__(NT, MINB)
    attention_resident_kernel(const __nv_bfloat16* __restrict__ q, long long ld_q, const __nv_bfloat16* __restrict__ k,
                              long long ld_k, const __nv_bfloat16* __restrict__ v, long long ld_v,
                              __nv_bfloat16* __restrict__ out, long long ld_o, int lq, int lk, float scale_log2,
                              float* __restrict__ lse, int qiter) {
    pdl_enter();
    constexpr int PITCH = AttRes<D8>::PITCH;
    constexpr int K16 = D8 / 2;        // k16 chunks of the head dim
    constexpr bool TAIL = (D8 & 1);    // one trailing k8 chunk
    extern __shared__ __align__(16) uint8_t att_smem[];
    const int nthreads = blockDim.x, bq = (nthreads >> 5) * 16;
    const int lk_pad = (lk + 63) & ~63;
    __nv_bfloat16* sK = reinterpret_cast<__nv_bfloat16*>(att_smem);
    __nv_bfloat16* sV = sK + lk_pad * PITCH;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int head = blockIdx.y;
    const long long b = blockIdx.z;
    const __nv_bfloat16* qg = q + (b * lq) * ld_q + head * (D8 * 8);
    const __nv_bfloat16* kg = k + (b * lk) * ld_k + head * (D8 * 8);
    const __nv_bfloat16* vg = v + (b * lk) * ld_v + head * (D8 * 8);

    for (int i = tid; i < lk_pad * D8; i += nthreads) {
        const int r = i / D8, c = i - r * D8;
        const bool ok = r < lk;
        const long long ro = ok ? r : 0;
        cp_async16(sK + r * PITCH + c * 8, kg + ro * ld_k + c * 8, ok ? 16 : 0);
        cp_async16(sV + r * PITCH + c * 8, vg + ro * ld_v + c * 8, ok ? 16 : 0);
    }
    cp_async_commit();
    cp_async_wait<0>();
    __syncthreads();  // the only barrier

    const int ntiles = lk_pad >> 6;
    const int nfull = lk >> 6;
#pragma unroll 1
    for (int it = 0; it < qiter; ++it) {
    const int qw = (blockIdx.x * qiter + it) * bq + warp * 16;  // this warp's first query row
    if (qw >= lq) break;
    uint32_t qf[K16 > 0 ? K16 : 1][4];
    uint32_t qt[2] = {0u, 0u};
    {
        const int ra = qw + (lane >> 2), rb = ra + 8;
        const __nv_bfloat16* pa = qg + static_cast<long long>(ra < lq ? ra : lq - 1) * ld_q + (lane & 3) * 2;
        const __nv_bfloat16* pb = qg + static_cast<long long>(rb < lq ? rb : lq - 1) * ld_q + (lane & 3) * 2;
#pragma unroll
        for (int kb = 0; kb < K16; ++kb) {
            qf[kb][0] = ldg32(pa + kb * 16), qf[kb][1] = ldg32(pb + kb * 16);
            qf[kb][2] = ldg32(pa + kb * 16 + 8), qf[kb][3] = ldg32(pb + kb * 16 + 8);
        }
        if (TAIL) qt[0] = ldg32(pa + K16 * 16), qt[1] = ldg32(pb + K16 * 16);
    }

    float o[D8][4];
#pragma unroll
    for (int i = 0; i < D8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    // One 64-key tile. NPV = 0: a full tile (no masking); NPV = 1..4: the last, partial tile with that many 16-key
    // pairs present — every loop bound is a compile-time constant (lk = 77 costs 64 + 16 keys of straight-line code).
    auto tile = [&](int t, auto npv_tag) {
        constexpr int NPV = decltype(npv_tag)::value;
        constexpr bool FULL = NPV == 0;
        constexpr int NPX = FULL ? 4 : NPV;  // 16-key pairs touched
        const __nv_bfloat16* tK = sK + t * 64 * PITCH;
        const __nv_bfloat16* tV = sV + t * 64 * PITCH;
        const int rem = lk - t * 64;  // keys left (>= 1)
        float s[2 * NPX][4];
#pragma unroll
        for (int i = 0; i < 2 * NPX; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
        for (int np = 0; np < NPX; ++np) {
#pragma unroll
            for (int kb = 0; kb < K16; ++kb) {
                uint32_t bf[4];
                ldsm_x4(bf, tK + (np * 16 + (lane & 7) + 8 * (lane >> 4)) * PITCH + kb * 16 + 8 * ((lane >> 3) & 1));
                mma16816(s[2 * np], qf[kb], bf[0], bf[1]);
                mma16816(s[2 * np + 1], qf[kb], bf[2], bf[3]);
            }
        }
        if (TAIL) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (4 * h < 2 * NPX) {
                    uint32_t bf[4];
                    ldsm_x4(bf, tK + (h * 32 + lane) * PITCH + K16 * 16);
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (4 * h + i < 2 * NPX) mma1688(s[4 * h + i], qt[0], qt[1], bf[i]);
                }
            }
        }
        if (!FULL) {  // mask keys beyond lk
            const int kbase = (lane & 3) * 2;
#pragma unroll
            for (int nb = 0; nb < 2 * NPX; ++nb) {
                const int key = kbase + nb * 8;
                if (key >= rem) s[nb][0] = s[nb][2] = -INFINITY;
                if (key + 1 >= rem) s[nb][1] = s[nb][3] = -INFINITY;
            }
        }
        float mx0 = m0, mx1 = m1;
#pragma unroll
        for (int nb = 0; nb < 2 * NPX; ++nb) {
            mx0 = fmaxf(mx0, fmaxf(s[nb][0], s[nb][1]));
            mx1 = fmaxf(mx1, fmaxf(s[nb][2], s[nb][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float a0 = ex2((m0 - mx0) * scale_log2), a1 = ex2((m1 - mx1) * scale_log2);
        m0 = mx0, m1 = mx1;
        const float ms0 = mx0 * scale_log2, ms1 = mx1 * scale_log2;
        float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
        for (int nb = 0; nb < 2 * NPX; ++nb) {
            s[nb][0] = ex2(fmaf(s[nb][0], scale_log2, -ms0));
            s[nb][1] = ex2(fmaf(s[nb][1], scale_log2, -ms0));
            s[nb][2] = ex2(fmaf(s[nb][2], scale_log2, -ms1));
            s[nb][3] = ex2(fmaf(s[nb][3], scale_log2, -ms1));
            rs0 += s[nb][0] + s[nb][1];
            rs1 += s[nb][2] + s[nb][3];
        }
        l0 = l0 * a0 + rs0;
        l1 = l1 * a1 + rs1;
#pragma unroll
        for (int i = 0; i < D8; ++i) {
            o[i][0] *= a0, o[i][1] *= a0;
            o[i][2] *= a1, o[i][3] *= a1;
        }
        // O += P V (16 keys per step; V fragments transposed on load)
#pragma unroll
        for (int kb = 0; kb < NPX; ++kb) {
            uint32_t pf[4];
            pf[0] = pack_bf16(s[2 * kb][0], s[2 * kb][1]);
            pf[1] = pack_bf16(s[2 * kb][2], s[2 * kb][3]);
            pf[2] = pack_bf16(s[2 * kb + 1][0], s[2 * kb + 1][1]);
            pf[3] = pack_bf16(s[2 * kb + 1][2], s[2 * kb + 1][3]);
            const __nv_bfloat16* vrow = tV + (kb * 16 + (lane & 7) + 8 * ((lane >> 3) & 1)) * PITCH;
#pragma unroll
            for (int dp = 0; dp < K16; ++dp) {
                uint32_t bf[4];
                ldsm_x4_t(bf, vrow + dp * 16 + 8 * (lane >> 4));
                mma16816(o[2 * dp], pf, bf[0], bf[1]);
                mma16816(o[2 * dp + 1], pf, bf[2], bf[3]);
            }
            if (TAIL) {
                uint32_t bf[2];
                ldsm_x2_t(bf, vrow + K16 * 16);
                mma16816(o[D8 - 1], pf, bf[0], bf[1]);
            }
        }
    };
#pragma unroll 1
    for (int t = 0; t < nfull; ++t) tile(t, std::integral_constant<int, 0>{});
    if (nfull < ntiles) {
        switch ((lk - nfull * 64 + 15) >> 4) {  // 16-key pairs in the last tile (warp-uniform)
            case 1: tile(nfull, std::integral_constant<int, 1>{}); break;
            case 2: tile(nfull, std::integral_constant<int, 2>{}); break;
            case 3: tile(nfull, std::integral_constant<int, 3>{}); break;
            default: tile(nfull, std::integral_constant<int, 4>{}); break;
        }
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.f / l0, i1 = 1.f / l1;
    const int r0 = qw + (lane >> 2), r1 = r0 + 8;
    if (lse && (lane & 3) == 0) {  // log2-domain log-sum-exp of the scaled scores, for the backward pass
        float* lp = lse + (b * gridDim.y + head) * lq;
        if (r0 < lq) lp[r0] = m0 * scale_log2 + log2f(l0);
        if (r1 < lq) lp[r1] = m1 * scale_log2 + log2f(l1);
    }
    __nv_bfloat16* og = out + (b * lq) * ld_o + head * (D8 * 8);
#pragma unroll
    for (int nb = 0; nb < D8; ++nb) {
        const int col = nb * 8 + (lane & 3) * 2;
        if (r0 < lq)
            *reinterpret_cast<__nv_bfloat162*>(og + (long long)r0 * ld_o + col) =
                __floats2bfloat162_rn(o[nb][0] * i0, o[nb][1] * i0);
        if (r1 < lq)
            *reinterpret_cast<__nv_bfloat162*>(og + (long long)r1 * ld_o + col) =
                __floats2bfloat162_rn(o[nb][2] * i1, o[nb][3] * i1);
    }
    }  // query tiles
}

static size_t resident_smem(int d8, int lk) {
    const int p16 = (d8 & 1) ? d8 : d8 + 1;
    const size_t lk_pad = (size_t)((lk + 63) & ~63);
    return 2 * lk_pad * p16 * 16;
}

template <int D8>
static int launch_resident(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v, int64_t ld_v,
                           void* out, int64_t ld_o, int64_t batch, int heads, int lq, int lk, float scale,
                           float* lse, cudaStream_t st) {
    constexpr int MB256 = D8 <= 4 ? 3 : 0;  // 80 registers -> three 8-warp CTAs per SM for the small head dims
    const size_t smem = resident_smem(D8, lk);
    // 16 warps x 2 CTAs per SM where the 64-register budget holds without heavy spilling (head dim <= 32)
    static const int big_min = [] { const char* f = getenv("B200SD_ATT_BIG_MINLQ"); return f ? atoi(f) : 512; }();
    const bool big = D8 <= 4 && lq >= big_min && smem <= 112 * 1024;
    const int warps = big ? 16 : lq >= 128 ? 8 : lq > 32 ? 4 : lq > 16 ? 2 : 1;
    const int bq = warps * 16;
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(attention_resident_kernel<D8, 256, MB256>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             200 * 1024);
        cudaFuncSetAttribute(attention_resident_kernel<D8, 512, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             112 * 1024);
        configured = true;
    }
    // query tiles per CTA: as many as leave >= 6 resident waves of CTAs (K/V staging is paid once per CTA)
    const int qtiles = (lq + bq - 1) / bq;
    const long long cap = (long long)sm_count() * (big ? 2 : 3);
    int qiter = 1;
    while (qiter * 2 <= qtiles && (long long)((qtiles + 2 * qiter - 1) / (2 * qiter)) * heads * batch >= 6 * cap)
        qiter *= 2;
    dim3 grid((qtiles + qiter - 1) / qiter, heads, (unsigned)batch);
    if (big)
        launch_k(attention_resident_kernel<D8, 512, 2>, grid, warps * 32, smem, st,
                 static_cast<const __nv_bfloat16*>(q), ld_q, static_cast<const __nv_bfloat16*>(k), ld_k,
                 static_cast<const __nv_bfloat16*>(v), ld_v, static_cast<__nv_bfloat16*>(out), ld_o, lq, lk,
                 scale * 1.4426950408889634f, lse, qiter);
    else
        launch_k(attention_resident_kernel<D8, 256, MB256>, grid, warps * 32, smem, st,
                 static_cast<const __nv_bfloat16*>(q), ld_q, static_cast<const __nv_bfloat16*>(k), ld_k,
                 static_cast<const __nv_bfloat16*>(v), ld_v, static_cast<__nv_bfloat16*>(out), ld_o, lq, lk,
                 scale * 1.4426950408889634f, lse, qiter);
    B200_CHECK_LAUNCH("attention_resident");
    return B200SD_OK;
}

template <int DP>
static int launch_attention(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v, int64_t ld_v,
                            void* out, int64_t ld_o, int64_t batch, int heads, int lq, int lk, int d, float scale,
                            cudaStream_t st) {
    constexpr int PITCH = DP + 8;
    const size_t smem = (size_t)(ATT_BQ + 4 * ATT_BKV) * PITCH * 2;
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(attention_kernel<DP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        configured = true;
    }
    dim3 grid((lq + ATT_BQ - 1) / ATT_BQ, heads, (unsigned)batch);
    launch_k(attention_kernel<DP>, grid, ATT_THREADS, smem, st, 
        static_cast<const __nv_bfloat16*>(q), ld_q, static_cast<const __nv_bfloat16*>(k), ld_k,
        static_cast<const __nv_bfloat16*>(v), ld_v, static_cast<__nv_bfloat16*>(out), ld_o, lq, lk, d,
        scale * 1.4426950408889634f);
    B200_CHECK_LAUNCH("attention");
    return B200SD_OK;
}


// ---------------------------------------------------------------------------------------------------------------
// Attention BACKWARD (resident operands, same structure as the forward). One kernel, two roles:
//   DKV = false ("dQ"):   rows = queries.  R1 = Q, R2 = dO, resident C1 = K, C2 = V.   outB = dQ = dS K
//   DKV = true  ("dK,dV"): rows = keys.    R1 = K, R2 = V,  resident C1 = Q, C2 = dO.  outB = dK = dS^T Q, outA = dV = P^T dO
// Per 64-column tile: S = R1 C1^T, dP = R2 C2^T, P = exp2(S*scale_log2 - L), dS = P o (dP - D) * scale, where L (the
// forward's log-sum-exp) and D = rowsum(dO o O) belong to the query: per ROW in the dQ role, per COLUMN in the dK/dV
// role. Scores are recomputed, never stored (the reference materialises B*H*L*L weights, multihead_attention.py:58-61).
template <int D8, bool DKV>
__global__ void __launch_bounds__(256)
    attention_bwd_kernel(const __nv_bfloat16* __restrict__ r1, long long ld_r1, const __nv_bfloat16* __restrict__ r2,
                         long long ld_r2, const __nv_bfloat16* __restrict__ c1, long long ld_c1,
                         const __nv_bfloat16* __restrict__ c2, long long ld_c2, const float* __restrict__ lse,
                         const float* __restrict__ dvec, __nv_bfloat16* __restrict__ out_b, long long ld_ob,
                         __nv_bfloat16* __restrict__ out_a, long long ld_oa, int lr, int lc, int lq, float scale,
                         float scale_log2) {
    pdl_enter();
    constexpr int PITCH = AttRes<D8>::PITCH;
    constexpr int K16 = D8 / 2;
    constexpr bool TAIL = (D8 & 1);
    extern __shared__ __align__(16) uint8_t att_smem[];
    const int nthreads = blockDim.x, bq = (nthreads >> 5) * 16;
    const int lc_pad = (lc + 63) & ~63;
    __nv_bfloat16* sC1 = reinterpret_cast<__nv_bfloat16*>(att_smem);
    __nv_bfloat16* sC2 = sC1 + lc_pad * PITCH;
    __nv_bfloat16* sR1 = sC2 + lc_pad * PITCH;
    __nv_bfloat16* sR2 = sR1 + bq * PITCH;
    float* sL = reinterpret_cast<float*>(sR2 + bq * PITCH);  // [lc_pad] (DKV only)
    float* sD = sL + lc_pad;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int row0 = blockIdx.x * bq;
    const int head = blockIdx.y;
    const long long b = blockIdx.z;
    const int hoff = head * (D8 * 8);
    const __nv_bfloat16* r1g = r1 + (b * lr) * ld_r1 + hoff;
    const __nv_bfloat16* r2g = r2 + (b * lr) * ld_r2 + hoff;
    const __nv_bfloat16* c1g = c1 + (b * lc) * ld_c1 + hoff;
    const __nv_bfloat16* c2g = c2 + (b * lc) * ld_c2 + hoff;
    const float* lg = lse + (b * gridDim.y + head) * lq;
    const float* dg = dvec + (b * gridDim.y + head) * lq;

    for (int i = tid; i < lc_pad * D8; i += nthreads) {
        const int r = i / D8, c = i - r * D8;
        const bool ok = r < lc;
        const long long ro = ok ? r : 0;
        cp_async16(sC1 + r * PITCH + c * 8, c1g + ro * ld_c1 + c * 8, ok ? 16 : 0);
        cp_async16(sC2 + r * PITCH + c * 8, c2g + ro * ld_c2 + c * 8, ok ? 16 : 0);
    }
    for (int i = tid; i < bq * D8; i += nthreads) {
        const int r = i / D8, c = i - r * D8;
        const bool ok = (row0 + r) < lr;
        const long long ro = ok ? row0 + r : 0;
        cp_async16(sR1 + r * PITCH + c * 8, r1g + ro * ld_r1 + c * 8, ok ? 16 : 0);
        cp_async16(sR2 + r * PITCH + c * 8, r2g + ro * ld_r2 + c * 8, ok ? 16 : 0);
    }
    cp_async_commit();
    if (DKV) {
        for (int i = tid; i < lc_pad; i += nthreads) {
            sL[i] = i < lc ? lg[i] : 0.f;
            sD[i] = i < lc ? dg[i] : 0.f;
        }
    }
    cp_async_wait<0>();
    __syncthreads();
    if (row0 + warp * 16 >= lr) return;

    uint32_t f1[K16 > 0 ? K16 : 1][4], f2[K16 > 0 ? K16 : 1][4];
    uint32_t t1[2] = {0u, 0u}, t2[2] = {0u, 0u};
    {
        const int ro = (warp * 16 + (lane & 7) + 8 * ((lane >> 3) & 1)) * PITCH;
#pragma unroll
        for (int kb = 0; kb < K16; ++kb) {
            ldsm_x4(f1[kb], sR1 + ro + kb * 16 + 8 * (lane >> 4));
            ldsm_x4(f2[kb], sR2 + ro + kb * 16 + 8 * (lane >> 4));
        }
        if (TAIL) {
            ldsm_x2(t1, sR1 + ro + K16 * 16);
            ldsm_x2(t2, sR2 + ro + K16 * 16);
        }
    }
    const int rr0 = row0 + warp * 16 + (lane >> 2), rr1 = rr0 + 8;
    float L0 = 0.f, L1 = 0.f, D0 = 0.f, D1 = 0.f;
    if (!DKV) {
        if (rr0 < lr) L0 = lg[rr0], D0 = dg[rr0];
        if (rr1 < lr) L1 = lg[rr1], D1 = dg[rr1];
    }
    float acc_b[D8][4], acc_a[DKV ? D8 : 1][4];
#pragma unroll
    for (int i = 0; i < D8; ++i) acc_b[i][0] = acc_b[i][1] = acc_b[i][2] = acc_b[i][3] = 0.f;
#pragma unroll
    for (int i = 0; i < (DKV ? D8 : 1); ++i) acc_a[i][0] = acc_a[i][1] = acc_a[i][2] = acc_a[i][3] = 0.f;

    const int ntiles = lc_pad >> 6;
    for (int t = 0; t < ntiles; ++t) {
        const __nv_bfloat16* tC1 = sC1 + t * 64 * PITCH;
        const __nv_bfloat16* tC2 = sC2 + t * 64 * PITCH;
        const int rem = lc - t * 64;
        const int nb_valid = rem >= 64 ? 8 : (rem + 7) >> 3;
        float s[8][4], dp[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
            dp[i][0] = dp[i][1] = dp[i][2] = dp[i][3] = 0.f;
        }
#pragma unroll
        for (int np = 0; np < 4; ++np) {
            if (2 * np < nb_valid) {
                const int off = (np * 16 + (lane & 7) + 8 * (lane >> 4)) * PITCH + 8 * ((lane >> 3) & 1);
#pragma unroll
                for (int kb = 0; kb < K16; ++kb) {
                    uint32_t bf[4];
                    ldsm_x4(bf, tC1 + off + kb * 16);
                    mma16816(s[2 * np], f1[kb], bf[0], bf[1]);
                    mma16816(s[2 * np + 1], f1[kb], bf[2], bf[3]);
                    ldsm_x4(bf, tC2 + off + kb * 16);
                    mma16816(dp[2 * np], f2[kb], bf[0], bf[1]);
                    mma16816(dp[2 * np + 1], f2[kb], bf[2], bf[3]);
                }
            }
        }
        if (TAIL) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (4 * h < nb_valid) {
                    uint32_t bf[4];
                    ldsm_x4(bf, tC1 + (h * 32 + lane) * PITCH + K16 * 16);
#pragma unroll
                    for (int i = 0; i < 4; ++i) mma1688(s[4 * h + i], t1[0], t1[1], bf[i]);
                    ldsm_x4(bf, tC2 + (h * 32 + lane) * PITCH + K16 * 16);
#pragma unroll
                    for (int i = 0; i < 4; ++i) mma1688(dp[4 * h + i], t2[0], t2[1], bf[i]);
                }
            }
        }
        // P and dS (dS overwrites dp)
        const int cbase = t * 64 + (lane & 3) * 2;
#pragma unroll
        for (int nb = 0; nb < 8; ++nb) {
            if (nb < nb_valid) {
                const int col = cbase + nb * 8;
                float la0 = L0, la1 = L0, lb0 = L1, lb1 = L1, da0 = D0, da1 = D0, db0 = D1, db1 = D1;
                if (DKV) {
                    la0 = lb0 = sL[col], la1 = lb1 = sL[col + 1];
                    da0 = db0 = sD[col], da1 = db1 = sD[col + 1];
                }
                const bool v0 = col < lc, v1 = col + 1 < lc;
                const float p0 = v0 ? ex2(fmaf(s[nb][0], scale_log2, -la0)) : 0.f;
                const float p1 = v1 ? ex2(fmaf(s[nb][1], scale_log2, -la1)) : 0.f;
                const float p2 = v0 ? ex2(fmaf(s[nb][2], scale_log2, -lb0)) : 0.f;
                const float p3 = v1 ? ex2(fmaf(s[nb][3], scale_log2, -lb1)) : 0.f;
                s[nb][0] = p0, s[nb][1] = p1, s[nb][2] = p2, s[nb][3] = p3;
                dp[nb][0] = p0 * (dp[nb][0] - da0) * scale;
                dp[nb][1] = p1 * (dp[nb][1] - da1) * scale;
                dp[nb][2] = p2 * (dp[nb][2] - db0) * scale;
                dp[nb][3] = p3 * (dp[nb][3] - db1) * scale;
            } else {
                s[nb][0] = s[nb][1] = s[nb][2] = s[nb][3] = 0.f;
                dp[nb][0] = dp[nb][1] = dp[nb][2] = dp[nb][3] = 0.f;
            }
        }
#pragma unroll
        for (int kb = 0; kb < 4; ++kb) {
            if (2 * kb < nb_valid) {
                uint32_t pf[4], df[4];
                df[0] = pack_bf16(dp[2 * kb][0], dp[2 * kb][1]);
                df[1] = pack_bf16(dp[2 * kb][2], dp[2 * kb][3]);
                df[2] = pack_bf16(dp[2 * kb + 1][0], dp[2 * kb + 1][1]);
                df[3] = pack_bf16(dp[2 * kb + 1][2], dp[2 * kb + 1][3]);
                if (DKV) {
                    pf[0] = pack_bf16(s[2 * kb][0], s[2 * kb][1]);
                    pf[1] = pack_bf16(s[2 * kb][2], s[2 * kb][3]);
                    pf[2] = pack_bf16(s[2 * kb + 1][0], s[2 * kb + 1][1]);
                    pf[3] = pack_bf16(s[2 * kb + 1][2], s[2 * kb + 1][3]);
                }
                const int off = (kb * 16 + (lane & 7) + 8 * ((lane >> 3) & 1)) * PITCH;
#pragma unroll
                for (int d2 = 0; d2 < K16; ++d2) {
                    uint32_t bf[4];
                    ldsm_x4_t(bf, tC1 + off + d2 * 16 + 8 * (lane >> 4));
                    mma16816(acc_b[2 * d2], df, bf[0], bf[1]);
                    mma16816(acc_b[2 * d2 + 1], df, bf[2], bf[3]);
                    if (DKV) {
                        ldsm_x4_t(bf, tC2 + off + d2 * 16 + 8 * (lane >> 4));
                        mma16816(acc_a[2 * d2], pf, bf[0], bf[1]);
                        mma16816(acc_a[2 * d2 + 1], pf, bf[2], bf[3]);
                    }
                }
                if (TAIL) {
                    uint32_t bf[2];
                    ldsm_x2_t(bf, tC1 + off + K16 * 16);
                    mma16816(acc_b[D8 - 1], df, bf[0], bf[1]);
                    if (DKV) {
                        ldsm_x2_t(bf, tC2 + off + K16 * 16);
                        mma16816(acc_a[D8 - 1], pf, bf[0], bf[1]);
                    }
                }
            }
        }
    }
    __nv_bfloat16* ob = out_b + (b * lr) * ld_ob + hoff;
    __nv_bfloat16* oa = DKV ? out_a + (b * lr) * ld_oa + hoff : nullptr;
#pragma unroll
    for (int nb = 0; nb < D8; ++nb) {
        const int col = nb * 8 + (lane & 3) * 2;
        if (rr0 < lr) {
            *reinterpret_cast<__nv_bfloat162*>(ob + (long long)rr0 * ld_ob + col) =
                __floats2bfloat162_rn(acc_b[nb][0], acc_b[nb][1]);
            if (DKV)
                *reinterpret_cast<__nv_bfloat162*>(oa + (long long)rr0 * ld_oa + col) =
                    __floats2bfloat162_rn(acc_a[nb][0], acc_a[nb][1]);
        }
        if (rr1 < lr) {
            *reinterpret_cast<__nv_bfloat162*>(ob + (long long)rr1 * ld_ob + col) =
                __floats2bfloat162_rn(acc_b[nb][2], acc_b[nb][3]);
            if (DKV)
                *reinterpret_cast<__nv_bfloat162*>(oa + (long long)rr1 * ld_oa + col) =
                    __floats2bfloat162_rn(acc_a[nb][2], acc_a[nb][3]);
        }
    }
}

// D[b, h, i] = sum_d dO[i, h*d + j] * O[i, h*d + j]
__global__ void attn_dot_kernel(const __nv_bfloat16* __restrict__ o, long long ld_o, const __nv_bfloat16* __restrict__ d_o,
                                long long ld_do, float* __restrict__ dvec, long long batch, int heads, int lq, int d) {
    pdl_enter();
    const long long total = batch * heads * lq;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int row = static_cast<int>(i % lq);
        const long long bh = i / lq;
        const int h = static_cast<int>(bh % heads);
        const long long b = bh / heads;
        const __nv_bfloat16* po = o + (b * lq + row) * ld_o + h * d;
        const __nv_bfloat16* pd = d_o + (b * lq + row) * ld_do + h * d;
        float acc = 0.f;
        for (int j = 0; j < d; j += 8) {
            const uint4 uo = *reinterpret_cast<const uint4*>(po + j), ud = *reinterpret_cast<const uint4*>(pd + j);
            const __nv_bfloat162* ho = reinterpret_cast<const __nv_bfloat162*>(&uo);
            const __nv_bfloat162* hd = reinterpret_cast<const __nv_bfloat162*>(&ud);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 a = __bfloat1622float2(ho[e]), c = __bfloat1622float2(hd[e]);
                acc += a.x * c.x + a.y * c.y;
            }
        }
        dvec[i] = acc;
    }
}

static size_t bwd_smem(int d8, int lc, int bq) {
    const int p16 = (d8 & 1) ? d8 : d8 + 1;
    const size_t lc_pad = (size_t)((lc + 63) & ~63);
    return (2 * lc_pad + 2 * bq) * p16 * 16 + 2 * lc_pad * sizeof(float);
}

template <int D8, bool DKV>
static int launch_bwd(const void* r1, int64_t ld_r1, const void* r2, int64_t ld_r2, const void* c1, int64_t ld_c1,
                      const void* c2, int64_t ld_c2, const float* lse, const float* dvec, void* out_b, int64_t ld_ob,
                      void* out_a, int64_t ld_oa, int64_t batch, int heads, int lr, int lc, int lq, float scale,
                      cudaStream_t st) {
    const int warps = lr >= 128 ? 8 : lr > 32 ? 4 : lr > 16 ? 2 : 1;
    const int bq = warps * 16;
    const size_t smem = bwd_smem(D8, lc, bq);
    static bool configured = false;
    if (!configured) {
        cudaFuncSetAttribute(attention_bwd_kernel<D8, DKV>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        configured = true;
    }
    dim3 grid((lr + bq - 1) / bq, heads, (unsigned)batch);
    launch_k(attention_bwd_kernel<D8, DKV>, grid, warps * 32, smem, st, 
        static_cast<const __nv_bfloat16*>(r1), ld_r1, static_cast<const __nv_bfloat16*>(r2), ld_r2,
        static_cast<const __nv_bfloat16*>(c1), ld_c1, static_cast<const __nv_bfloat16*>(c2), ld_c2, lse, dvec,
        static_cast<__nv_bfloat16*>(out_b), ld_ob, static_cast<__nv_bfloat16*>(out_a), ld_oa, lr, lc, lq, scale,
        scale * 1.4426950408889634f);
    B200_CHECK_LAUNCH("attention_bwd");
    return B200SD_OK;
}

template <int D8>
static int attention_bwd_d8(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v, int64_t ld_v,
                            const void* d_o, int64_t ld_do, const float* lse, const float* dvec, void* dq, int64_t ld_dq,
                            void* dk, int64_t ld_dk, void* dv, int64_t ld_dv, int64_t batch, int heads, int lq, int lk,
                            float scale, cudaStream_t st) {
    int rc = launch_bwd<D8, false>(q, ld_q, d_o, ld_do, k, ld_k, v, ld_v, lse, dvec, dq, ld_dq, nullptr, 0, batch, heads,
                                   lq, lk, lq, scale, st);
    if (rc) return rc;
    return launch_bwd<D8, true>(k, ld_k, v, ld_v, q, ld_q, d_o, ld_do, lse, dvec, dk, ld_dk, dv, ld_dv, batch, heads, lk,
                                lq, lq, scale, st);
}

}  // namespace b200

using namespace b200;

namespace b200 {
int attention_tc(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v, int64_t ld_v, void* out,
                 int64_t ld_o, int64_t batch, int heads, int lq, int lk, int d, float scale, float* lse,
                 cudaStream_t st);   // attention_tc.cu; > 0: not taken
}

static int attention_fwd(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v, int64_t ld_v,
                         void* out, int64_t ld_o, int64_t batch, int heads, int lq, int lk, int d, float scale,
                         float* lse, void* stream) {
    B200_REQUIRE(q && k && v && out, "attention: null pointer");
    B200_REQUIRE(batch >= 0 && heads >= 1 && lq >= 1 && lk >= 1, "attention: bad sizes");
    B200_REQUIRE(batch <= 65535 && heads <= 65535, "attention: batch/heads exceed grid limits");
    if (d % 8 != 0 || d < 8 || d > 64) {
        set_error("attention: head dim %d unsupported (need a multiple of 8 in [8, 64])", d);
        return B200SD_ERR_UNSUPPORTED;
    }
    B200_REQUIRE(ld_q % 8 == 0 && ld_k % 8 == 0 && ld_v % 8 == 0 && ld_o % 2 == 0, "attention: leading dims unaligned");
    B200_REQUIRE(((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(v)) & 15) == 0 &&
                     (reinterpret_cast<uintptr_t>(out) & 3) == 0,
                 "attention: pointers unaligned");
    if (batch == 0) return B200SD_OK;
    cudaStream_t st = as_stream(stream);
    // tcgen05 / TMEM kernel (attention_tc.cu) for query counts that fill its 128-row tiles; B200SD_ATTN_TC=0 switches it
    // off, B200SD_ATTN_TC_MINLQ moves the threshold (tuning aids)
    // measured on B200 (profiles/r2c_kernel_bench_attention.txt): it wins where the head dim gives the tensor pipe real
    // work (d >= 32: 148 vs 333 us at L = 1024, d = 64; 45 vs 49 us at L = 256, d = 32) and loses at d = 8 / 16 / 24,
    // where both kernels are bound by the exponentials and the mma.sync kernel keeps P in registers
    static const int tc_mode = [] {
        const char* on = getenv("B200SD_ATTN_TC");   // 0: never, 1: whenever it applies, default: by shape
        return on ? (on[0] == '0' ? 0 : 1) : 2;
    }();
    if (tc_mode == 1 || (tc_mode == 2 && d >= 32 && lq >= 128 && lk >= 128)) {
        const int rc = attention_tc(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, d, scale, lse, st);
        if (rc <= 0) return rc;
    }
    if (resident_smem(d / 8, lk) <= 110 * 1024) {  // two CTAs per SM still fit
        switch (d / 8) {
            case 1: return launch_resident<1>(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, scale, lse, st);
            case 2: return launch_resident<2>(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, scale, lse, st);
            case 3: return launch_resident<3>(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, scale, lse, st);
            case 4: return launch_resident<4>(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, scale, lse, st);
            case 5: return launch_resident<5>(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, scale, lse, st);
            case 6: return launch_resident<6>(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, scale, lse, st);
            case 7: return launch_resident<7>(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, scale, lse, st);
            default: return launch_resident<8>(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, scale, lse, st);
        }
    }
    if (lse) {
        set_error("attention: log-sum-exp output needs the resident-KV kernel (lk=%d, d=%d do not fit)", lk, d);
        return B200SD_ERR_UNSUPPORTED;
    }
    const int dp = (d + 15) / 16 * 16;
    switch (dp) {
        case 16: return launch_attention<16>(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, d, scale, st);
        case 32: return launch_attention<32>(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, d, scale, st);
        case 48: return launch_attention<48>(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, d, scale, st);
        default: return launch_attention<64>(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, d, scale, st);
    }
}

extern "C" int b200sd_attention(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v, int64_t ld_v,
                                void* out, int64_t ld_o, int64_t batch, int heads, int lq, int lk, int d, float scale,
                                void* stream) {
    return attention_fwd(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, d, scale, nullptr, stream);
}

/* The tcgen05 / TMEM / TMA kernel (csrc/attention_tc.cu) explicitly, for any shape it supports. */
extern "C" int b200sd_attention_tc(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v, int64_t ld_v,
                                   void* out, int64_t ld_o, float* lse, int64_t batch, int heads, int lq, int lk, int d,
                                   float scale, void* stream) {
    B200_REQUIRE(q && k && v && out, "attention_tc: null pointer");
    B200_REQUIRE(batch >= 0 && heads >= 1 && lq >= 1 && lk >= 1 && heads <= 65535, "attention_tc: bad sizes");
    B200_REQUIRE(((reinterpret_cast<uintptr_t>(q) | reinterpret_cast<uintptr_t>(k) | reinterpret_cast<uintptr_t>(v)) & 15) == 0,
                 "attention_tc: pointers unaligned");
    if (batch == 0) return B200SD_OK;
    const int rc = attention_tc(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, d, scale, lse, as_stream(stream));
    if (rc > 0) {
        set_error("attention_tc: head dim %d / pitches (%lld, %lld, %lld, %lld) / scale %g not supported", d,
                  (long long)ld_q, (long long)ld_k, (long long)ld_v, (long long)ld_o, scale);
        return B200SD_ERR_UNSUPPORTED;
    }
    return rc;
}

extern "C" int b200sd_attention_lse(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v,
                                    int64_t ld_v, void* out, int64_t ld_o, float* lse, int64_t batch, int heads, int lq,
                                    int lk, int d, float scale, void* stream) {
    B200_REQUIRE(lse, "attention_lse: null lse");
    return attention_fwd(q, ld_q, k, ld_k, v, ld_v, out, ld_o, batch, heads, lq, lk, d, scale, lse, stream);
}

/* Backward of softmax(Q K^T * scale) V: dq/dk/dv (bf16, same row/head layout as q/k/v) from d_o, the forward output
 * `o` and the forward's log-sum-exp `lse` (b200sd_attention_lse). dvec: fp32 scratch [batch, heads, lq]. */
extern "C" int b200sd_attention_bwd(const void* q, int64_t ld_q, const void* k, int64_t ld_k, const void* v,
                                    int64_t ld_v, const void* o, int64_t ld_o, const void* d_o, int64_t ld_do,
                                    const float* lse, float* dvec, void* dq, int64_t ld_dq, void* dk, int64_t ld_dk,
                                    void* dv, int64_t ld_dv, int64_t batch, int heads, int lq, int lk, int d,
                                    float scale, void* stream) {
    B200_REQUIRE(q && k && v && o && d_o && lse && dvec && dq && dk && dv, "attention_bwd: null pointer");
    B200_REQUIRE(batch >= 0 && heads >= 1 && lq >= 1 && lk >= 1 && batch <= 65535 && heads <= 65535,
                 "attention_bwd: bad sizes");
    if (d % 8 != 0 || d < 8 || d > 64) {
        set_error("attention_bwd: head dim %d unsupported (need a multiple of 8 in [8, 64])", d);
        return B200SD_ERR_UNSUPPORTED;
    }
    for (int64_t ld : {ld_q, ld_k, ld_v, ld_o, ld_do, ld_dq, ld_dk, ld_dv})
        B200_REQUIRE(ld % 8 == 0, "attention_bwd: leading dims must be multiples of 8");
    if (batch == 0) return B200SD_OK;
    const int lmax = lq > lk ? lq : lk;
    if (bwd_smem(d / 8, lmax, 128) > 220 * 1024) {
        set_error("attention_bwd: lq=%d lk=%d d=%d do not fit the resident kernel", lq, lk, d);
        return B200SD_ERR_UNSUPPORTED;
    }
    cudaStream_t st = as_stream(stream);
    {
        const long long total = batch * heads * (long long)lq;
        long long blocks = (total + 255) / 256;
        if (blocks > 8ll * sm_count()) blocks = 8ll * sm_count();
        launch_k(attn_dot_kernel, (int)blocks, 256, 0, st, static_cast<const __nv_bfloat16*>(o), ld_o,
                                                     static_cast<const __nv_bfloat16*>(d_o), ld_do, dvec, batch, heads, lq, d);
        B200_CHECK_LAUNCH("attention_dot");
    }
#define B200_BWD(N) \
    case N: return attention_bwd_d8<N>(q, ld_q, k, ld_k, v, ld_v, d_o, ld_do, lse, dvec, dq, ld_dq, dk, ld_dk, dv, ld_dv, \
                                       batch, heads, lq, lk, scale, st);
    switch (d / 8) {
        B200_BWD(1) B200_BWD(2) B200_BWD(3) B200_BWD(4) B200_BWD(5) B200_BWD(6) B200_BWD(7)
        default: return attention_bwd_d8<8>(q, ld_q, k, ld_k, v, ld_v, d_o, ld_do, lse, dvec, dq, ld_dq, dk, ld_dk, dv,
                                            ld_dv, batch, heads, lq, lk, scale, st);
    }
#undef B200_BWD
}

// Implicit-GEMM core on tcgen05 / TMEM / TMA (SURVEY K1-K4, K6, K9-K13 projections).
//
// One CTA computes a 128 (output pixels) x BN (output channels) tile:
//   warp 0      : TMA producer — per 64-channel K chunk, one 4-D box load of the activation tile (the tap shift is
//                 a coordinate offset; image borders are TMA out-of-bounds zero fill = conv padding) and one 2-D
//                 box load of the weight tile, both landing 128B-swizzled, K-major in shared memory
//   warp 1      : MMA issuer — lane 0 issues 4 x tcgen05.mma (M=128, N=BN, K=16) per chunk, accumulating fp32 in
//                 TMEM, and releases the smem stage with tcgen05.commit
//   warps 2..5  : epilogue — tcgen05.ld the accumulator (one TMEM lane = one output pixel per thread), add bias /
//                 time-embedding row bias, activation, gate, residual, convert and store
// smem ring: STAGES x (16 KB A + BN*128 B B); full/empty mbarriers; one accumulator-ready mbarrier.
#include "host.h"
#include "ptx.cuh"
#include <cudaTypedefs.h>
#include <mutex>

namespace b200 {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int IGEMM_THREADS = 192;
constexpr int MAX_SEG = 20;

struct SegDev {
    int src, dy, dx, c0, nchunk;
};

struct alignas(64) IgemmArgs {
    CUtensorMap a_map[4];
    CUtensorMap b_map;
    SegDev seg[MAX_SEG];
    int n_seg, total_chunks;
    int n_out;
    int on, oh, ow;  // output pixel grid
    int bw, bh, bn;  // tile box (bw*bh*bn == 128)
    int tiles_x, tiles_y, tiles_img, tiles_col;
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
};

__device__ __forceinline__ float apply_act(float v, int act) {
    if (act == B200SD_ACT_SILU) return v / (1.f + __expf(-v));
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

template <int BN, int STAGES>
struct IgemmSmem {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int BAR_OFF = STAGES * STAGE_BYTES;
    static constexpr int TOTAL = BAR_OFF + (2 * STAGES + 1) * 8 + 16 + 1024;  // +1024: manual alignment slack
};

template <int BN, int STAGES, int MIN_CTAS>
__global__ void __launch_bounds__(IGEMM_THREADS, MIN_CTAS) igemm_kernel(const __grid_constant__ IgemmArgs args) {
    using L = IgemmSmem<BN, STAGES>;
    constexpr uint32_t TMEM_COLS = BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* accum_bar = empty_bar + STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // tile coordinates: column tiles vary fastest so CTAs that share an activation tile are co-scheduled (L2 reuse)
    const int tile = blockIdx.x;
    const int col_t = tile % args.tiles_col;
    int m_t = tile / args.tiles_col;
    const int tx = m_t % args.tiles_x;
    m_t /= args.tiles_x;
    const int ty = m_t % args.tiles_y;
    const int tn = m_t / args.tiles_y;
    const int x0 = tx * args.bw, y0 = ty * args.bh, n0 = tn * args.bn;

    if (warp == 0 && lane == 0) {
        for (int i = 0; i < 4; ++i) tma_prefetch_desc(&args.a_map[i]);
        tma_prefetch_desc(&args.b_map);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(accum_bar, 1);
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

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int it = 0;
            for (int s = 0; s < args.n_seg; ++s) {
                const SegDev sg = args.seg[s];
                for (int ch = 0; ch < sg.nchunk; ++ch, ++it) {
                    const int stage = it % STAGES;
                    const uint32_t phase = (it / STAGES) & 1;
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * L::STAGE_BYTES;
                    uint8_t* sb = sa + L::A_BYTES;
                    mbar_expect_tx(&full_bar[stage], L::STAGE_BYTES);
                    tma_load_4d(sa, &args.a_map[sg.src], &full_bar[stage], sg.c0 + ch * BK, x0 + sg.dx, y0 + sg.dy,
                                n0);
                    tma_load_2d(sb, &args.b_map, &full_bar[stage], it * BK, col_t * BN);
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
        for (int it = 0; it < args.total_chunks; ++it) {
            const int stage = it % STAGES;
            const uint32_t phase = (it / STAGES) & 1;
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            if (lane == 0) {
                const uint32_t sa = smem_u32(smem + stage * L::STAGE_BYTES);
                const uint64_t da = umma_desc_k128(sa);
                const uint64_t db = umma_desc_k128(sa + L::A_BYTES);
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) {
                    // +32 B per K=16 step inside the 128 B swizzle atom: +2 in the (>>4) start-address field
                    umma_bf16(tmem_base, da + 2 * k, db + 2 * k, idesc, (it | k) != 0);
                }
                umma_commit(&empty_bar[stage]);
            }
            __syncwarp();
        }
        if (lane == 0) umma_commit(accum_bar);
        __syncwarp();
    } else {
        // ------------------------------------------------------------------ epilogue (warps 2..5)
        const int q = warp & 3;  // TMEM lane quarter this warp may read
        const int r = q * 32 + lane;
        const int lx = r % args.bw;
        const int ly = (r / args.bw) % args.bh;
        const int ln = r / (args.bw * args.bh);
        const int px = x0 + lx, py = y0 + ly, pn = n0 + ln;
        const bool row_ok = (px < args.ow) && (py < args.oh) && (pn < args.on);
        const long long o_off = pn * args.out_sn + py * args.out_sy + px * args.out_sx;
        const long long r_off = pn * args.res_sn + py * args.res_sy + px * args.res_sx;

        mbar_wait(accum_bar, 0);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
            const int col0 = col_t * BN + c * 32;
            if (col0 >= args.n_out) break;  // warp-uniform
            uint32_t acc[32];
            tmem_ld32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, acc);
            tmem_ld_wait();
            if (!row_ok) continue;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const int col = col0 + g * 8;
                if (col >= args.n_out) break;
                const int nv = min(8, args.n_out - col);
                float v[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(acc[g * 8 + j]);
                if (args.bias) {
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (j < nv) v[j] += __ldg(args.bias + col + j);
                }
                if (args.rowbias) {
                    const __nv_bfloat16* rb = args.rowbias + pn * args.rowbias_ld + col;
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (j < nv) v[j] += __bfloat162float(rb[j]);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = apply_act(v[j], args.act);
                if (args.gate) {
                    const __nv_bfloat16* gp = args.gate + pn * args.gate_ld + col;
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (j < nv) v[j] *= __bfloat162float(gp[j]);
                }
                if (args.residual) {
                    const __nv_bfloat16* rp = args.residual + r_off + col;
                    if (nv == 8 && ((reinterpret_cast<uintptr_t>(rp) & 15) == 0)) {
                        float f[8];
                        unpack8(*reinterpret_cast<const uint4*>(rp), f);
#pragma unroll
                        for (int j = 0; j < 8; ++j) v[j] += f[j];
                    } else {
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            if (j < nv) v[j] += __bfloat162float(rp[j]);
                    }
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
                        uint4 pk;
                        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                        for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
                        *reinterpret_cast<uint4*>(op) = pk;
                    } else {
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            if (j < nv) op[j] = __float2bfloat16_rn(v[j]);
                    }
                }
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
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

template <int BN, int STAGES, int MIN_CTAS>
static int launch_igemm(const IgemmArgs& a, int grid, cudaStream_t st) {
    using L = IgemmSmem<BN, STAGES>;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(igemm_kernel<BN, STAGES, MIN_CTAS>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL);
        if (e != cudaSuccess) {
            set_error("igemm: cannot set %d B dynamic smem: %s", L::TOTAL, cudaGetErrorString(e));
            return B200SD_ERR_CUDA;
        }
        configured = true;
    }
    igemm_kernel<BN, STAGES, MIN_CTAS><<<grid, IGEMM_THREADS, L::TOTAL, st>>>(a);
    B200_CHECK_LAUNCH("igemm");
    return B200SD_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200sd_igemm(const b200sd_igemm_params* p, void* stream) {
    B200_REQUIRE(p, "igemm: null params");
    B200_REQUIRE(p->n_src >= 1 && p->n_src <= 4, "igemm: n_src=%d not in [1,4]", p->n_src);
    B200_REQUIRE(p->n_seg >= 1 && p->n_seg <= MAX_SEG, "igemm: n_seg=%d not in [1,%d]", p->n_seg, MAX_SEG);
    B200_REQUIRE(p->weight && p->out, "igemm: null weight/out");
    B200_REQUIRE(p->n_out >= 1, "igemm: n_out=%d", p->n_out);
    B200_REQUIRE(p->on >= 1 && p->oh >= 1 && p->ow >= 1, "igemm: empty output grid %dx%dx%d", p->on, p->oh, p->ow);
    B200_REQUIRE(p->out_dtype == B200SD_DT_BF16 || p->out_dtype == B200SD_DT_F32, "igemm: bad out dtype");
    B200_REQUIRE(p->stats == nullptr, "igemm: fused statistics are not built in this version");
    if (p->stats) return B200SD_ERR_UNSUPPORTED;

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

    for (int i = 0; i < p->n_src; ++i) {
        const b200sd_src& s = p->srcs[i];
        B200_REQUIRE(s.base && (reinterpret_cast<uintptr_t>(s.base) & 15) == 0, "igemm: source %d base unaligned", i);
        for (int d = 0; d < 3; ++d)
            B200_REQUIRE(s.strides[d] > 0 && s.strides[d] % 16 == 0,
                         "igemm: source %d stride[%d]=%lld B is not a positive multiple of 16", i, d,
                         (long long)s.strides[d]);
        uint64_t dims[4] = {(uint64_t)s.dims[0], (uint64_t)s.dims[1], (uint64_t)s.dims[2], (uint64_t)s.dims[3]};
        uint64_t str[3] = {(uint64_t)s.strides[0], (uint64_t)s.strides[1], (uint64_t)s.strides[2]};
        uint32_t box[4] = {(uint32_t)BK, (uint32_t)bw, (uint32_t)bh, (uint32_t)bn};
        int rc = encode_map(&a.a_map[i], s.base, 4, dims, str, box);
        if (rc) return rc;
    }
    for (int i = p->n_src; i < 4; ++i) a.a_map[i] = a.a_map[0];

    // column tile: largest of {256,192,128,96,64,32} that divides n_out, else by size
    const int n = p->n_out;
    int BNsel;
    const long long m_tiles = (long long)a.tiles_x * a.tiles_y * a.tiles_img;
    if (n % 256 == 0 && m_tiles * (n / 256) >= 2 * sm_count()) BNsel = 256;
    else if (n % 192 == 0 && m_tiles * (n / 192) >= 2 * sm_count()) BNsel = 192;
    else if (n % 128 == 0 && m_tiles * (n / 128) >= sm_count()) BNsel = 128;
    else if (n % 96 == 0 && n % 64 != 0) BNsel = 96;
    else if (n >= 64 && m_tiles * ((n + 63) / 64) >= sm_count() / 2) BNsel = 64;
    else if (n > 32 && n % 64 == 0) BNsel = 64;
    else BNsel = n > 64 ? 64 : 32;
    if (n <= 32) BNsel = 32;
    a.tiles_col = (n + BNsel - 1) / BNsel;

    {
        B200_REQUIRE((reinterpret_cast<uintptr_t>(p->weight) & 15) == 0, "igemm: weight unaligned");
        uint64_t dims[2] = {(uint64_t)total * BK, (uint64_t)n};
        uint64_t str[1] = {(uint64_t)total * BK * 2};
        uint32_t box[2] = {(uint32_t)BK, (uint32_t)BNsel};
        int rc = encode_map(&a.b_map, p->weight, 2, dims, str, box);
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

    const long long grid_ll = m_tiles * a.tiles_col;
    B200_REQUIRE(grid_ll < (1ll << 31), "igemm: grid too large");
    const int grid = (int)grid_ll;
    cudaStream_t st = as_stream(stream);
    switch (BNsel) {
        case 256: return launch_igemm<256, 4, 1>(a, grid, st);
        case 192: return launch_igemm<192, 4, 1>(a, grid, st);
        case 128: return launch_igemm<128, 3, 2>(a, grid, st);
        case 96: return launch_igemm<96, 3, 2>(a, grid, st);
        case 64: return launch_igemm<64, 4, 2>(a, grid, st);
        default: return launch_igemm<32, 4, 2>(a, grid, st);
    }
}

// Token-row kernels of the DiT path (reference models/transformer_layer.py:80-106, models/transformer.py:153-213):
// adaLN-modulated LayerNorm forward/backward, the adaLN-Zero gated residual (training form; inference fuses it into
// the igemm epilogue), ReLU backward, and the patch (un)folding at the model boundary.
//
// All of these are HBM-bound row streams: one warp owns one token row (C <= 1024 bf16 = at most four 16-byte chunks
// per lane, held in registers, so a row is read exactly once), statistics in fp32 with the two-pass variance the
// reference's nn.LayerNorm uses. Column sums needed by the backward (d shift, d scale, d gate) are accumulated in
// registers across the rows a CTA owns, combined through shared memory in a fixed order and written as partials
// [n, P, ...] that b200sd_reduce_f32 folds — deterministic, no atomics.
#include "host.h"
#include <cuda_bf16.h>

namespace b200 {

constexpr int ROW_WARPS = 8;          // warps (= rows in flight) per CTA
constexpr int ROW_MAX_CHUNKS = 4;     // 16-byte chunks per lane -> C <= 1024
constexpr int ROWS_PER_CTA = 32;      // rows of one image a CTA folds into one partial slot

__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float* f) {
    const uint4 u = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float2 v = __bfloat1622float2(h[j]);
        f[2 * j] = v.x, f[2 * j + 1] = v.y;
    }
}

__device__ __forceinline__ void st8(__nv_bfloat16* p, const float* f) {
    uint4 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
    *reinterpret_cast<uint4*>(p) = u;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// y = LN(x) * (1 + scale[img]) + shift[img]   (scale/shift NULL: plain LN without affine)
// LPR lanes share one row: 32 (C <= 1024) or 16 (C <= 512: two rows per warp in flight — the DiT's C = 288 is 36
// chunks, which left 28 lanes idle in the second pass of the one-row-per-warp form and one load in flight per lane).
template <int LPR>
__global__ void __launch_bounds__(ROW_WARPS * 32)
    ln_mod_kernel(const __nv_bfloat16* __restrict__ x, long long ld_x, __nv_bfloat16* __restrict__ y, long long ld_y,
                  const __nv_bfloat16* __restrict__ shift, const __nv_bfloat16* __restrict__ scale, long long ld_m,
                  long long rows, long long rows_per_img, int c, float eps, float* __restrict__ stats) {
    pdl_enter();
    constexpr int RPW = 32 / LPR;                     // rows per warp
    constexpr int NCH = LPR == 32 ? ROW_MAX_CHUNKS : 4;   // 16-byte chunks per lane (16 lanes x 4 x 8 = 512 channels)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sl = lane % LPR, sub = lane / LPR;
    const int nchunk = c >> 3;
    const float inv_c = 1.f / static_cast<float>(c);
    auto row_sum = [](float v) {
#pragma unroll
        for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        return v;
    };
    for (long long r0 = (static_cast<long long>(blockIdx.x) * ROW_WARPS + warp) * RPW; r0 < rows;
         r0 += static_cast<long long>(gridDim.x) * ROW_WARPS * RPW) {
        const long long r = r0 + sub;
        const bool live = r < rows;                   // (an odd row count leaves the last half-warp idle; it still shuffles)
        float v[NCH][8];
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
            const int ch = sl + LPR * k;
            if (live && ch < nchunk) {
                ld8(x + r * ld_x + ch * 8, v[k]);
#pragma unroll
                for (int j = 0; j < 8; ++j) s += v[k][j];
            }
        }
        const float mean = row_sum(s) * inv_c;
        float q = 0.f;
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
            if (live && sl + LPR * k < nchunk) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float d = v[k][j] - mean;
                    q += d * d;
                }
            }
        }
        const float rstd = rsqrtf(row_sum(q) * inv_c + eps);
        if (!live) continue;
        if (stats && sl == 0) stats[2 * r] = mean, stats[2 * r + 1] = rstd;
        const long long img = r / rows_per_img;
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
            const int ch = sl + LPR * k;
            if (ch < nchunk) {
                float o[8];
                if (scale) {
                    float sc[8], sh[8];
                    ld8(scale + img * ld_m + ch * 8, sc);
                    ld8(shift + img * ld_m + ch * 8, sh);
#pragma unroll
                    for (int j = 0; j < 8; ++j) o[j] = (v[k][j] - mean) * rstd * (1.f + sc[j]) + sh[j];
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) o[j] = (v[k][j] - mean) * rstd;
                }
                st8(y + r * ld_y + ch * 8, o);
            }
        }
    }
}

// Backward of the above. g = dy * (1 + scale); dx = rstd * (g - mean_c(g) - xhat * mean_c(g * xhat)) (+ dres).
// part[img][slot][0][c] = sum over the CTA's rows of dy, part[img][slot][1][c] = sum of dy * xhat.
__global__ void __launch_bounds__(ROW_WARPS * 32)
    ln_mod_bwd_kernel(const __nv_bfloat16* __restrict__ dy, long long ld_dy, const __nv_bfloat16* __restrict__ x,
                      long long ld_x, const float* __restrict__ stats, const __nv_bfloat16* __restrict__ scale,
                      long long ld_m, const __nv_bfloat16* __restrict__ dres, long long ld_r,
                      __nv_bfloat16* __restrict__ dx, long long ld_dx, long long rows_per_img, int c, int slots,
                      float* __restrict__ part) {
    pdl_enter();
    extern __shared__ float row_sm[];  // [ROW_WARPS][2][c]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nchunk = c >> 3;
    const float inv_c = 1.f / static_cast<float>(c);
    const long long img = blockIdx.y;
    const long long r_begin = static_cast<long long>(blockIdx.x) * ROWS_PER_CTA;
    long long r_end = r_begin + ROWS_PER_CTA;
    if (r_end > rows_per_img) r_end = rows_per_img;
    float a_dy[ROW_MAX_CHUNKS][8], a_dx[ROW_MAX_CHUNKS][8];
#pragma unroll
    for (int k = 0; k < ROW_MAX_CHUNKS; ++k)
#pragma unroll
        for (int j = 0; j < 8; ++j) a_dy[k][j] = a_dx[k][j] = 0.f;
    for (long long rr = r_begin + warp; rr < r_end; rr += ROW_WARPS) {
        const long long r = img * rows_per_img + rr;
        const float mean = stats[2 * r], rstd = stats[2 * r + 1];
        float g[ROW_MAX_CHUNKS][8], xh[ROW_MAX_CHUNKS][8];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < ROW_MAX_CHUNKS; ++k) {
            const int ch = lane + 32 * k;
            if (ch < nchunk) {
                float d[8], xv[8], sc[8];
                ld8(dy + r * ld_dy + ch * 8, d);
                ld8(x + r * ld_x + ch * 8, xv);
                if (scale) ld8(scale + img * ld_m + ch * 8, sc);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    xh[k][j] = (xv[j] - mean) * rstd;
                    g[k][j] = scale ? d[j] * (1.f + sc[j]) : d[j];
                    a_dy[k][j] += d[j];
                    a_dx[k][j] += d[j] * xh[k][j];
                    s1 += g[k][j];
                    s2 += g[k][j] * xh[k][j];
                }
            }
        }
        const float m1 = warp_sum(s1) * inv_c, m2 = warp_sum(s2) * inv_c;
#pragma unroll
        for (int k = 0; k < ROW_MAX_CHUNKS; ++k) {
            const int ch = lane + 32 * k;
            if (ch < nchunk) {
                float o[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] = rstd * (g[k][j] - m1 - xh[k][j] * m2);
                if (dres) {
                    float e[8];
                    ld8(dres + r * ld_r + ch * 8, e);
#pragma unroll
                    for (int j = 0; j < 8; ++j) o[j] += e[j];
                }
                st8(dx + r * ld_dx + ch * 8, o);
            }
        }
    }
    if (!part) return;
#pragma unroll
    for (int k = 0; k < ROW_MAX_CHUNKS; ++k) {
        const int ch = lane + 32 * k;
        if (ch < nchunk) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                row_sm[(warp * 2 + 0) * c + ch * 8 + j] = a_dy[k][j];
                row_sm[(warp * 2 + 1) * c + ch * 8 + j] = a_dx[k][j];
            }
        }
    }
    __syncthreads();
    float* out = part + (img * slots + blockIdx.x) * 2ll * c;
    for (int i = threadIdx.x; i < 2 * c; i += blockDim.x) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < ROW_WARPS; ++w) s += row_sm[w * 2 * c + i];
        out[i] = s;
    }
}

// y = x + gate[img] * h
__global__ void gate_residual_kernel(const __nv_bfloat16* __restrict__ x, long long ld_x,
                                     const __nv_bfloat16* __restrict__ h, long long ld_h,
                                     const __nv_bfloat16* __restrict__ gate, long long ld_g,
                                     __nv_bfloat16* __restrict__ y, long long ld_y, long long rows,
                                     long long rows_per_img, int c) {
    pdl_enter();
    const int nchunk = c >> 3;
    const long long total = rows * nchunk;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / nchunk;
        const int ch = static_cast<int>(i - r * nchunk);
        float a[8], b[8], g[8];
        ld8(x + r * ld_x + ch * 8, a);
        ld8(h + r * ld_h + ch * 8, b);
        ld8(gate + (r / rows_per_img) * ld_g + ch * 8, g);
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] += g[j] * b[j];
        st8(y + r * ld_y + ch * 8, a);
    }
}

// dh = gate[img] * dy; part[img][slot][c] = sum over the CTA's rows of dy * h
__global__ void __launch_bounds__(ROW_WARPS * 32)
    gate_residual_bwd_kernel(const __nv_bfloat16* __restrict__ dy, long long ld_dy,
                             const __nv_bfloat16* __restrict__ h, long long ld_h,
                             const __nv_bfloat16* __restrict__ gate, long long ld_g, __nv_bfloat16* __restrict__ dh,
                             long long ld_dh, long long rows_per_img, int c, int slots, float* __restrict__ part) {
    pdl_enter();
    extern __shared__ float row_sm[];  // [ROW_WARPS][c]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nchunk = c >> 3;
    const long long img = blockIdx.y;
    const long long r_begin = static_cast<long long>(blockIdx.x) * ROWS_PER_CTA;
    long long r_end = r_begin + ROWS_PER_CTA;
    if (r_end > rows_per_img) r_end = rows_per_img;
    float acc[ROW_MAX_CHUNKS][8];
#pragma unroll
    for (int k = 0; k < ROW_MAX_CHUNKS; ++k)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[k][j] = 0.f;
    for (long long rr = r_begin + warp; rr < r_end; rr += ROW_WARPS) {
        const long long r = img * rows_per_img + rr;
#pragma unroll
        for (int k = 0; k < ROW_MAX_CHUNKS; ++k) {
            const int ch = lane + 32 * k;
            if (ch < nchunk) {
                float d[8], hv[8], g[8];
                ld8(dy + r * ld_dy + ch * 8, d);
                ld8(h + r * ld_h + ch * 8, hv);
                ld8(gate + img * ld_g + ch * 8, g);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    acc[k][j] += d[j] * hv[j];
                    g[j] *= d[j];
                }
                st8(dh + r * ld_dh + ch * 8, g);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < ROW_MAX_CHUNKS; ++k) {
        const int ch = lane + 32 * k;
        if (ch < nchunk) {
#pragma unroll
            for (int j = 0; j < 8; ++j) row_sm[warp * c + ch * 8 + j] = acc[k][j];
        }
    }
    __syncthreads();
    float* out = part + (img * slots + blockIdx.x) * static_cast<long long>(c);
    for (int i = threadIdx.x; i < c; i += blockDim.x) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < ROW_WARPS; ++w) s += row_sm[w * c + i];
        out[i] = s;
    }
}

// dx = dy where y > 0 (y = relu(pre-activation) as saved by the forward)
__global__ void relu_bwd_kernel(const __nv_bfloat16* __restrict__ dy, long long ld_dy,
                                const __nv_bfloat16* __restrict__ y, long long ld_y, __nv_bfloat16* __restrict__ dx,
                                long long ld_dx, long long rows, int c) {
    pdl_enter();
    const int nchunk = c >> 3;
    const long long total = rows * nchunk;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / nchunk;
        const int ch = static_cast<int>(i - r * nchunk);
        float d[8], v[8];
        ld8(dy + r * ld_dy + ch * 8, d);
        ld8(y + r * ld_y + ch * 8, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) d[j] = v[j] > 0.f ? d[j] : 0.f;
        st8(dx + r * ld_dx + ch * 8, d);
    }
}

// tokens[(img, gy, gx)][(py, px, ch)] <-> image[img][ch][gy*ph+py][gx*pw+px]
// (einops 'b (nh nw) (ph pw c) <-> b c (nh ph) (nw pw)', transformer.py:207-211, patch_embed.py:84-86)
template <typename T>
__global__ void unpatchify_kernel(const T* __restrict__ tok, long long ld, float* __restrict__ img, long long n, int ch,
                                  int gh, int gw, int ph, int pw) {
    pdl_enter();
    const int H = gh * ph, W = gw * pw;
    const long long total = n * ch * H * W;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int xw = static_cast<int>(i % W);
        const int yh = static_cast<int>((i / W) % H);
        const int k = static_cast<int>((i / (static_cast<long long>(W) * H)) % ch);
        const long long b = i / (static_cast<long long>(W) * H * ch);
        const int gy = yh / ph, py = yh - gy * ph, gx = xw / pw, px = xw - gx * pw;
        img[i] = static_cast<float>(tok[((b * gh + gy) * gw + gx) * ld + (py * pw + px) * ch + k]);
    }
}

__global__ void patchify_kernel(const float* __restrict__ img, __nv_bfloat16* __restrict__ tok, long long ld,
                                long long n, int ch, int gh, int gw, int ph, int pw) {
    pdl_enter();
    const int H = gh * ph, W = gw * pw;
    const long long total = n * gh * gw * ld;
    const int width = ph * pw * ch;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int col = static_cast<int>(i % ld);
        const long long t = i / ld;
        float v = 0.f;
        if (col < width) {
            const int k = col % ch, pp = col / ch, px = pp % pw, py = pp / pw;
            const int gx = static_cast<int>(t % gw), gy = static_cast<int>((t / gw) % gh);
            const long long b = t / (static_cast<long long>(gw) * gh);
            v = img[((b * ch + k) * H + gy * ph + py) * W + gx * pw + px];
        }
        tok[i] = __float2bfloat16_rn(v);
    }
}

static int grid_for(long long items, int per_block) {
    long long b = (items + per_block - 1) / per_block;
    const long long cap = 16ll * sm_count();
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return static_cast<int>(b);
}

static bool row_ok(int c, long long ld) { return c > 0 && c % 8 == 0 && c <= 8 * 32 * ROW_MAX_CHUNKS && ld % 8 == 0 && ld >= c; }

}  // namespace b200

using namespace b200;

extern "C" int b200sd_rows_per_slot(void) { return ROWS_PER_CTA; }

extern "C" int b200sd_layernorm_mod(const void* x, int64_t ld_x, void* y, int64_t ld_y, const void* shift,
                                    const void* scale, int64_t ld_m, int64_t rows, int64_t rows_per_img, int32_t c,
                                    float eps, float* stats, void* stream) {
    B200_REQUIRE(x && y && rows >= 0 && rows_per_img > 0, "layernorm_mod: bad arguments");
    B200_REQUIRE(row_ok(c, ld_x) && row_ok(c, ld_y), "layernorm_mod: channels must be a multiple of 8 and <= 1024, pitches multiples of 8");
    B200_REQUIRE((shift == nullptr) == (scale == nullptr), "layernorm_mod: shift and scale come together");
    B200_REQUIRE(!scale || (ld_m % 8 == 0 && rows % rows_per_img == 0), "layernorm_mod: modulation pitch / rows_per_img");
    if (rows == 0) return B200SD_OK;
    if (c <= 512)
        launch_k(ln_mod_kernel<16>, grid_for((rows + 1) / 2, ROW_WARPS), ROW_WARPS * 32, 0, as_stream(stream),
                 static_cast<const __nv_bfloat16*>(x), ld_x, static_cast<__nv_bfloat16*>(y), ld_y,
                 static_cast<const __nv_bfloat16*>(shift), static_cast<const __nv_bfloat16*>(scale), ld_m, rows,
                 rows_per_img, c, eps, stats);
    else
        launch_k(ln_mod_kernel<32>, grid_for(rows, ROW_WARPS), ROW_WARPS * 32, 0, as_stream(stream),
                 static_cast<const __nv_bfloat16*>(x), ld_x, static_cast<__nv_bfloat16*>(y), ld_y,
                 static_cast<const __nv_bfloat16*>(shift), static_cast<const __nv_bfloat16*>(scale), ld_m, rows,
                 rows_per_img, c, eps, stats);
    B200_CHECK_LAUNCH("layernorm_mod");
    return B200SD_OK;
}

extern "C" int b200sd_layernorm_mod_bwd(const void* dy, int64_t ld_dy, const void* x, int64_t ld_x, const float* stats,
                                        const void* scale, int64_t ld_m, const void* dres, int64_t ld_r, void* dx,
                                        int64_t ld_dx, int64_t n, int64_t rows_per_img, int32_t c, float* part,
                                        void* stream) {
    B200_REQUIRE(dy && x && stats && dx && n >= 0 && rows_per_img > 0, "layernorm_mod_bwd: bad arguments");
    B200_REQUIRE(row_ok(c, ld_dy) && row_ok(c, ld_x) && row_ok(c, ld_dx), "layernorm_mod_bwd: channels / pitches");
    B200_REQUIRE(!dres || row_ok(c, ld_r), "layernorm_mod_bwd: residual-gradient pitch");
    B200_REQUIRE(!scale || ld_m % 8 == 0, "layernorm_mod_bwd: modulation pitch");
    if (n == 0) return B200SD_OK;
    const int slots = static_cast<int>((rows_per_img + ROWS_PER_CTA - 1) / ROWS_PER_CTA);
    const size_t smem = sizeof(float) * ROW_WARPS * 2 * c;
    static bool attr = false;
    if (!attr) {
        cudaFuncSetAttribute(ln_mod_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 2 * 1024 * 4);
        attr = true;
    }
    launch_k(ln_mod_bwd_kernel, dim3(slots, static_cast<unsigned>(n)), ROW_WARPS * 32, smem, as_stream(stream), 
        static_cast<const __nv_bfloat16*>(dy), ld_dy, static_cast<const __nv_bfloat16*>(x), ld_x, stats,
        static_cast<const __nv_bfloat16*>(scale), ld_m, static_cast<const __nv_bfloat16*>(dres), ld_r,
        static_cast<__nv_bfloat16*>(dx), ld_dx, rows_per_img, c, slots, part);
    B200_CHECK_LAUNCH("layernorm_mod_bwd");
    return B200SD_OK;
}

extern "C" int b200sd_gate_residual(const void* x, int64_t ld_x, const void* h, int64_t ld_h, const void* gate,
                                    int64_t ld_g, void* y, int64_t ld_y, int64_t rows, int64_t rows_per_img, int32_t c,
                                    void* stream) {
    B200_REQUIRE(x && h && gate && y && rows >= 0 && rows_per_img > 0, "gate_residual: bad arguments");
    B200_REQUIRE(row_ok(c, ld_x) && row_ok(c, ld_h) && row_ok(c, ld_y) && ld_g % 8 == 0, "gate_residual: channels / pitches");
    if (rows == 0) return B200SD_OK;
    launch_k(gate_residual_kernel, grid_for(rows * (c / 8), 256), 256, 0, as_stream(stream), 
        static_cast<const __nv_bfloat16*>(x), ld_x, static_cast<const __nv_bfloat16*>(h), ld_h,
        static_cast<const __nv_bfloat16*>(gate), ld_g, static_cast<__nv_bfloat16*>(y), ld_y, rows, rows_per_img, c);
    B200_CHECK_LAUNCH("gate_residual");
    return B200SD_OK;
}

extern "C" int b200sd_gate_residual_bwd(const void* dy, int64_t ld_dy, const void* h, int64_t ld_h, const void* gate,
                                        int64_t ld_g, void* dh, int64_t ld_dh, int64_t n, int64_t rows_per_img,
                                        int32_t c, float* part, void* stream) {
    B200_REQUIRE(dy && h && gate && dh && part && n >= 0 && rows_per_img > 0, "gate_residual_bwd: bad arguments");
    B200_REQUIRE(row_ok(c, ld_dy) && row_ok(c, ld_h) && row_ok(c, ld_dh) && ld_g % 8 == 0, "gate_residual_bwd: channels / pitches");
    if (n == 0) return B200SD_OK;
    const int slots = static_cast<int>((rows_per_img + ROWS_PER_CTA - 1) / ROWS_PER_CTA);
    launch_k(gate_residual_bwd_kernel, dim3(slots, static_cast<unsigned>(n)), ROW_WARPS * 32, sizeof(float) * ROW_WARPS * c, as_stream(stream), 
        static_cast<const __nv_bfloat16*>(dy), ld_dy, static_cast<const __nv_bfloat16*>(h), ld_h,
        static_cast<const __nv_bfloat16*>(gate), ld_g, static_cast<__nv_bfloat16*>(dh), ld_dh, rows_per_img, c, slots,
        part);
    B200_CHECK_LAUNCH("gate_residual_bwd");
    return B200SD_OK;
}

extern "C" int b200sd_relu_bwd(const void* dy, int64_t ld_dy, const void* y, int64_t ld_y, void* dx, int64_t ld_dx,
                               int64_t rows, int32_t c, void* stream) {
    B200_REQUIRE(dy && y && dx && rows >= 0 && c > 0 && c % 8 == 0, "relu_bwd: bad arguments");
    B200_REQUIRE(ld_dy % 8 == 0 && ld_y % 8 == 0 && ld_dx % 8 == 0, "relu_bwd: pitches must be multiples of 8");
    if (rows == 0) return B200SD_OK;
    launch_k(relu_bwd_kernel, grid_for(rows * (c / 8), 256), 256, 0, as_stream(stream), 
        static_cast<const __nv_bfloat16*>(dy), ld_dy, static_cast<const __nv_bfloat16*>(y), ld_y,
        static_cast<__nv_bfloat16*>(dx), ld_dx, rows, c);
    B200_CHECK_LAUNCH("relu_bwd");
    return B200SD_OK;
}

extern "C" int b200sd_unpatchify(const void* tokens, int32_t tokens_f32, int64_t ld, float* image, int64_t n,
                                 int32_t channels, int32_t grid_h, int32_t grid_w, int32_t patch_h, int32_t patch_w,
                                 void* stream) {
    B200_REQUIRE(tokens && image && n >= 0 && channels > 0 && grid_h > 0 && grid_w > 0 && patch_h > 0 && patch_w > 0,
                 "unpatchify: bad arguments");
    B200_REQUIRE(ld >= static_cast<int64_t>(patch_h) * patch_w * channels, "unpatchify: token pitch too small");
    const long long total = n * channels * grid_h * patch_h * static_cast<long long>(grid_w) * patch_w;
    if (total == 0) return B200SD_OK;
    if (tokens_f32)
        launch_k(unpatchify_kernel<float>, grid_for(total, 256), 256, 0, as_stream(stream), 
            static_cast<const float*>(tokens), ld, image, n, channels, grid_h, grid_w, patch_h, patch_w);
    else
        launch_k(unpatchify_kernel<__nv_bfloat16>, grid_for(total, 256), 256, 0, as_stream(stream), 
            static_cast<const __nv_bfloat16*>(tokens), ld, image, n, channels, grid_h, grid_w, patch_h, patch_w);
    B200_CHECK_LAUNCH("unpatchify");
    return B200SD_OK;
}

extern "C" int b200sd_patchify(const float* image, void* tokens, int64_t ld, int64_t n, int32_t channels,
                               int32_t grid_h, int32_t grid_w, int32_t patch_h, int32_t patch_w, void* stream) {
    B200_REQUIRE(tokens && image && n >= 0 && channels > 0 && grid_h > 0 && grid_w > 0 && patch_h > 0 && patch_w > 0,
                 "patchify: bad arguments");
    B200_REQUIRE(ld >= static_cast<int64_t>(patch_h) * patch_w * channels, "patchify: token pitch too small");
    const long long total = n * grid_h * static_cast<long long>(grid_w) * ld;
    if (total == 0) return B200SD_OK;
    launch_k(patchify_kernel, grid_for(total, 256), 256, 0, as_stream(stream), 
        image, static_cast<__nv_bfloat16*>(tokens), ld, n, channels, grid_h, grid_w, patch_h, patch_w);
    B200_CHECK_LAUNCH("patchify");
    return B200SD_OK;
}

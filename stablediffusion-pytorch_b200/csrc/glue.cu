// Layout and conditioning glue (SURVEY K14 + boundary conversions). All HBM-bound, coalesced on the side that
// dominates traffic; tiny next to the GEMMs but each replaces several ATen launches in the reference.
#include "host.h"
#include "ptx.cuh"
#include <cuda_bf16.h>

namespace b200 {

__device__ __forceinline__ float glue_act(float v, int act) {
    if (act == B200SD_ACT_SILU) return silu_mufu(v);
    if (act == B200SD_ACT_RELU) return fmaxf(v, 0.f);
    return v;
}

// src NCHW fp32 -> dst NHWC bf16 slice. Thread per (n, p): reads are coalesced over p for each c.
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long ld,
                                    int c_off, long long n, int c, long long hw) {
    pdl_enter();
    const long long total = n * hw;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long img = i / hw, p = i - img * hw;
        const float* s = src + img * c * hw + p;
        __nv_bfloat16* d = dst + i * ld + c_off;
        for (int k = 0; k < c; ++k) d[k] = __float2bfloat16_rn(s[(long long)k * hw]);
    }
}

template <typename T>
__global__ void nhwc_to_nchw_kernel(const T* __restrict__ src, long long ld, float* __restrict__ dst, long long n,
                                    int c, long long hw) {
    pdl_enter();
    const long long total = n * hw;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long img = i / hw, p = i - img * hw;
        const T* s = src + i * ld;
        float* d = dst + img * c * hw + p;
        for (int k = 0; k < c; ++k) d[(long long)k * hw] = static_cast<float>(s[k]);
    }
}

// Mask path: nearest resize + 1x1 conv (no bias) + write into the conv_in input slice. Thread per output pixel.
// F.interpolate(mode='nearest') picks src = floor(dst * scale) with scale = in/out computed in fp32.
__global__ void mask_cond_kernel(const float* __restrict__ mask, const float* __restrict__ w,
                                 __nv_bfloat16* __restrict__ dst, long long ld, int c_off, long long n, int cin,
                                 int cout, int mh, int mw, int h, int wd) {
    pdl_enter();
    extern __shared__ float sw[];  // [cout, cin]
    for (int i = threadIdx.x; i < cout * cin; i += blockDim.x) sw[i] = w[i];
    __syncthreads();
    const float sy = (float)mh / (float)h, sx = (float)mw / (float)wd;
    const long long total = n * h * wd;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int x = (int)(i % wd);
        const int yy = (int)((i / wd) % h);
        const long long img = i / ((long long)wd * h);
        const int srcy = min((int)floorf(yy * sy), mh - 1);
        const int srcx = min((int)floorf(x * sx), mw - 1);
        const float* m = mask + (img * cin * mh + srcy) * (long long)mw + srcx;
        float acc[8];
        for (int o = 0; o < cout; ++o) acc[o] = 0.f;
        for (int k = 0; k < cin; ++k) {
            // autocast semantics: the conv sees bf16 inputs and weights, accumulates fp32
            const float mv = __bfloat162float(__float2bfloat16_rn(m[(long long)k * mh * mw]));
            for (int o = 0; o < cout; ++o)
                acc[o] += mv * __bfloat162float(__float2bfloat16_rn(sw[o * cin + k]));
        }
        __nv_bfloat16* d = dst + i * ld + c_off;
        for (int o = 0; o < cout; ++o) d[o] = __float2bfloat16_rn(acc[o]);
    }
}

__global__ void copy_act_kernel(const __nv_bfloat16* __restrict__ src, long long ld_src,
                                __nv_bfloat16* __restrict__ dst, long long ld_dst, int c_off, long long rows, int c,
                                int act, int vec) {
    pdl_enter();
    if (vec) {  // 8 channels (16 B) per thread
        const int vpr = c >> 3;
        const long long total = rows * vpr;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
             i += (long long)gridDim.x * blockDim.x) {
            const long long r = i / vpr;
            const int v = (int)(i - r * vpr);
            uint4 u = *reinterpret_cast<const uint4*>(src + r * ld_src + v * 8);
            if (act != B200SD_ACT_NONE) {
                __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    float2 f = __bfloat1622float2(h[j]);
                    h[j] = __floats2bfloat162_rn(glue_act(f.x, act), glue_act(f.y, act));
                }
            }
            *reinterpret_cast<uint4*>(dst + r * ld_dst + c_off + v * 8) = u;
        }
    } else {
        const long long total = rows * c;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
             i += (long long)gridDim.x * blockDim.x) {
            const long long r = i / c;
            const int k = (int)(i - r * c);
            dst[r * ld_dst + c_off + k] =
                __float2bfloat16_rn(glue_act(__bfloat162float(src[r * ld_src + k]), act));
        }
    }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n4,
                                     long long n) {
    pdl_enter();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
         i += (long long)gridDim.x * blockDim.x) {
        float4 f = reinterpret_cast<const float4*>(src)[i];
        uint2 u;
        *reinterpret_cast<__nv_bfloat162*>(&u.x) = __floats2bfloat162_rn(f.x, f.y);
        *reinterpret_cast<__nv_bfloat162*>(&u.y) = __floats2bfloat162_rn(f.z, f.w);
        reinterpret_cast<uint2*>(dst)[i] = u;
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        const long long j = (n & ~3ll) + threadIdx.x;
        dst[j] = __float2bfloat16_rn(src[j]);
    }
}

__global__ void add_bf16_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, uint4* __restrict__ o,
                                long long n8) {
    pdl_enter();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8;
         i += (long long)gridDim.x * blockDim.x) {
        uint4 x = a[i], y = b[i], r;
        const __nv_bfloat162* hx = reinterpret_cast<const __nv_bfloat162*>(&x);
        const __nv_bfloat162* hy = reinterpret_cast<const __nv_bfloat162*>(&y);
        __nv_bfloat162* hr = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float2 fx = __bfloat1622float2(hx[j]), fy = __bfloat1622float2(hy[j]);
            hr[j] = __floats2bfloat162_rn(fx.x + fy.x, fx.y + fy.y);
        }
        o[i] = r;
    }
}

static inline int glue_grid(long long items, int threads) {
    long long b = (items + threads - 1) / threads;
    long long cap = (long long)sm_count() * 8;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

// nn.Unfold of an NHWC bf16 tensor into GEMM rows in the (c, ky, kx) column order of weight.view(Cout, -1) — the operand
// layout of the compute-in-memory convolutions (cim_layers/layers_qn_lsq_adda_cim.py:92-94,198: the mapped weight blocks
// are row ranges of that K axis, so the explicit matrix is what the array sees). Thread = (output pixel, channel):
// reads one input channel's k*k window (border taps zero), writes k*k consecutive columns.
__global__ void im2col_kernel(const __nv_bfloat16* __restrict__ x, long long ld_x, __nv_bfloat16* __restrict__ rows,
                              long long ld_r, long long n, int h, int w, int c, int oh, int ow, int k, int stride,
                              int pad) {
    pdl_enter();
    const long long total = n * oh * ow * c;
    const int kk = k * k;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int ch = static_cast<int>(i % c);
        const long long p = i / c;
        const int ox = static_cast<int>(p % ow), oy = static_cast<int>((p / ow) % oh);
        const long long img = p / (static_cast<long long>(ow) * oh);
        __nv_bfloat16* dst = rows + p * ld_r + static_cast<long long>(ch) * kk;
        for (int ky = 0; ky < k; ++ky) {
            const int iy = oy * stride - pad + ky;
            for (int kx = 0; kx < k; ++kx) {
                const int ix = ox * stride - pad + kx;
                __nv_bfloat16 v = __float2bfloat16(0.f);
                if (iy >= 0 && iy < h && ix >= 0 && ix < w) v = x[((img * h + iy) * w + ix) * ld_x + ch];
                dst[ky * k + kx] = v;
            }
        }
    }
}

// Its adjoint (nn.Fold): dx[n, y, x, c] = sum over the (ky, kx) whose window covers (y, x) of drows[(n, oy, ox),
// c*k*k + ky*k + kx], gathered per input element in fp32 (fixed order: deterministic, no atomics).
__global__ void col2im_kernel(const __nv_bfloat16* __restrict__ drows, long long ld_r, __nv_bfloat16* __restrict__ dx,
                              long long ld_x, long long n, int h, int w, int c, int oh, int ow, int k, int stride,
                              int pad) {
    pdl_enter();
    const long long total = n * h * w * c;
    const int kk = k * k;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int ch = static_cast<int>(i % c);
        const long long p = i / c;
        const int ix = static_cast<int>(p % w), iy = static_cast<int>((p / w) % h);
        const long long img = p / (static_cast<long long>(w) * h);
        float acc = 0.f;
        for (int ky = 0; ky < k; ++ky) {
            const int ty = iy + pad - ky;
            if (ty < 0 || ty % stride) continue;
            const int oy = ty / stride;
            if (oy >= oh) continue;
            for (int kx = 0; kx < k; ++kx) {
                const int tx = ix + pad - kx;
                if (tx < 0 || tx % stride) continue;
                const int ox = tx / stride;
                if (ox >= ow) continue;
                acc += __bfloat162float(drows[((img * oh + oy) * ow + ox) * ld_r + static_cast<long long>(ch) * kk + ky * k + kx]);
            }
        }
        dx[p * ld_x + ch] = __float2bfloat16_rn(acc);
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200sd_nchw_to_nhwc(const float* src, void* dst, int64_t ld_dst, int c_off, int64_t n, int c,
                                   int64_t hw, void* stream) {
    B200_REQUIRE(src && dst, "nchw_to_nhwc: null pointer");
    B200_REQUIRE(c_off >= 0 && c_off + c <= ld_dst, "nchw_to_nhwc: channel slice [%d,%d) exceeds ld %lld", c_off,
                 c_off + c, (long long)ld_dst);
    if (n * hw == 0) return B200SD_OK;
    launch_k(nchw_to_nhwc_kernel, glue_grid(n * hw, 256), 256, 0, as_stream(stream), 
        src, static_cast<__nv_bfloat16*>(dst), ld_dst, c_off, n, c, hw);
    B200_CHECK_LAUNCH("nchw_to_nhwc");
    return B200SD_OK;
}

extern "C" int b200sd_nhwc_to_nchw(const void* src, int src_dtype, int64_t ld_src, float* dst, int64_t n, int c,
                                   int64_t hw, void* stream) {
    B200_REQUIRE(src && dst, "nhwc_to_nchw: null pointer");
    B200_REQUIRE(c <= ld_src, "nhwc_to_nchw: c=%d exceeds ld %lld", c, (long long)ld_src);
    if (n * hw == 0) return B200SD_OK;
    const int grid = glue_grid(n * hw, 256);
    if (src_dtype == B200SD_DT_BF16)
        launch_k(nhwc_to_nchw_kernel<__nv_bfloat16>, grid, 256, 0, as_stream(stream), static_cast<const __nv_bfloat16*>(src), ld_src, dst, n, c, hw);
    else if (src_dtype == B200SD_DT_F32)
        launch_k(nhwc_to_nchw_kernel<float>, grid, 256, 0, as_stream(stream), static_cast<const float*>(src), ld_src, dst, n, c, hw);
    else {
        set_error("nhwc_to_nchw: bad dtype %d", src_dtype);
        return B200SD_ERR_ARG;
    }
    B200_CHECK_LAUNCH("nhwc_to_nchw");
    return B200SD_OK;
}

extern "C" int b200sd_mask_cond(const float* mask, const float* w, void* dst, int64_t ld_dst, int c_off, int64_t n,
                                int cin, int cout, int mh, int mw, int h, int wd, void* stream) {
    B200_REQUIRE(mask && w && dst, "mask_cond: null pointer");
    B200_REQUIRE(cout >= 1 && cout <= 8, "mask_cond: %d output channels (max 8)", cout);
    B200_REQUIRE(c_off >= 0 && c_off + cout <= ld_dst, "mask_cond: channel slice exceeds ld");
    B200_REQUIRE(cin >= 1 && cin * cout * 4 <= 48 * 1024, "mask_cond: weight too large for shared memory");
    if (n == 0) return B200SD_OK;
    launch_k(mask_cond_kernel, glue_grid(n * h * wd, 128), 128, cin * cout * sizeof(float), as_stream(stream), 
        mask, w, static_cast<__nv_bfloat16*>(dst), ld_dst, c_off, n, cin, cout, mh, mw, h, wd);
    B200_CHECK_LAUNCH("mask_cond");
    return B200SD_OK;
}

extern "C" int b200sd_copy_act(const void* src, int64_t ld_src, void* dst, int64_t ld_dst, int c_off, int64_t rows,
                               int c, int act, void* stream) {
    B200_REQUIRE(src && dst, "copy_act: null pointer");
    B200_REQUIRE(c <= ld_src && c_off >= 0 && c_off + c <= ld_dst, "copy_act: slice exceeds leading dimension");
    if (rows * c == 0) return B200SD_OK;
    const int vec = (c % 8 == 0) && (ld_src % 8 == 0) && (ld_dst % 8 == 0) && (c_off % 8 == 0) &&
                    ((reinterpret_cast<uintptr_t>(src) & 15) == 0) && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0);
    const long long items = vec ? rows * (c / 8) : rows * c;
    launch_k(copy_act_kernel, glue_grid(items, 256), 256, 0, as_stream(stream), 
        static_cast<const __nv_bfloat16*>(src), ld_src, static_cast<__nv_bfloat16*>(dst), ld_dst, c_off, rows, c, act,
        vec);
    B200_CHECK_LAUNCH("copy_act");
    return B200SD_OK;
}

extern "C" int b200sd_cast_f32_bf16(const float* src, void* dst, int64_t n, void* stream) {
    B200_REQUIRE(src && dst, "cast: null pointer");
    B200_REQUIRE(((reinterpret_cast<uintptr_t>(src) & 15) == 0) && ((reinterpret_cast<uintptr_t>(dst) & 7) == 0),
                 "cast: pointers unaligned");
    if (n == 0) return B200SD_OK;
    launch_k(cast_f32_bf16_kernel, glue_grid((n + 3) / 4, 256), 256, 0, as_stream(stream), 
        src, static_cast<__nv_bfloat16*>(dst), n / 4, n);
    B200_CHECK_LAUNCH("cast_f32_bf16");
    return B200SD_OK;
}

extern "C" int b200sd_add_bf16(const void* a, const void* b, void* out, int64_t n, void* stream) {
    B200_REQUIRE(a && b && out, "add: null pointer");
    B200_REQUIRE(n % 8 == 0, "add: element count %lld is not a multiple of 8", (long long)n);
    B200_REQUIRE(((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(out)) & 15) == 0,
                 "add: pointers unaligned");
    if (n == 0) return B200SD_OK;
    launch_k(add_bf16_kernel, glue_grid(n / 8, 256), 256, 0, as_stream(stream), 
        static_cast<const uint4*>(a), static_cast<const uint4*>(b), static_cast<uint4*>(out), n / 8);
    B200_CHECK_LAUNCH("add_bf16");
    return B200SD_OK;
}

extern "C" int b200sd_im2col(const void* x, int64_t ld_x, void* rows, int64_t ld_r, int64_t n, int h, int w, int c, int k,
                             int stride, int pad, void* stream) {
    B200_REQUIRE(x && rows, "im2col: null pointer");
    B200_REQUIRE(k >= 1 && stride >= 1 && pad >= 0 && h + 2 * pad >= k && w + 2 * pad >= k, "im2col: bad geometry");
    B200_REQUIRE(ld_x >= c && ld_r >= (int64_t)c * k * k, "im2col: pitches too small");
    const int oh = (h + 2 * pad - k) / stride + 1, ow = (w + 2 * pad - k) / stride + 1;
    if (n * oh * ow * c == 0) return B200SD_OK;
    launch_k(im2col_kernel, glue_grid(n * oh * ow * c, 256), 256, 0, as_stream(stream), static_cast<const __nv_bfloat16*>(x),
             ld_x, static_cast<__nv_bfloat16*>(rows), ld_r, n, h, w, c, oh, ow, k, stride, pad);
    B200_CHECK_LAUNCH("im2col");
    return B200SD_OK;
}

extern "C" int b200sd_col2im(const void* drows, int64_t ld_r, void* dx, int64_t ld_x, int64_t n, int h, int w, int c,
                             int k, int stride, int pad, void* stream) {
    B200_REQUIRE(drows && dx, "col2im: null pointer");
    B200_REQUIRE(k >= 1 && stride >= 1 && pad >= 0 && h + 2 * pad >= k && w + 2 * pad >= k, "col2im: bad geometry");
    B200_REQUIRE(ld_x >= c && ld_r >= (int64_t)c * k * k, "col2im: pitches too small");
    const int oh = (h + 2 * pad - k) / stride + 1, ow = (w + 2 * pad - k) / stride + 1;
    if (n * h * w * c == 0) return B200SD_OK;
    launch_k(col2im_kernel, glue_grid(n * h * w * c, 256), 256, 0, as_stream(stream), static_cast<const __nv_bfloat16*>(drows),
             ld_r, static_cast<__nv_bfloat16*>(dx), ld_x, n, h, w, c, oh, ow, k, stride, pad);
    B200_CHECK_LAUNCH("col2im");
    return B200SD_OK;
}

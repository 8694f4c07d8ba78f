// Train-side preprocess chain, fused: cubic resize (OpenCV arithmetic) -> BGR2GRAY -> *1/255.
//
// Replaces, for a batch of decoded uploads resident in HBM:
//   cv2.resize(im1,(28,28),interpolation=cv2.INTER_CUBIC)      apps/preprocess/preprocess.py:29
//   gray conversion of cv2.imread(dir,0)                        apps/preprocess/preprocess.py:40
//   images.astype(float32) * (1/255)      apps/construction/util/construct_distribute.py:503-504
//
// One CTA per picture.  A non-antialiased cubic resize touches only 4x4 source pixels per
// output pixel (28*28*16*C bytes of a 2.36 MB upload), so the kernel is a gather: it is bound by
// DRAM sectors and load issue, not by streaming bandwidth.  Stage 1 computes the horizontal pass
// for the 112 (dy,tap) source rows into shared memory as exact int32; stage 2 the vertical pass
// (float32 for the first 80 interleaved elements of a row -- OpenCV's 128-bit SIMD body -- and
// integer for the tail, see oracle/pp_restate.py:cv_resize_cubic_u8); stage 3 gray + normalise.
#include "nc_internal.h"
#include <stdlib.h>
#include <string.h>

namespace {

struct TrainTabs {
    CvCubicAxis x, y;
};
struct IppTabs {          // mode 1: float32 coefficients of the IPP arithmetic
    IppCubicAxis x, y;
};

constexpr int kThreads = 256;

// IPP = true: the arithmetic of the wheel's default cv2.resize (Intel IPP): float32 coefficients that are never quantised,
// float32 products and sums tap by tap in both passes (horizontal first), round half to even, saturate.
template <int C, bool IPP>
__global__ void __launch_bounds__(kThreads)
train_chain_kernel(const uint8_t* __restrict__ src, long long img_stride, int h, int w,
                   const __grid_constant__ TrainTabs tabs, uint8_t* __restrict__ out_u8,
                   float* __restrict__ out_f32, int identity, uint8_t* __restrict__ out_color, const IppTabs* __restrict__ ft) {
    constexpr int E = 28 * C;               // interleaved elements per output row
    constexpr int NSIMD = (E / 8) * 8;      // OpenCV v_int16x8 body
    __shared__ int hp[112][E];
    __shared__ uint8_t col[28][E];
    __shared__ int s_xofs[28][4];
    __shared__ int s_xcoef[28][4];
    __shared__ int s_yofs[28][4];
    __shared__ int s_ycoef[28][4];

    const int tid = threadIdx.x;
    const uint8_t* base = src + (long long)blockIdx.x * img_stride;

    if (!identity) {
        for (int i = tid; i < 112; i += kThreads) {
            if (IPP) {
                s_xofs[i >> 2][i & 3] = ft->x.ofs[i >> 2][i & 3] * C;
                s_xcoef[i >> 2][i & 3] = __float_as_int(ft->x.coef[i >> 2][i & 3]);
                s_yofs[i >> 2][i & 3] = ft->y.ofs[i >> 2][i & 3];
                s_ycoef[i >> 2][i & 3] = __float_as_int(ft->y.coef[i >> 2][i & 3]);
            } else {
                s_xofs[i >> 2][i & 3] = tabs.x.ofs[i >> 2][i & 3] * C;
                s_xcoef[i >> 2][i & 3] = tabs.x.coef[i >> 2][i & 3];
                s_yofs[i >> 2][i & 3] = tabs.y.ofs[i >> 2][i & 3];
                s_ycoef[i >> 2][i & 3] = tabs.y.coef[i >> 2][i & 3];
            }
        }
        __syncthreads();

        // ---- stage 1: horizontal pass, HResizeCubic<uchar,int,short>
        const long long row_bytes = (long long)w * C;
        for (int idx = tid; idx < 112 * 28; idx += kThreads) {
            const int slot = idx / 28, dx = idx - slot * 28;
            const uint8_t* rp = base + (long long)s_yofs[slot >> 2][slot & 3] * row_bytes;
            int acc[C];
#pragma unroll
            for (int c = 0; c < C; ++c) acc[c] = 0;
            unsigned char px[4][C];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const uint8_t* p = rp + s_xofs[dx][t];
#pragma unroll
                for (int c = 0; c < C; ++c) px[t][c] = __ldg(p + c);
            }
            if (IPP) {
                float fa[C];
#pragma unroll
                for (int c = 0; c < C; ++c) fa[c] = 0.0f;
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const float a = __int_as_float(s_xcoef[dx][t]);
#pragma unroll
                    for (int c = 0; c < C; ++c) fa[c] = __fadd_rn(fa[c], __fmul_rn((float)px[t][c], a));
                }
#pragma unroll
                for (int c = 0; c < C; ++c) hp[slot][dx * C + c] = __float_as_int(fa[c]);
            } else {
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const int a = s_xcoef[dx][t];
#pragma unroll
                    for (int c = 0; c < C; ++c) acc[c] += (int)px[t][c] * a;
                }
#pragma unroll
                for (int c = 0; c < C; ++c) hp[slot][dx * C + c] = acc[c];
            }
        }
        __syncthreads();

        // ---- stage 2: vertical pass, VResizeCubic<uchar,int,short,FixedPtCast<22>,VResizeCubicVec_32s8u>
        const float scale = 1.0f / (2048.0f * 2048.0f);
        for (int idx = tid; idx < 28 * E; idx += kThreads) {
            const int dy = idx / E, e = idx - dy * E;
            const int S0 = hp[dy * 4 + 0][e], S1 = hp[dy * 4 + 1][e];
            const int S2 = hp[dy * 4 + 2][e], S3 = hp[dy * 4 + 3][e];
            const int b0 = s_ycoef[dy][0], b1 = s_ycoef[dy][1], b2 = s_ycoef[dy][2], b3 = s_ycoef[dy][3];
            int v;
            if (IPP) {
                float t = __fmul_rn(__int_as_float(S0), __int_as_float(b0));        // (0 + S0 b0) + S1 b1 ...: tap order
                t = __fadd_rn(t, __fmul_rn(__int_as_float(S1), __int_as_float(b1)));
                t = __fadd_rn(t, __fmul_rn(__int_as_float(S2), __int_as_float(b2)));
                t = __fadd_rn(t, __fmul_rn(__int_as_float(S3), __int_as_float(b3)));
                v = __float2int_rn(fminf(fmaxf(t, -1.0f), 256.0f));
            } else if (e < NSIMD) {
                // v_round(f(S0)*b0 + (f(S1)*b1 + (f(S2)*b2 + f(S3)*b3))), separate mul / add
                const float f0 = __fmul_rn((float)b0, scale), f1 = __fmul_rn((float)b1, scale);
                const float f2 = __fmul_rn((float)b2, scale), f3 = __fmul_rn((float)b3, scale);
                float t = __fmul_rn(__int2float_rn(S3), f3);
                t = __fadd_rn(__fmul_rn(__int2float_rn(S2), f2), t);
                t = __fadd_rn(__fmul_rn(__int2float_rn(S1), f1), t);
                t = __fadd_rn(__fmul_rn(__int2float_rn(S0), f0), t);
                v = __float2int_rn(t);
            } else {
                // 32-bit wrap-around arithmetic exactly as the C code (int)
                const int sum = S0 * b0 + S1 * b1 + S2 * b2 + S3 * b3;
                v = (sum + (1 << 21)) >> 22;
            }
            v = v < 0 ? 0 : (v > 255 ? 255 : v);
            col[dy][e] = (uint8_t)v;
        }
        __syncthreads();
    }

    if (out_color) {   // the resized picture itself (what preprocess.resize writes back, preprocess.py:29-30)
        for (int idx = tid; idx < 28 * E; idx += kThreads)
            out_color[(long long)blockIdx.x * 28 * E + idx] = identity ? __ldg(base + idx) : (&col[0][0])[idx];
        if (!out_u8 && !out_f32) return;
    }
    // ---- stage 3: BGR2GRAY (cv2 4.13: 15-bit coefficients) and float32 normalise
    const float inv255 = (float)(1.0 / 255.0);
    for (int idx = tid; idx < 784; idx += kThreads) {
        int g;
        if (C == 3) {
            int b, gg, r;
            if (identity) {
                const uint8_t* p = base + idx * 3;
                b = __ldg(p); gg = __ldg(p + 1); r = __ldg(p + 2);
            } else {
                const uint8_t* p = &col[0][0] + idx * 3;
                b = p[0]; gg = p[1]; r = p[2];
            }
            g = (3735 * b + 19235 * gg + 9798 * r + (1 << 14)) >> 15;
        } else {
            g = identity ? (int)__ldg(base + idx) : (int)(&col[0][0])[idx];
        }
        if (out_u8) out_u8[(long long)blockIdx.x * 784 + idx] = (uint8_t)g;
        if (out_f32) out_f32[(long long)blockIdx.x * 784 + idx] = __fmul_rn((float)g, inv255);
    }
}


// ---------------------------------------------------------------------------------------------
// Row-streaming variant.  At 1024 -> 28 the 4-tap groups of a source row are 110 B apart, so DRAM
// (64 B bursts) already moves the whole 112 source rows of a picture (measured: 344 KB / picture =
// 112 x 3072 B).  This kernel therefore streams exactly those rows: each warp owns 4 of the 28
// output rows; for an output row it bulk-copies the <= 4 consecutive source rows (one
// cp.async.bulk, 12 KB at 1024 x 3) into its private shared-memory stage, waits on its mbarrier,
// and each lane computes one output pixel (16 taps x 3 channels, vertical pass, gray) from shared
// memory.  Same arithmetic, same bytes out; DRAM pages are read sequentially and there is no
// per-byte global load left.  Works on device memory and on pinned host memory (zero copy).
// ---------------------------------------------------------------------------------------------
constexpr int kRowWarps = 7;                 // 28 output rows / 7 warps = 4 rows per warp
constexpr int kRowThreads = kRowWarps * 32;

__global__ void __launch_bounds__(kRowThreads)
train_chain_rows_kernel(const uint8_t* __restrict__ src, long long img_stride, int h, int w,
                        const __grid_constant__ TrainTabs tabs, uint8_t* __restrict__ out_u8,
                        float* __restrict__ out_f32, unsigned stage_bytes) {
    constexpr int C = 3, E = 28 * C, NSIMD = (E / 8) * 8;
    extern __shared__ __align__(128) unsigned char stage_mem[];
    __shared__ int s_xofs[28][4];
    __shared__ int s_xcoef[28][4];
    __shared__ __align__(8) unsigned long long s_bar[kRowWarps];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < 112; i += kRowThreads) {
        s_xofs[i >> 2][i & 3] = tabs.x.ofs[i >> 2][i & 3] * C;
        s_xcoef[i >> 2][i & 3] = tabs.x.coef[i >> 2][i & 3];
    }
    const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_bar[warp]);
    unsigned char* stage = stage_mem + (size_t)warp * stage_bytes;
    const unsigned stage_addr = (unsigned)__cvta_generic_to_shared(stage);
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const uint8_t* base = src + (long long)blockIdx.x * img_stride;
    const unsigned row_bytes = (unsigned)w * C;
    const float inv255 = (float)(1.0 / 255.0);
    const float scale = 1.0f / (2048.0f * 2048.0f);

    for (int j = 0; j < 4; ++j) {
        const int dy = warp * 4 + j;
        const int r_lo = tabs.y.ofs[dy][0], r_hi = tabs.y.ofs[dy][3];     // clamped, non-decreasing
        if (lane == 0) {
            const unsigned bytes = (unsigned)(r_hi - r_lo + 1) * row_bytes;
            if (j > 0) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(stage_addr), "l"(base + (long long)r_lo * row_bytes), "r"(bytes), "r"(bar) : "memory");
        }
        // wait for the bytes of this output row (phase parity j & 1)
        {
            const unsigned parity = (unsigned)(j & 1);
            asm volatile(
                "{\n\t.reg .pred p;\n\t"
                "W_LOOP:\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                "@p bra W_DONE;\n\t"
                "bra W_LOOP;\n\t"
                "W_DONE:\n\t}\n" ::"r"(bar), "r"(parity) : "memory");
        }
        if (lane < 28) {
            const int dx = lane;
            int S[4][C];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const unsigned char* rp = stage + (size_t)(tabs.y.ofs[dy][k] - r_lo) * row_bytes;
                int acc[C] = {0, 0, 0};
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const unsigned char* p = rp + s_xofs[dx][t];
                    const int a = s_xcoef[dx][t];
#pragma unroll
                    for (int c = 0; c < C; ++c) acc[c] += (int)p[c] * a;
                }
#pragma unroll
                for (int c = 0; c < C; ++c) S[k][c] = acc[c];
            }
            const int b0 = tabs.y.coef[dy][0], b1 = tabs.y.coef[dy][1], b2 = tabs.y.coef[dy][2], b3 = tabs.y.coef[dy][3];
            const float f0 = __fmul_rn((float)b0, scale), f1 = __fmul_rn((float)b1, scale);
            const float f2 = __fmul_rn((float)b2, scale), f3 = __fmul_rn((float)b3, scale);
            int col[C];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const int e = dx * C + c;
                int v;
                if (e < NSIMD) {
                    float t = __fmul_rn(__int2float_rn(S[3][c]), f3);
                    t = __fadd_rn(__fmul_rn(__int2float_rn(S[2][c]), f2), t);
                    t = __fadd_rn(__fmul_rn(__int2float_rn(S[1][c]), f1), t);
                    t = __fadd_rn(__fmul_rn(__int2float_rn(S[0][c]), f0), t);
                    v = __float2int_rn(t);
                } else {
                    const int sum = S[0][c] * b0 + S[1][c] * b1 + S[2][c] * b2 + S[3][c] * b3;
                    v = (sum + (1 << 21)) >> 22;
                }
                col[c] = v < 0 ? 0 : (v > 255 ? 255 : v);
            }
            const int g = (3735 * col[0] + 19235 * col[1] + 9798 * col[2] + (1 << 14)) >> 15;
            const long long o = (long long)blockIdx.x * 784 + dy * 28 + dx;
            if (out_u8) out_u8[o] = (uint8_t)g;
            if (out_f32) out_f32[o] = __fmul_rn((float)g, inv255);
        }
        __syncwarp();     // every lane is done with the stage before lane 0 overwrites it
    }
}

}  // namespace

static int launch_train_chain(nc_ctx* ctx, const uint8_t* src, int n, int h, int w, int c,
                              int64_t img_stride, uint8_t* out_u8, float* out_f32, cudaStream_t st,
                              uint8_t* out_color = nullptr, int mode = 0) {
    TrainTabs tabs;
    const int identity = (h == 28 && w == 28) ? 1 : 0;
    if (mode == 1 && !identity) {
        // IPP arithmetic: float tables per shape, cached on the device like the Pillow tables (key: negative sizes)
        nc_dev_table& t = ctx->pil_tables[std::make_pair(-h, -w)];
        if (!t.dptr) {
            IppTabs ft;
            nc_ipp_cubic_axis(w, 28, &ft.x);
            nc_ipp_cubic_axis(h, 28, &ft.y);
            NC_CUDA(ctx, cudaMalloc(&t.dptr, sizeof(IppTabs)));
            NC_CUDA(ctx, cudaMemcpy(t.dptr, &ft, sizeof(IppTabs), cudaMemcpyHostToDevice));
            t.bytes = sizeof(IppTabs);
        }
        memset(&tabs, 0, sizeof(tabs));
        const IppTabs* dft = static_cast<const IppTabs*>(t.dptr);
        if (c == 3) train_chain_kernel<3, true><<<n, kThreads, 0, st>>>(src, img_stride, h, w, tabs, out_u8, out_f32, 0, out_color, dft);
        else train_chain_kernel<1, true><<<n, kThreads, 0, st>>>(src, img_stride, h, w, tabs, out_u8, out_f32, 0, out_color, dft);
        NC_LAUNCH_CHECK(ctx);
        return NC_OK;
    }
    if (!identity) {
        nc_cv_cubic_axis(w, 28, &tabs.x);
        nc_cv_cubic_axis(h, 28, &tabs.y);
    } else {
        memset(&tabs, 0, sizeof(tabs));
    }
    // row-streaming kernel: 3 channels, 16-byte aligned rows, source rows close enough together that
    // DRAM would fetch them whole anyway (tap spacing <= 192 B), and the 7 stages fit shared memory
    const long long row_bytes = (long long)w * c;
    if (c == 3 && !identity && !out_color && (row_bytes % 16) == 0 && (reinterpret_cast<uintptr_t>(src) % 16) == 0 && (img_stride % 16) == 0 &&
        h >= 4 && row_bytes * 4 * kRowWarps <= 200 * 1024 && row_bytes <= 28 * 192 && !getenv("NC_TRAIN_NO_ROWS")) {
        const unsigned stage_bytes = (unsigned)(row_bytes * 4);
        const size_t smem = (size_t)stage_bytes * kRowWarps;
        NC_CUDA(ctx, cudaFuncSetAttribute(train_chain_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        train_chain_rows_kernel<<<n, kRowThreads, smem, st>>>(src, img_stride, h, w, tabs, out_u8, out_f32, stage_bytes);
        NC_LAUNCH_CHECK(ctx);
        return NC_OK;
    }
    if (c == 3)
        train_chain_kernel<3, false><<<n, kThreads, 0, st>>>(src, img_stride, h, w, tabs, out_u8, out_f32, identity, out_color, nullptr);
    else
        train_chain_kernel<1, false><<<n, kThreads, 0, st>>>(src, img_stride, h, w, tabs, out_u8, out_f32, identity, out_color, nullptr);
    NC_LAUNCH_CHECK(ctx);
    return NC_OK;
}

static int check_train_args(nc_ctx* ctx, const void* src, int n, int h, int w, int c, int64_t img_stride,
                            int mode, const void* o1, const void* o2) {
    if (!ctx) return nc_set_error(nullptr, NC_EINVAL, "ctx is NULL");
    if (n < 0 || h <= 0 || w <= 0) return nc_set_error(ctx, NC_EINVAL, "bad shape n=%d h=%d w=%d", n, h, w);
    if (c != 1 && c != 3) return nc_set_error(ctx, NC_EINVAL, "channels must be 1 or 3, got %d", c);
    if (img_stride < (int64_t)h * w * c) return nc_set_error(ctx, NC_EINVAL, "img_stride %lld < h*w*c", (long long)img_stride);
    if (mode != 0 && mode != 1) return nc_set_error(ctx, NC_EINVAL, "resize mode %d (0 = OpenCV-native arithmetic, 1 = IPP arithmetic)", mode);
    if (n > 0 && !src) return nc_set_error(ctx, NC_EINVAL, "src is NULL");
    if (n > 0 && !o1 && !o2) return nc_set_error(ctx, NC_EINVAL, "both outputs are NULL");
    return NC_OK;
}

extern "C" int nc_pp_resize_gray_norm(nc_ctx* ctx, const uint8_t* src, int n, int h, int w, int c,
                                      int64_t img_stride, int mode, uint8_t* out_u8, float* out_f32,
                                      void* stream) {
    int rc = check_train_args(ctx, src, n, h, w, c, img_stride, mode, out_u8, out_f32);
    if (rc != NC_OK) return rc;
    if (n == 0) return NC_OK;
    NC_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_train_chain(ctx, src, n, h, w, c, img_stride, out_u8, out_f32, (cudaStream_t)stream, nullptr, mode);
}

extern "C" int nc_pp_resize_gray_norm_host(nc_ctx* ctx, const uint8_t* host_src, int n, int h, int w, int c,
                                           int64_t img_stride, int mode, uint8_t* out_u8, float* out_f32,
                                           int* staged, void* stream) {
    int rc = check_train_args(ctx, host_src, n, h, w, c, img_stride, mode, out_u8, out_f32);
    if (rc != NC_OK) return rc;
    if (staged) *staged = 0;
    if (n == 0) return NC_OK;
    NC_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    cudaPointerAttributes attr;
    cudaError_t e = cudaPointerGetAttributes(&attr, host_src);
    if (e != cudaSuccess) { cudaGetLastError(); attr.type = cudaMemoryTypeUnregistered; }
    if (attr.type == cudaMemoryTypeHost && attr.devicePointer != nullptr) {
        // pinned + mapped: gather the taps straight from host memory over PCIe (zero copy)
        return launch_train_chain(ctx, (const uint8_t*)attr.devicePointer, n, h, w, c, img_stride, out_u8, out_f32, st, nullptr, mode);
    }
    // pageable: stage the whole batch through the workspace
    // (the caller's buffer may end right behind the last picture: copy (n-1) strides + one picture, not n strides)
    const size_t bytes = (size_t)(n - 1) * (size_t)img_stride + (size_t)h * w * c;
    rc = nc_ws_reserve(ctx, (size_t)n * (size_t)img_stride);
    if (rc != NC_OK) return rc;
    NC_CUDA(ctx, cudaMemcpyAsync(ctx->ws, host_src, bytes, cudaMemcpyHostToDevice, st));
    if (staged) *staged = 1;
    return launch_train_chain(ctx, (const uint8_t*)ctx->ws, n, h, w, c, img_stride, out_u8, out_f32, st, nullptr, mode);
}

extern "C" int nc_pp_resize_cubic(nc_ctx* ctx, const uint8_t* src, int n, int h, int w, int c, int64_t img_stride,
                                  int mode, uint8_t* out, void* stream) {
    int rc = check_train_args(ctx, src, n, h, w, c, img_stride, mode, out, out);
    if (rc != NC_OK) return rc;
    if (n == 0) return NC_OK;
    NC_CUDA(ctx, cudaSetDevice(ctx->device));
    return launch_train_chain(ctx, src, n, h, w, c, img_stride, nullptr, nullptr, (cudaStream_t)stream, out, mode);
}

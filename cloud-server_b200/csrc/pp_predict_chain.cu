// Predict-side preprocess chain: 'L' conversion -> Pillow antialiased BICUBIC to 20x20 ->
// threshold onto a 28x28 canvas -> float32.
//
// Replaces apps/construction/util/construct_inference.py:312-330 for a batch of decoded pictures
// resident in HBM.  Unlike the train-side chain this one reads EVERY source byte (the antialias
// filter support is 2*scale: 205 x 154 taps for a 1024x768 upload), so it is the streaming,
// HBM-bound kernel of the preprocess half: 2 362 432 algorithmic bytes per 1024x768x3 picture.
//
// One CTA (320 threads) per picture, rows processed in blocks of `rb`:
//   phase 1  coalesced 32-bit loads -> rgb2l with dp4a on the packed byte stream (no byte
//            extraction) -> gray block in shared memory (row stride odd in words: conflict free)
//   phase 2  horizontal pass: each thread owns (row pair, output column); weights pre-split on the
//            host into three byte planes so that 4 taps cost 3 dp4a: sum p*k = sum p*k0 +
//            2^8 sum p*k1 + 2^16 sum p*k2 (exact in wrap-around int32, as Pillow's C int)
//   phase 3  vertical pass over the u8 intermediate [h][20] kept in shared memory
//   phase 4  threshold / canvas (compat: the reference's in-place 220 write + IndexError)
#include "nc_internal.h"
#include <string.h>
#include <stdlib.h>
#include <algorithm>

namespace {

constexpr int kThreads = 320;
constexpr int kOut = 20;

struct PredHdr {                 // per-shape table header (device + host copy)
    int a0[kOut];                // horizontal window start, aligned down to 4 pixels
    int ng[kOut];                // number of 4-tap groups
    int goff[kOut];              // offset of the first group
    int ymin[kOut], ycnt[kOut];  // vertical window
    int ksize_y, total_groups;
};

struct GrayCoef {                // rgb2l byte-plane masks for the 4 pixel phases of a 12-byte group
    unsigned lo[6], hi[6];       // order: p0(w0) p1(w0) p1(w1) p2(w1) p2(w2) p3(w2)
};

__device__ __forceinline__ unsigned dp4a_uu(unsigned a, unsigned b, unsigned c) {
    unsigned d;
    asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp4a_us(unsigned a, unsigned b_signed, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b_signed), "r"(c));
    return d;
}

__device__ __forceinline__ int clip8(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }

// phases 3 + 4, shared by both kernels: vertical pass over s_temp[h][20], then the threshold /
// canvas logic of construct_inference.py:317-330
__device__ __forceinline__ void finish_picture(const PredHdr& hdr, const unsigned char* s_temp, unsigned char* s_small,
                                               unsigned char* s_canvas, int* s_err, const int* __restrict__ g_kky, int compat,
                                               float* __restrict__ out_f32, uint8_t* __restrict__ out_small20,
                                               int* __restrict__ status, int nthreads) {
    const int tid = threadIdx.x;
    // ---------------- phase 3: vertical pass (ImagingResampleVertical_8bpc)
    for (int o = tid; o < kOut * kOut; o += nthreads) {
        const int yy = o / kOut, xx = o - yy * kOut;
        const int y0 = hdr.ymin[yy], cnt = hdr.ycnt[yy];
        const int* kk = g_kky + yy * hdr.ksize_y;
        unsigned acc = 1u << 21;
        for (int i = 0; i < cnt; ++i) acc += (unsigned)s_temp[(y0 + i) * kOut + xx] * (unsigned)__ldg(kk + i);
        s_small[o] = (unsigned char)clip8((int)acc >> 22);
    }
    for (int i = tid; i < 784; i += nthreads) s_canvas[i] = 0;
    if (tid == 0) *s_err = 0;
    __syncthreads();

    // out_small20 = the picture as im.resize((20,20)) returns it (:315), before the compat scan below patches 220s into it
    if (out_small20) {
        for (int o = tid; o < kOut * kOut; o += nthreads) out_small20[(long long)blockIdx.x * 400 + o] = s_small[o];
        if (compat) __syncthreads();      // (block-uniform) the scan of thread 0 writes s_small
    }

    // ---------------- phase 4: construct_inference.py:317-330
    if (!compat) {
        for (int o = tid; o < kOut * kOut; o += nthreads) {
            const int i = o / kOut, j = o - i * kOut;
            if (s_small[o] > 150) s_canvas[(i + 4) * 28 + (j + 4)] = 1;
        }
    } else if (tid == 0) {
        // sequential: the `elif arr[i][j] > 100: arr[bizar+i][bizar+j] = 220` statement feeds
        // later iterations of the same scan and raises IndexError past row/column 15
        int err = 0;
        for (int i = 0; i < kOut && !err; ++i)
            for (int j = 0; j < kOut; ++j) {
                const int v = s_small[i * kOut + j];
                if (v > 150) {
                    s_canvas[(i + 4) * 28 + (j + 4)] = 1;
                } else if (v > 100) {
                    if (i + 4 >= kOut || j + 4 >= kOut) { err = 1; break; }
                    s_small[(i + 4) * kOut + (j + 4)] = 220;
                }
            }
        *s_err = err;
    }
    __syncthreads();
    const int err = *s_err;
    if (tid == 0 && status) status[blockIdx.x] = err;
    const float v254 = (float)(254.0 / 255.0);
    for (int i = tid; i < 784; i += nthreads)
        out_f32[(long long)blockIdx.x * 784 + i] = (!err && s_canvas[i]) ? v254 : 0.0f;
}


// ---------------------------------------------------------------------------------------------
// TMA-staged variant (C == 3, rows 16-byte aligned): raw rows arrive in shared memory through
// cp.async.bulk (one bulk copy per 8-row stage, double buffered, mbarrier completion), so no warp
// ever waits on a global load; rgb2l reads the staged bytes with conflict-free LDS.128 and uses
// dp2a (16-bit coefficient pairs x pixel bytes): 8 dp2a + 3 prmt per 4 pixels.
// ---------------------------------------------------------------------------------------------
constexpr int kStageRows = 8;
constexpr int kBlockRows = 16;    // gray block consumed by one horizontal pass
constexpr int kSPB = kBlockRows / kStageRows;   // raw stages per gray block

__device__ __forceinline__ unsigned dp2a_lo(unsigned coef16x2, unsigned px, unsigned c) {
    unsigned d;
    asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(coef16x2), "r"(px), "r"(c));
    return d;
}
__device__ __forceinline__ unsigned dp2a_hi(unsigned coef16x2, unsigned px, unsigned c) {
    unsigned d;
    asm("dp2a.hi.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(coef16x2), "r"(px), "r"(c));
    return d;
}
// 12 stream bytes (4 pixels) -> 4 gray bytes packed in one word
__device__ __forceinline__ unsigned gray4(unsigned w0, unsigned w1, unsigned w2, unsigned A01, unsigned A2_, unsigned A_0,
                                          unsigned A12) {
    const unsigned s0 = dp2a_hi(A2_, w0, dp2a_lo(A01, w0, 0x8000u));     // bytes 0,1,2 of w0
    const unsigned s1 = dp2a_lo(A12, w1, dp2a_hi(A_0, w0, 0x8000u));     // byte 3 of w0, bytes 0,1 of w1
    const unsigned s2 = dp2a_lo(A2_, w2, dp2a_hi(A01, w1, 0x8000u));     // bytes 2,3 of w1, byte 0 of w2
    const unsigned s3 = dp2a_hi(A12, w2, dp2a_lo(A_0, w2, 0x8000u));     // bytes 1,2,3 of w2
    // L = sum >> 16 = byte 2 of each sum (sum < 2^24)
    const unsigned t01 = __byte_perm(s0, s1, 0x0062), t23 = __byte_perm(s2, s3, 0x0062);
    return __byte_perm(t01, t23, 0x5410);
}

__device__ __forceinline__ void mbar_wait(unsigned bar_addr, unsigned parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "DONE:\n\t}\n" ::"r"(bar_addr), "r"(parity) : "memory");
}

__global__ void __launch_bounds__(kThreads)
predict_chain_tma_kernel(const uint8_t* __restrict__ src, long long img_stride, int h, int w,
                         const PredHdr* __restrict__ g_hdr, const uint4* __restrict__ g_groups,
                         const int* __restrict__ g_kky, unsigned A01, unsigned A2_, unsigned A_0, unsigned A12,
                         int compat, float* __restrict__ out_f32, uint8_t* __restrict__ out_small20,
                         int* __restrict__ status, int wq, unsigned ipr, unsigned ipr_magic) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ PredHdr hdr;
    __shared__ unsigned char s_small[kOut * kOut];
    __shared__ unsigned char s_canvas[784];
    __shared__ int s_err;
    __shared__ __align__(8) unsigned long long s_bar[2];

    const int tid = threadIdx.x;
    for (int i = tid; i < (int)(sizeof(PredHdr) / 4); i += kThreads) ((int*)&hdr)[i] = ((const int*)g_hdr)[i];
    const unsigned row_bytes = (unsigned)w * 3u;
    const unsigned stage_bytes = kStageRows * row_bytes;
    unsigned char* raw = smem;                                            // 2 stages of raw rows
    __syncthreads();
    uint4* s_groups = reinterpret_cast<uint4*>(smem + 2 * stage_bytes);
    unsigned* s_gbuf = reinterpret_cast<unsigned*>(smem + 2 * stage_bytes + (size_t)hdr.total_groups * 16);
    unsigned char* s_temp = reinterpret_cast<unsigned char*>(s_gbuf + (size_t)kBlockRows * wq);

    const uint8_t* base = src + (long long)blockIdx.x * img_stride;
    const int nstages = (h + kStageRows - 1) / kStageRows;
    const unsigned bar0 = (unsigned)__cvta_generic_to_shared(&s_bar[0]);
    const unsigned raw0 = (unsigned)__cvta_generic_to_shared(raw);

    auto issue = [&](int s) {      // one elected thread: arm the barrier, start the bulk copy of stage s
        const int r0 = s * kStageRows;
        const unsigned bytes = (unsigned)min(kStageRows, h - r0) * row_bytes;
        const unsigned bar = bar0 + 8u * (s & 1);
        const unsigned dst = raw0 + (s & 1) * stage_bytes;
        const uint8_t* gsrc = base + (long long)r0 * row_bytes;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(gsrc), "r"(bytes), "r"(bar) : "memory");
    };

    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8u));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        issue(0);
        if (nstages > 1) issue(1);
    }
    for (int i = tid; i < hdr.total_groups; i += kThreads) s_groups[i] = g_groups[i];
    for (int i = tid; i < kBlockRows * wq; i += kThreads) s_gbuf[i] = 0u;
    __syncthreads();

    // horizontal-pass task of this thread: (row pair, output column, half of the taps)
    const int half = tid & 1, pr = (tid >> 1) & 7, xx = tid >> 4;
    const int ng = hdr.ng[xx], gh = (ng + 1) >> 1;
    const int g_begin = half ? gh : 0, g_end = half ? ng : gh;
    const unsigned* ra0 = s_gbuf + (size_t)(2 * pr) * wq + (hdr.a0[xx] >> 2);
    const uint4* wg = s_groups + hdr.goff[xx];

    for (int s = 0; s < nstages; ++s) {
        const int r0 = s * kStageRows;
        const int rows = min(kStageRows, h - r0);
        mbar_wait(bar0 + 8u * (s & 1), (unsigned)((s >> 1) & 1));
        // ---------------- phase 1: staged bytes -> rgb2l -> gray block rows (s&1)*8 ..
        const unsigned char* rbuf = raw + (size_t)(s & 1) * stage_bytes;
        unsigned* gdst = s_gbuf + (size_t)((s % kSPB) * kStageRows) * wq;
        const unsigned items = (unsigned)rows * ipr;                      // ipr = 16-pixel items per row
        for (unsigned it = tid; it < items; it += kThreads) {
            const unsigned rl = ipr_magic ? __umulhi(it, ipr_magic) : it;
            const unsigned g16 = it - rl * ipr;
            const uint4* p = reinterpret_cast<const uint4*>(rbuf + (size_t)rl * row_bytes + (size_t)g16 * 48);
            const uint4 q0 = p[0], q1 = p[1], q2 = p[2];
            uint2 o0, o1;
            o0.x = gray4(q0.x, q0.y, q0.z, A01, A2_, A_0, A12);
            o0.y = gray4(q0.w, q1.x, q1.y, A01, A2_, A_0, A12);
            o1.x = gray4(q1.z, q1.w, q2.x, A01, A2_, A_0, A12);
            o1.y = gray4(q2.y, q2.z, q2.w, A01, A2_, A_0, A12);
            uint2* d = reinterpret_cast<uint2*>(gdst + (size_t)rl * wq + g16 * 4);   // wq even: 8-byte aligned
            d[0] = o0;
            d[1] = o1;
        }
        __syncthreads();          // raw stage consumed, gray rows visible
        if (tid == 0 && s + 2 < nstages) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic reads before async writes
            issue(s + 2);
        }
        if ((s % kSPB) == kSPB - 1 || s == nstages - 1) {
            // ---------------- phase 2: horizontal pass over the 16-row gray block
            const int b0 = (s / kSPB) * kBlockRows;               // first picture row of this block
            const int brow = min(kBlockRows, h - b0);
            const unsigned* ra = ra0;
            const unsigned* rb = ra0 + wq;
            unsigned s0a = 0, s1a = 0, s0b = 0, s1b = 0;
            int s2a = 0, s2b = 0;
#pragma unroll 4
            for (int i = g_begin; i < g_end; ++i) {
                const uint4 k = wg[i];
                const unsigned pa = ra[i], pb = rb[i];
                s0a = dp4a_uu(pa, k.x, s0a); s1a = dp4a_uu(pa, k.y, s1a); s2a = dp4a_us(pa, k.z, s2a);
                s0b = dp4a_uu(pb, k.x, s0b); s1b = dp4a_uu(pb, k.y, s1b); s2b = dp4a_us(pb, k.z, s2b);
            }
            unsigned va = s0a + (s1a << 8) + ((unsigned)s2a << 16);
            unsigned vb = s0b + (s1b << 8) + ((unsigned)s2b << 16);
            va += __shfl_xor_sync(0xffffffffu, va, 1);
            vb += __shfl_xor_sync(0xffffffffu, vb, 1);
            const int row = 2 * pr + half;
            if (row < brow) {
                const unsigned v = half ? vb : va;
                s_temp[(b0 + row) * kOut + xx] = (unsigned char)clip8((int)(v + (1u << 21)) >> 22);
            }
            __syncthreads();      // gray block free for the next two stages
        }
    }
    finish_picture(hdr, s_temp, s_small, s_canvas, &s_err, g_kky, compat, out_f32, out_small20, status, kThreads);
}

template <int C>
__global__ void __launch_bounds__(kThreads)
predict_chain_kernel(const uint8_t* __restrict__ src, long long img_stride, int h, int w,
                     const PredHdr* __restrict__ g_hdr, const uint4* __restrict__ g_groups,
                     const int* __restrict__ g_kky, const __grid_constant__ GrayCoef gc,
                     int r_coef, int g_coef, int b_coef,   // applied to stream bytes 0,1,2
                     int compat, float* __restrict__ out_f32, uint8_t* __restrict__ out_small20,
                     int* __restrict__ status, int rb, int wq, unsigned wq_src, unsigned wq_magic) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ PredHdr hdr;
    __shared__ unsigned char s_small[kOut * kOut];
    __shared__ unsigned char s_canvas[784];
    __shared__ int s_err;

    const int tid = threadIdx.x;
    for (int i = tid; i < (int)(sizeof(PredHdr) / 4); i += kThreads) ((int*)&hdr)[i] = ((const int*)g_hdr)[i];
    __syncthreads();

    uint4* s_groups = reinterpret_cast<uint4*>(smem);
    unsigned* s_gbuf = reinterpret_cast<unsigned*>(smem + (size_t)hdr.total_groups * 16);
    unsigned char* s_temp = reinterpret_cast<unsigned char*>(s_gbuf + (size_t)rb * wq);

    for (int i = tid; i < hdr.total_groups; i += kThreads) s_groups[i] = g_groups[i];
    // zero the gray block once: pad words beyond the row end are multiplied by zero weights but
    // must hold defined bytes
    for (int i = tid; i < rb * wq; i += kThreads) s_gbuf[i] = 0u;
    __syncthreads();

    const uint8_t* base = src + (long long)blockIdx.x * img_stride;
    const long long row_bytes = (long long)w * C;
    const bool aligned = ((reinterpret_cast<unsigned long long>(base) & 3ull) == 0) && ((w & 3) == 0);

    for (int r0 = 0; r0 < h; r0 += rb) {
        const int rows = min(rb, h - r0);
        // ---------------- phase 1: load + rgb2l -> s_gbuf
        const unsigned items = (unsigned)rows * wq_src;   // wq_src = ceil(w/4) groups per row
        for (unsigned it = tid; it < items; it += kThreads) {
            const unsigned rl = wq_magic ? __umulhi(it, wq_magic) : it;   // it / wq_src (exact for it < 2^21)
            const unsigned g = it - rl * wq_src;
            const uint8_t* rp = base + (long long)(r0 + rl) * row_bytes;
            unsigned packed;
            if (C == 1) {
                if (aligned) {
                    packed = __ldg(reinterpret_cast<const unsigned*>(rp) + g);
                } else {
                    packed = 0;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int x = 4 * g + k;
                        if (x < w) packed |= (unsigned)__ldg(rp + x) << (8 * k);
                    }
                }
            } else if (aligned) {
                const unsigned* wp3 = reinterpret_cast<const unsigned*>(rp) + 3 * g;
                const unsigned w0 = __ldg(wp3), w1 = __ldg(wp3 + 1), w2 = __ldg(wp3 + 2);
                // L = (19595 R + 38470 G + 7471 B + 0x8000) >> 16 with the 16-bit coefficients split
                // into byte planes: lo-plane sum + 256 * hi-plane sum
                unsigned l0 = dp4a_uu(w0, gc.lo[0], 0x8000u), h0 = dp4a_uu(w0, gc.hi[0], 0u);
                unsigned l1 = dp4a_uu(w1, gc.lo[2], dp4a_uu(w0, gc.lo[1], 0x8000u));
                unsigned h1 = dp4a_uu(w1, gc.hi[2], dp4a_uu(w0, gc.hi[1], 0u));
                unsigned l2 = dp4a_uu(w2, gc.lo[4], dp4a_uu(w1, gc.lo[3], 0x8000u));
                unsigned h2 = dp4a_uu(w2, gc.hi[4], dp4a_uu(w1, gc.hi[3], 0u));
                unsigned l3 = dp4a_uu(w2, gc.lo[5], 0x8000u), h3 = dp4a_uu(w2, gc.hi[5], 0u);
                const unsigned p0 = (l0 + (h0 << 8)) >> 16, p1 = (l1 + (h1 << 8)) >> 16;
                const unsigned p2 = (l2 + (h2 << 8)) >> 16, p3 = (l3 + (h3 << 8)) >> 16;
                packed = p0 | (p1 << 8) | (p2 << 16) | (p3 << 24);
            } else {
                packed = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int x = 4 * g + k;
                    if (x < w) {
                        const uint8_t* p = rp + 3 * x;
                        const unsigned v = (unsigned)(r_coef * (int)__ldg(p) + g_coef * (int)__ldg(p + 1) +
                                                      b_coef * (int)__ldg(p + 2) + 0x8000) >> 16;
                        packed |= v << (8 * k);
                    }
                }
            }
            s_gbuf[rl * wq + g] = packed;
        }
        __syncthreads();

        // ---------------- phase 2: horizontal pass (ImagingResampleHorizontal_8bpc)
        const int npairs = (rows + 1) >> 1;
        const int half = rb >> 1;                       // row pairs per block (rb is even)
        for (int t = tid; t < half * kOut; t += kThreads) {
            const int rp_ = t % half, xx = t / half;
            if (rp_ < npairs) {
                const unsigned* ra = s_gbuf + (size_t)(2 * rp_) * wq + (hdr.a0[xx] >> 2);
                const unsigned* rbp = ra + wq;
                const uint4* wg = s_groups + hdr.goff[xx];
                const int ng = hdr.ng[xx];
                unsigned s0a = 0, s1a = 0, s0b = 0, s1b = 0;
                int s2a = 0, s2b = 0;
#pragma unroll 4
                for (int i = 0; i < ng; ++i) {
                    const uint4 k = wg[i];
                    const unsigned pa = ra[i], pb = rbp[i];
                    s0a = dp4a_uu(pa, k.x, s0a); s1a = dp4a_uu(pa, k.y, s1a); s2a = dp4a_us(pa, k.z, s2a);
                    s0b = dp4a_uu(pb, k.x, s0b); s1b = dp4a_uu(pb, k.y, s1b); s2b = dp4a_us(pb, k.z, s2b);
                }
                const int ya = r0 + 2 * rp_;
                const int ta = (int)(s0a + (s1a << 8) + ((unsigned)s2a << 16) + (1u << 21));
                s_temp[ya * kOut + xx] = (unsigned char)clip8(ta >> 22);
                if (2 * rp_ + 1 < rows) {
                    const int tb = (int)(s0b + (s1b << 8) + ((unsigned)s2b << 16) + (1u << 21));
                    s_temp[(ya + 1) * kOut + xx] = (unsigned char)clip8(tb >> 22);
                }
            }
        }
        __syncthreads();
    }

    finish_picture(hdr, s_temp, s_small, s_canvas, &s_err, g_kky, compat, out_f32, out_small20, status, kThreads);
}

// ---------------------------------------------------------------------------------------------
// Tensor-core variant of the horizontal pass.  The pass is a banded integer matrix product
// (gray rows [16 x w] u8) x (weights [w x 20] s32); with the weights split into three signed
// byte digits (k = d0 + 256 d1 + 65536 d2, each in [-128,127]) it is exact in int32 on the
// u8 x s8 -> s32 warp MMA (m16n8k32).  Warp v owns output columns 2v and 2v+1: the 8 MMA
// columns are {o0d0, o0d1, o1d0, o1d1, o0d2, 0, o1d2, 0}, its K range is the union of the two tap
// windows (nch 32-pixel chunks from kw[v]), and its B fragments are constant for the whole
// picture, so they live in registers.  A fragments come from the gray block with one ldmatrix.x4
// per chunk (the 16 x 32-byte u8 tile is four 8 x 16-byte matrices; K ranges start on 16-pixel
// boundaries).  One raw stage = one gray block of `sr` <= 16 rows, gray
// blocks double buffered: a single __syncthreads per stage.
// Costs 9 MMA + 9 LDSM per warp and block where the dp4a pass needs ~270 instructions.
// ---------------------------------------------------------------------------------------------
struct MmaPlan {
    int kw[kOut / 2];        // first gray word (4 pixels) of the warp's K range, a multiple of 4 (16-byte aligned rows for ldmatrix)
    int nch;                 // 32-pixel chunks in every warp's range
};

__device__ __forceinline__ void mma_u8s8(int& c0, int& c1, int& c2, int& c3, unsigned a0, unsigned a1, unsigned a2, unsigned a3,
                                         unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c0), "+r"(c1), "+r"(c2), "+r"(c3)
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// FUSE: the train-side chain (pp_train_chain.cu: OpenCV cubic 28x28 -> BGR2GRAY -> *1/255) computed from the same
// staged rows, so the upload is read from HBM once for both chains.  The 112 (output row, tap) source rows are met in
// increasing order while the picture streams through; each one's horizontal pass (84 int32) goes to a ring of 8 row
// slots (two output rows), and an output row is finished -- vertical pass, gray, store -- by warp 0 in the stage after
// its fourth tap row.  The host only selects FUSE when the tap rows are sorted and two output rows apart never share
// two consecutive stages, which makes the ring race free with the one __syncthreads per stage.
struct FuseArgs {
    CvCubicAxis x, y;
    uint8_t* out_u8;          // [n][784] gray, nullable
    float* out_f32;           // [n][784] gray / 255, nullable
};

// VResizeCubic (float32 body for the first 80 interleaved elements, integer tail) + BGR2GRAY + *1/255 of output row dy,
// pixel dx, from the four tap rows in the ring; same arithmetic as train_chain_rows_kernel (pp_train_chain.cu)
__device__ __forceinline__ void train_row_finish(const FuseArgs& fa, const int (*hbuf)[84], int dy, int dx) {
    constexpr int NSIMD = (84 / 8) * 8;
    const float scale = 1.0f / (2048.0f * 2048.0f);
    const int b0 = fa.y.coef[dy][0], b1 = fa.y.coef[dy][1], b2 = fa.y.coef[dy][2], b3 = fa.y.coef[dy][3];
    const float f0 = __fmul_rn((float)b0, scale), f1 = __fmul_rn((float)b1, scale);
    const float f2 = __fmul_rn((float)b2, scale), f3 = __fmul_rn((float)b3, scale);
    const int q = (4 * dy) & 7;
    int col[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const int e = dx * 3 + c;
        const int S0 = hbuf[q][e], S1 = hbuf[q + 1][e], S2 = hbuf[q + 2][e], S3 = hbuf[q + 3][e];
        int v;
        if (e < NSIMD) {
            float t = __fmul_rn(__int2float_rn(S3), f3);
            t = __fadd_rn(__fmul_rn(__int2float_rn(S2), f2), t);
            t = __fadd_rn(__fmul_rn(__int2float_rn(S1), f1), t);
            t = __fadd_rn(__fmul_rn(__int2float_rn(S0), f0), t);
            v = __float2int_rn(t);
        } else {
            const int sum = S0 * b0 + S1 * b1 + S2 * b2 + S3 * b3;
            v = (sum + (1 << 21)) >> 22;
        }
        col[c] = v < 0 ? 0 : (v > 255 ? 255 : v);
    }
    const int g = (3735 * col[0] + 19235 * col[1] + 9798 * col[2] + (1 << 14)) >> 15;
    const long long o = (long long)blockIdx.x * 784 + dy * 28 + dx;
    if (fa.out_u8) fa.out_u8[o] = (uint8_t)g;
    if (fa.out_f32) fa.out_f32[o] = __fmul_rn((float)g, (float)(1.0 / 255.0));
}

template <int NCH, bool FUSE>
__global__ void __launch_bounds__(kThreads + 32, 2)
predict_chain_mma_kernel(const uint8_t* __restrict__ src, long long img_stride, int h, int w,
                         const PredHdr* __restrict__ g_hdr, const MmaPlan* __restrict__ g_plan,
                         const uint2* __restrict__ g_bfrag,
                         const int* __restrict__ g_kky, unsigned A01, unsigned A2_, unsigned A_0, unsigned A12,
                         int compat, float* __restrict__ out_f32, uint8_t* __restrict__ out_small20,
                         int* __restrict__ status, int pitch, int sr, unsigned ipr, const __grid_constant__ FuseArgs fa) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ int s_hbuf[FUSE ? 8 : 1][84];
    __shared__ int s_xofs3[FUSE ? 28 : 1][4], s_xcoef[FUSE ? 28 : 1][4];
    __shared__ PredHdr hdr;
    __shared__ unsigned char s_small[kOut * kOut];
    __shared__ unsigned char s_canvas[784];
    __shared__ int s_err;
    __shared__ __align__(8) unsigned long long s_bar[2];

    // FUSE adds an 11th warp that finishes the train-side output rows (vertical pass, gray, stores), so that serial
    // piece never sits between a worker warp and the stage barrier; the tap sums themselves are spread over the workers
    constexpr int kAll = FUSE ? kThreads + 32 : kThreads;
    const int tid = threadIdx.x;
    const bool worker = !FUSE || tid < kThreads;
    for (int i = tid; i < (int)(sizeof(PredHdr) / 4); i += kAll) ((int*)&hdr)[i] = ((const int*)g_hdr)[i];
    if (FUSE)
        for (int i = tid; i < 112; i += kAll) {
            s_xofs3[i >> 2][i & 3] = fa.x.ofs[i >> 2][i & 3] * 3;
            s_xcoef[i >> 2][i & 3] = fa.x.coef[i >> 2][i & 3];
        }
    const unsigned row_bytes = (unsigned)w * 3u;
    const unsigned stage_bytes = (unsigned)sr * row_bytes;
    unsigned char* raw = smem;                                            // 2 stages of raw rows
    // 2 gray blocks of sr rows x pitch words + (16 - sr) rows of slack: the MMA always reads 16 rows, the rows past sr
    // belong to the other block / the slack and only feed accumulator rows that are dropped
    unsigned* s_gbuf = reinterpret_cast<unsigned*>(smem + 2 * stage_bytes);
    const unsigned gblock = (unsigned)sr * (unsigned)pitch;
    const unsigned gwords = (unsigned)(sr + 16) * (unsigned)pitch;
    unsigned char* s_temp = reinterpret_cast<unsigned char*>(s_gbuf + gwords);

    const uint8_t* base = src + (long long)blockIdx.x * img_stride;
    const int nstages = (h + sr - 1) / sr;
    const unsigned bar0 = (unsigned)__cvta_generic_to_shared(&s_bar[0]);
    const unsigned raw0 = (unsigned)__cvta_generic_to_shared(raw);

    auto issue = [&](int s) {
        const int r0 = s * sr;
        const unsigned bytes = (unsigned)min(sr, h - r0) * row_bytes;
        const unsigned bar = bar0 + 8u * (s & 1);
        const unsigned dst = raw0 + (s & 1) * stage_bytes;
        const uint8_t* gsrc = base + (long long)r0 * row_bytes;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(gsrc), "r"(bytes), "r"(bar) : "memory");
    };
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8u));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        issue(0);
        if (nstages > 1) issue(1);
    }
    // rows sr..15 of a gray block are never written by phase 1 but are MMA operands (their results are dropped)
    for (unsigned i = tid; i < gwords; i += kAll) s_gbuf[i] = 0u;

    // this warp's B fragments and K range
    const int wv = worker ? tid >> 5 : 0, lane = tid & 31, gid = lane >> 2, tig = lane & 3;
    unsigned bx[NCH], by[NCH];
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
        const uint2 b = __ldg(g_bfrag + ((size_t)wv * NCH + c) * 32 + lane);
        bx[c] = b.x; by[c] = b.y;
    }
    // ldmatrix address of this lane: the A fragment of one chunk is four 8-row x 16-byte matrices (rows 0-7 / 8-15, pixel
    // bytes 0-15 / 16-31 of the chunk); lanes 8m .. 8m+7 supply the row addresses of matrix m.  Byte offset in a gray block.
    const unsigned a_off = ((unsigned)((lane & 7) + ((lane >> 3) & 1) * 8) * (unsigned)pitch + (unsigned)g_plan->kw[wv]) * 4u +
                           (unsigned)(lane >> 4) * 16u;

    // phase-1 item walk: item = 16 pixels = 48 contiguous staged bytes -> 4 gray words; items of a stage are
    // contiguous in the raw buffer (row_bytes = 48 ipr), the gray rows have `pitch - 4 ipr` words of padding
    const unsigned it_g0 = (unsigned)tid % ipr, it_r0 = (unsigned)tid / ipr;
    const unsigned dg = (unsigned)kThreads % ipr, dr = (unsigned)kThreads / ipr;
    const unsigned padw = (unsigned)pitch - 4u * ipr;
    const unsigned d_dst = 4u * (unsigned)kThreads + dr * padw;          // words per step without wrap
    const unsigned dst0 = it_r0 * (unsigned)pitch + 4u * it_g0;
    const unsigned full_items = (unsigned)sr * ipr, last_items = (unsigned)(h - (nstages - 1) * sr) * ipr;
    const bool two_items = full_items == 2u * (unsigned)kThreads;
    const unsigned dst1 = dst0 + d_dst + ((it_g0 + dg >= ipr) ? padw : 0u);       // gray offset of this thread's second item
    // per-buffer bases (buffer = s & 1), selected per stage instead of recomputed
    const uint4* p_b0 = reinterpret_cast<const uint4*>(raw) + 3u * (unsigned)tid;
    const uint4* p_b1 = reinterpret_cast<const uint4*>(raw + stage_bytes) + 3u * (unsigned)tid;
    unsigned* g_b0 = s_gbuf;
    unsigned* g_b1 = s_gbuf + gblock;
    const int out_col = 2 * wv + (tig & 1);
    const int out_row = gid + ((tig & 2) ? 8 : 0);
    unsigned char* t_out = s_temp + out_row * kOut + out_col;
    int ns = 0, nv = 0;       // FUSE: next tap-row slot to fill (0..112), next output row to finish (0..28); uniform
    int next_row = FUSE ? fa.y.ofs[0][0] : 0x7fffffff;        // source row of slot ns
    __syncthreads();

    for (int s = 0; s < nstages; ++s) {
        const int odd = s & 1;
        const bool last = s == nstages - 1;
        if (worker) mbar_wait(bar0 + 8u * odd, (unsigned)((s >> 1) & 1));
        // ---------------- phase 1: staged bytes -> rgb2l -> gray block (s & 1)
        unsigned* gblk = odd ? g_b1 : g_b0;
        if (worker && !last && two_items) {
            // the common shape (1024 wide, 10-row stages): exactly two items per thread, straight-line, all six loads first
            const uint4* p = odd ? p_b1 : p_b0;
            const uint4 q0 = p[0], q1 = p[1], q2 = p[2];
            const uint4 r0 = p[3 * kThreads], r1 = p[3 * kThreads + 1], r2 = p[3 * kThreads + 2];
            uint4 o;
            o.x = gray4(q0.x, q0.y, q0.z, A01, A2_, A_0, A12);
            o.y = gray4(q0.w, q1.x, q1.y, A01, A2_, A_0, A12);
            o.z = gray4(q1.z, q1.w, q2.x, A01, A2_, A_0, A12);
            o.w = gray4(q2.y, q2.z, q2.w, A01, A2_, A_0, A12);
            *reinterpret_cast<uint4*>(gblk + dst0) = o;
            o.x = gray4(r0.x, r0.y, r0.z, A01, A2_, A_0, A12);
            o.y = gray4(r0.w, r1.x, r1.y, A01, A2_, A_0, A12);
            o.z = gray4(r1.z, r1.w, r2.x, A01, A2_, A_0, A12);
            o.w = gray4(r2.y, r2.z, r2.w, A01, A2_, A_0, A12);
            *reinterpret_cast<uint4*>(gblk + dst1) = o;
        } else if (worker) {
            const unsigned items = last ? last_items : full_items;
            unsigned g = it_g0;
            unsigned dst = dst0;
            const uint4* p = odd ? p_b1 : p_b0;
            for (unsigned it = tid; it < items; it += kThreads) {
                const uint4 q0 = p[0], q1 = p[1], q2 = p[2];
                uint2 o0, o1;
                o0.x = gray4(q0.x, q0.y, q0.z, A01, A2_, A_0, A12);
                o0.y = gray4(q0.w, q1.x, q1.y, A01, A2_, A_0, A12);
                o1.x = gray4(q1.z, q1.w, q2.x, A01, A2_, A_0, A12);
                o1.y = gray4(q2.y, q2.z, q2.w, A01, A2_, A_0, A12);
                *reinterpret_cast<uint4*>(gblk + dst) = make_uint4(o0.x, o0.y, o1.x, o1.y);   // pitch % 4 == 0: 16-byte aligned
                p += 3 * kThreads;
                g += dg;
                dst += d_dst;
                if (g >= ipr) { g -= ipr; dst += padw; }
            }
        }
        if (FUSE) {
            // train-side horizontal pass (HResizeCubic<uchar,int,short>) of the tap rows that are in this stage:
            // (slot, element) pairs spread over the worker threads, 4 taps each.  About every second stage has none.
            const int r0 = s * sr;
            const int rend = last ? h : r0 + sr;
            if (next_row < rend) {
                const int ns0 = ns;
                do {
                    ++ns;
                    next_row = ns < 112 ? fa.y.ofs[ns >> 2][ns & 3] : 0x7fffffff;
                } while (next_row < rend);
                if (worker) {
                    const unsigned char* rb = raw + (size_t)odd * stage_bytes;
                    const int pairs = (ns - ns0) * 84;
                    for (int pi = tid; pi < pairs; pi += kThreads) {
                        const int sl = pi / 84, e = pi - sl * 84;
                        const int slot = ns0 + sl;
                        const int dx = e / 3, ch = e - dx * 3;
                        const unsigned char* rp = rb + (size_t)(fa.y.ofs[slot >> 2][slot & 3] - r0) * row_bytes + ch;
                        int acc = 0;
#pragma unroll
                        for (int t = 0; t < 4; ++t) acc += (int)rp[s_xofs3[dx][t]] * s_xcoef[dx][t];
                        s_hbuf[slot & 7][e] = acc;
                    }
                }
            }
        }
        __syncthreads();          // raw stage consumed; gray block (s&1) complete; everyone is past phase 2 of stage s-1
        if (FUSE && !worker) {
            while (4 * nv + 3 < ns) {                     // ns <= 112, so nv stops at 28
                if (lane < 28) train_row_finish(fa, s_hbuf, nv, lane);
                ++nv;
            }
        }
        if (tid == 0 && s + 2 < nstages) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic reads before async writes
            issue(s + 2);
        }
        // ---------------- phase 2: horizontal pass, one m16n8k32 MMA per 32-pixel chunk
        if (worker) {
            const unsigned a_addr = (unsigned)__cvta_generic_to_shared(gblk) + a_off;
            int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                unsigned a0, a1, a2, a3;                                   // one ldmatrix = the whole 16 x 32-byte A fragment
                asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                             : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3) : "r"(a_addr + 32u * (unsigned)c));
                mma_u8s8(c0, c1, c2, c3, a0, a1, a2, a3, bx[c], by[c]);
            }
            // columns held: (2 tig, 2 tig + 1) = tig 0: o0 d0,d1 | tig 1: o1 d0,d1 | tig 2: o0 d2,0 | tig 3: o1 d2,0
            int p_lo = tig < 2 ? c0 + c1 * 256 : c0 * 65536;
            int p_hi = tig < 2 ? c2 + c3 * 256 : c2 * 65536;
            p_lo += __shfl_xor_sync(0xffffffffu, p_lo, 2);
            p_hi += __shfl_xor_sync(0xffffffffu, p_hi, 2);
            const int v = (tig & 2) ? p_hi : p_lo;
            if (out_row < (last ? h - s * sr : sr)) *t_out = (unsigned char)clip8((v + (1 << 21)) >> 22);
            t_out += sr * kOut;
        }
    }
    __syncthreads();
    if (FUSE && !worker) {
        while (4 * nv + 3 < ns) {
            if (lane < 28) train_row_finish(fa, s_hbuf, nv, lane);
            ++nv;
        }
    }
    finish_picture(hdr, s_temp, s_small, s_canvas, &s_err, g_kky, compat, out_f32, out_small20, status, kAll);
}

struct PredTabDev {
    PredHdr hdr;
    PredHdr* d_hdr;
    uint4* d_groups;
    int* d_kky;
    MmaPlan* d_plan;
    uint2* d_bfrag;
    int bstride;             // chunks per warp in d_bfrag (= max nch)
};

// B fragments of the tensor-core horizontal pass: [warp v][chunk c][lane] -> (b0, b1), see predict_chain_mma_kernel
void build_mma_plan(const PilAxisHost& ax, int w, MmaPlan* plan, std::vector<uint2>* frag) {
    int maxch = 1;
    int k0s[kOut / 2];
    for (int v = 0; v < kOut / 2; ++v) {
        const int lo = std::min(ax.xmin[2 * v], ax.xmin[2 * v + 1]);
        const int hi = std::max(ax.xmin[2 * v] + ax.cnt[2 * v], ax.xmin[2 * v + 1] + ax.cnt[2 * v + 1]);
        k0s[v] = lo & ~15;
        maxch = std::max(maxch, (hi - k0s[v] + 31) / 32);
    }
    plan->nch = maxch;
    // every warp runs maxch chunks; slide a range down where it would pass the 32 pixels of slack behind the row
    for (int v = 0; v < kOut / 2; ++v) {
        int k0 = k0s[v];
        const int limit = (w + 32 - 32 * maxch) & ~15;
        if (k0 > limit) k0 = limit;
        if (k0 < 0) k0 = 0;
        plan->kw[v] = k0 / 4;
    }
    frag->assign((size_t)(kOut / 2) * maxch * 32, make_uint2(0u, 0u));
    auto digit = [&](int o, int x, int d) -> int {      // signed byte digit d of output o's weight at pixel x
        const int xmin = ax.xmin[o], cnt = ax.cnt[o];
        if (x < xmin || x >= xmin + cnt) return 0;
        int k = ax.kk[(size_t)o * ax.ksize + (x - xmin)];
        for (int i = 0;; ++i) {
            const int dg = ((k + 128) & 255) - 128;
            if (i == d) return dg;
            k = (k - dg) / 256;                          // exact: k - dg is a multiple of 256
        }
    };
    for (int v = 0; v < kOut / 2; ++v)
        for (int c = 0; c < maxch; ++c)
            for (int lane = 0; lane < 32; ++lane) {
                const int n = lane >> 2, tig = lane & 3;
                // column n: {o0 d0, o0 d1, o1 d0, o1 d1, o0 d2, 0, o1 d2, 0}
                static const int col_o[8] = {0, 0, 1, 1, 0, -1, 1, -1}, col_d[8] = {0, 1, 0, 1, 2, 0, 2, 0};
                uint2 q = make_uint2(0u, 0u);
                if (col_o[n] >= 0) {
                    const int o = 2 * v + col_o[n];
                    for (int i = 0; i < 4; ++i) {
                        // K slots 4 tig + i (b0) and 16 + 4 tig + i (b1) are those pixels of the chunk
                        const int x0 = plan->kw[v] * 4 + 32 * c + 4 * tig + i;
                        q.x |= ((unsigned)digit(o, x0, col_d[n]) & 255u) << (8 * i);
                        q.y |= ((unsigned)digit(o, x0 + 16, col_d[n]) & 255u) << (8 * i);
                    }
                }
                (*frag)[((size_t)v * maxch + c) * 32 + lane] = q;
            }
}
}  // namespace

// Build (or fetch from the ctx cache) the per-shape tables.
static int get_pred_tables(nc_ctx* ctx, int h, int w, PredTabDev* out) {
    auto key = std::make_pair(h, w);
    auto it = ctx->pil_tables.find(key);
    if (it == ctx->pil_tables.end()) {
        PilAxisHost ax, ay;
        nc_pil_bicubic_axis(w, kOut, &ax);
        nc_pil_bicubic_axis(h, kOut, &ay);
        PredHdr hdr;
        memset(&hdr, 0, sizeof(hdr));
        std::vector<uint4> groups;
        for (int xx = 0; xx < kOut; ++xx) {
            const int xmin = ax.xmin[xx], cnt = ax.cnt[xx];
            const int a0 = xmin & ~3;
            const int ng = (xmin + cnt - a0 + 3) / 4;
            hdr.a0[xx] = a0;
            hdr.ng[xx] = ng;
            hdr.goff[xx] = (int)groups.size();
            for (int g = 0; g < ng; ++g) {
                uint4 q = make_uint4(0, 0, 0, 0);
                for (int k = 0; k < 4; ++k) {
                    const int x = a0 + 4 * g + k;
                    int kv = 0;
                    if (x >= xmin && x < xmin + cnt) kv = ax.kk[(size_t)xx * ax.ksize + (x - xmin)];
                    const unsigned u = (unsigned)kv;
                    q.x |= (u & 255u) << (8 * k);
                    q.y |= ((u >> 8) & 255u) << (8 * k);
                    q.z |= ((u >> 16) & 255u) << (8 * k);   // signed byte: bits 16..23 of the 2's complement
                    // weights are < 2^23 in magnitude (they are normalised to sum 2^22)
                }
                groups.push_back(q);
            }
            hdr.ymin[xx] = ay.xmin[xx];
            hdr.ycnt[xx] = ay.cnt[xx];
        }
        hdr.ksize_y = ay.ksize;
        hdr.total_groups = (int)groups.size();
        MmaPlan plan;
        std::vector<uint2> frag;
        build_mma_plan(ax, w, &plan, &frag);
        const int bstride = plan.nch;
        const size_t b_hdr = (sizeof(PredHdr) + 15) / 16 * 16;
        const size_t b_groups = groups.size() * sizeof(uint4);
        const size_t b_kky = (ay.kk.size() * sizeof(int) + 15) / 16 * 16;
        const size_t b_plan = (sizeof(MmaPlan) + 15) / 16 * 16;
        const size_t b_frag = frag.size() * sizeof(uint2);
        nc_dev_table tab;
        tab.bytes = b_hdr + b_groups + b_kky + b_plan + b_frag;
        NC_CUDA(ctx, cudaMalloc(&tab.dptr, tab.bytes));
        std::vector<unsigned char> blob(tab.bytes, 0);
        memcpy(blob.data(), &hdr, sizeof(hdr));
        memcpy(blob.data() + b_hdr, groups.data(), b_groups);
        memcpy(blob.data() + b_hdr + b_groups, ay.kk.data(), ay.kk.size() * sizeof(int));
        memcpy(blob.data() + b_hdr + b_groups + b_kky, &plan, sizeof(plan));
        memcpy(blob.data() + b_hdr + b_groups + b_kky + b_plan, frag.data(), b_frag);
        NC_CUDA(ctx, cudaMemcpy(tab.dptr, blob.data(), tab.bytes, cudaMemcpyHostToDevice));
        // host copy: header, then the sizes needed to find the other sections again
        tab.host.assign(blob.begin(), blob.begin() + sizeof(PredHdr));
        const int64_t extra[3] = {(int64_t)b_kky, (int64_t)bstride, 0};
        tab.host.insert(tab.host.end(), (const unsigned char*)extra, (const unsigned char*)extra + sizeof(extra));
        it = ctx->pil_tables.emplace(key, tab).first;
    }
    unsigned char* d = (unsigned char*)it->second.dptr;
    const size_t b_hdr = (sizeof(PredHdr) + 15) / 16 * 16;
    memcpy(&out->hdr, it->second.host.data(), sizeof(PredHdr));
    int64_t extra[3];
    memcpy(extra, it->second.host.data() + sizeof(PredHdr), sizeof(extra));
    const size_t b_groups = (size_t)out->hdr.total_groups * sizeof(uint4);
    out->d_hdr = (PredHdr*)d;
    out->d_groups = (uint4*)(d + b_hdr);
    out->d_kky = (int*)(d + b_hdr + b_groups);
    out->d_plan = (MmaPlan*)(d + b_hdr + b_groups + (size_t)extra[0]);
    out->d_bfrag = (uint2*)(d + b_hdr + b_groups + (size_t)extra[0] + (sizeof(MmaPlan) + 15) / 16 * 16);
    out->bstride = (int)extra[1];
    return NC_OK;
}

// fuse: when non-NULL and the tensor-core kernel with the fused train-side chain can take this shape, it is launched
// with FUSE and *fused is set; otherwise the predict-side chain alone runs and *fused stays 0.
static int predict_impl(nc_ctx* ctx, const uint8_t* src, int n, int h, int w, int c,
                        int64_t img_stride, int channel_order, int compat, float* out_f32,
                        uint8_t* out_small20, int* status, void* stream, FuseArgs* fuse, int* fused) {
    if (fused) *fused = 0;
    if (!ctx) return nc_set_error(nullptr, NC_EINVAL, "ctx is NULL");
    if (n < 0 || h <= 0 || w <= 0) return nc_set_error(ctx, NC_EINVAL, "bad shape n=%d h=%d w=%d", n, h, w);
    if (c != 1 && c != 3) return nc_set_error(ctx, NC_EINVAL, "channels must be 1 or 3, got %d", c);
    if (img_stride < (int64_t)h * w * c) return nc_set_error(ctx, NC_EINVAL, "img_stride too small");
    if (channel_order != 0 && channel_order != 1) return nc_set_error(ctx, NC_EINVAL, "channel_order must be 0 (RGB) or 1 (BGR)");
    if (!out_f32) return nc_set_error(ctx, NC_EINVAL, "out_f32 is NULL");
    if (compat && !status) return nc_set_error(ctx, NC_EINVAL, "compat mode needs a status buffer");
    if (h > 8192 || w > 8192) return nc_set_error(ctx, NC_EUNSUPPORTED, "pictures larger than 8192 on a side");
    if (n == 0) return NC_OK;
    if (!src) return nc_set_error(ctx, NC_EINVAL, "src is NULL");
    NC_CUDA(ctx, cudaSetDevice(ctx->device));
    PredTabDev tabs;
    int rc = get_pred_tables(ctx, h, w, &tabs);
    if (rc != NC_OK) return rc;

    const int wq_src = (w + 3) / 4;
    int wq = wq_src + 1;
    if ((wq & 1) == 0) wq += 1;          // odd word stride: row pairs hit distinct banks
    const size_t fixed = (size_t)tabs.hdr.total_groups * 16 + ((size_t)h * kOut + 15) / 16 * 16;
    const size_t budget = 200 * 1024;    // leave room for static smem; <= 227 KB per CTA on sm_100
    if (fixed + 2 * (size_t)wq * 4 > budget)
        return nc_set_error(ctx, NC_EUNSUPPORTED, "picture %dx%d does not fit the shared-memory plan", h, w);
    int rb = (int)((budget - fixed) / ((size_t)wq * 4));
    rb = rb > 32 ? 32 : rb;
    rb &= ~1;
    if (rb < 2) return nc_set_error(ctx, NC_EUNSUPPORTED, "picture too wide");
    const size_t smem = (size_t)tabs.hdr.total_groups * 16 + (size_t)rb * wq * 4 + ((size_t)h * kOut + 15) / 16 * 16;

    const int cr = 19595, cg = 38470, cb = 7471;
    const int c0 = channel_order == 0 ? cr : cb, c1 = cg, c2 = channel_order == 0 ? cb : cr;
    GrayCoef gc;
    auto plane = [&](int shift, unsigned* o) {
        const unsigned a = (c0 >> shift) & 255, b = (c1 >> shift) & 255, d = (c2 >> shift) & 255;
        o[0] = a | (b << 8) | (d << 16);   // pixel 0 in word 0
        o[1] = a << 24;                    // pixel 1: byte 3 of word 0
        o[2] = b | (d << 8);               //          bytes 0,1 of word 1
        o[3] = (a << 16) | (b << 24);      // pixel 2: bytes 2,3 of word 1
        o[4] = d;                          //          byte 0 of word 2
        o[5] = (a << 8) | (b << 16) | (d << 24);  // pixel 3: bytes 1..3 of word 2
    };
    plane(0, gc.lo);
    plane(8, gc.hi);
    const unsigned magic = wq_src == 1 ? 0u : (unsigned)((0x100000000ull + (unsigned)wq_src - 1) / (unsigned)wq_src);
    cudaStream_t st = (cudaStream_t)stream;
    // TMA-staged kernel: 3 channels, rows and pictures 16-byte aligned, the plan fits shared memory
    if (c == 3 && (w % 16) == 0 && (reinterpret_cast<uintptr_t>(src) % 16) == 0 && (img_stride % 16) == 0 && h >= 16 &&
        !getenv("NC_PREDICT_NO_TMA")) {
        // tensor-core horizontal pass: needs the warp's K range in <= 16 chunks (w <= ~2000) and a stage plan that fits
        if (!getenv("NC_PREDICT_NO_MMA") && tabs.bstride <= 16) {
            const unsigned ipr = (unsigned)w / 16;
            int pitch = w / 4 + 8;                       // >= 8 words of slack: the last chunk may overrun the row by < 32 pixels
            while ((pitch & 7) != 4) pitch += 4;         // = 4 mod 8 words: 16-byte item stores, and the 8 rows x 16 bytes of an
                                                         // ldmatrix 8x8 matrix fall on 8 distinct 16-byte bank groups
            // rows per stage: a multiple of the 320-thread item walk when possible (w = 1024: 10 rows = 2 items per thread)
            int sr = 16;
            while (sr > 1 && (size_t)sr * w * 3 > 32 * 1024) --sr;
            int best = sr; double best_eff = 0.0;
            for (int r = sr; r >= (sr + 1) / 2 && r >= 1; --r) {
                const unsigned items = (unsigned)r * ipr;
                const double eff = (double)items / (double)(((items + kThreads - 1) / kThreads) * kThreads);
                if (eff > best_eff + 1e-9) { best_eff = eff; best = r; }
            }
            sr = best;
            const size_t smem3 = 2 * (size_t)sr * w * 3 + (size_t)(sr + 16) * pitch * 4 + ((size_t)h * kOut + 15) / 16 * 16;
            if (smem3 <= 220 * 1024 && ((size_t)sr * w * 3) % 16 == 0) {
                // fused train-side chain: tap rows sorted, and output rows g, g + 2 never within two consecutive stages
                // (so the 8-slot ring of horizontal results is race free, see the kernel)
                // Measured on B200 (profiles/r01p_fused_chains.md): the single pass costs as much issue time as the
                // separate train-side launch costs DRAM time, so it is opt-in (it saves 13 % of the HBM traffic, no time).
                bool do_fuse = fuse != nullptr && !(h == 28 && w == 28) && getenv("NC_FUSED_CHAINS") != nullptr;
                FuseArgs fa;
                memset(&fa, 0, sizeof(fa));
                if (do_fuse) {
                    fa = *fuse;
                    nc_cv_cubic_axis(w, 28, &fa.x);
                    nc_cv_cubic_axis(h, 28, &fa.y);
                    const int* yo = &fa.y.ofs[0][0];
                    for (int i = 0; i + 1 < 112 && do_fuse; ++i) if (yo[i] > yo[i + 1]) do_fuse = false;
                    for (int g = 0; g + 2 < 28 && do_fuse; ++g) if (yo[4 * g + 8] - yo[4 * g + 3] < 2 * sr) do_fuse = false;
                }

                const unsigned A01 = (unsigned)c0 | ((unsigned)c1 << 16), A2_ = (unsigned)c2, A_0 = (unsigned)c0 << 16,
                               A12 = (unsigned)c1 | ((unsigned)c2 << 16);
#define NC_LAUNCH_MMA_F(NCH, F)                                                                                          \
    do {                                                                                                                 \
        NC_CUDA(ctx, cudaFuncSetAttribute(predict_chain_mma_kernel<NCH, F>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                          (int)smem3));                                                                  \
        predict_chain_mma_kernel<NCH, F><<<n, (F) ? kThreads + 32 : kThreads, smem3, st>>>(src, img_stride, h, w, tabs.d_hdr, tabs.d_plan,     \
                                                                      tabs.d_bfrag, tabs.d_kky, A01, A2_, A_0, A12,      \
                                                                      compat, out_f32, out_small20, status, pitch, sr,   \
                                                                      ipr, fa);                                          \
    } while (0)
#define NC_LAUNCH_MMA(NCH)                                                                                               \
    do {                                                                                                                 \
        if (do_fuse) NC_LAUNCH_MMA_F(NCH, true);                                                                         \
        else NC_LAUNCH_MMA_F(NCH, false);                                                                                \
    } while (0)
                switch (tabs.bstride) {                   // chunks per warp = registers held, so one instantiation each
                    case 1: NC_LAUNCH_MMA(1); break;   case 2: NC_LAUNCH_MMA(2); break;   case 3: NC_LAUNCH_MMA(3); break;
                    case 4: NC_LAUNCH_MMA(4); break;   case 5: NC_LAUNCH_MMA(5); break;   case 6: NC_LAUNCH_MMA(6); break;
                    case 7: NC_LAUNCH_MMA(7); break;   case 8: NC_LAUNCH_MMA(8); break;   case 9: NC_LAUNCH_MMA(9); break;
                    case 10: NC_LAUNCH_MMA(10); break; case 11: NC_LAUNCH_MMA(11); break; case 12: NC_LAUNCH_MMA(12); break;
                    case 13: NC_LAUNCH_MMA(13); break; case 14: NC_LAUNCH_MMA(14); break; case 15: NC_LAUNCH_MMA(15); break;
                    default: NC_LAUNCH_MMA(16); break;
                }
                if (do_fuse && fused) *fused = 1;
#undef NC_LAUNCH_MMA_F
#undef NC_LAUNCH_MMA
                NC_LAUNCH_CHECK(ctx);
                return NC_OK;
            }
        }
        const int wq2 = w / 4 + 2;                 // even, = 2 mod 4: 8-byte stores, row pairs on distinct banks
        const size_t stage = (size_t)kStageRows * w * 3;
        const size_t smem2 = 2 * stage + (size_t)tabs.hdr.total_groups * 16 + (size_t)16 * wq2 * 4 + ((size_t)h * kOut + 15) / 16 * 16;
        if (smem2 <= 220 * 1024) {
            const unsigned A01 = (unsigned)c0 | ((unsigned)c1 << 16), A2_ = (unsigned)c2, A_0 = (unsigned)c0 << 16,
                           A12 = (unsigned)c1 | ((unsigned)c2 << 16);
            const unsigned ipr = (unsigned)w / 16;
            const unsigned ipr_magic = ipr == 1 ? 0u : (unsigned)((0x100000000ull + ipr - 1) / ipr);
            NC_CUDA(ctx, cudaFuncSetAttribute(predict_chain_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
            predict_chain_tma_kernel<<<n, kThreads, smem2, st>>>(src, img_stride, h, w, tabs.d_hdr, tabs.d_groups, tabs.d_kky,
                                                                  A01, A2_, A_0, A12, compat, out_f32, out_small20, status,
                                                                  wq2, ipr, ipr_magic);
            NC_LAUNCH_CHECK(ctx);
            return NC_OK;
        }
    }
    if (c == 3) {
        NC_CUDA(ctx, cudaFuncSetAttribute(predict_chain_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        predict_chain_kernel<3><<<n, kThreads, smem, st>>>(src, img_stride, h, w, tabs.d_hdr, tabs.d_groups, tabs.d_kky,
                                                           gc, c0, c1, c2, compat, out_f32, out_small20, status, rb, wq,
                                                           (unsigned)wq_src, magic);
    } else {
        NC_CUDA(ctx, cudaFuncSetAttribute(predict_chain_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        predict_chain_kernel<1><<<n, kThreads, smem, st>>>(src, img_stride, h, w, tabs.d_hdr, tabs.d_groups, tabs.d_kky,
                                                           gc, c0, c1, c2, compat, out_f32, out_small20, status, rb, wq,
                                                           (unsigned)wq_src, magic);
    }
    NC_LAUNCH_CHECK(ctx);
    return NC_OK;
}

extern "C" int nc_pp_predict_chain(nc_ctx* ctx, const uint8_t* src, int n, int h, int w, int c,
                                   int64_t img_stride, int channel_order, int compat, float* out_f32,
                                   uint8_t* out_small20, int* status, void* stream) {
    return predict_impl(ctx, src, n, h, w, c, img_stride, channel_order, compat, out_f32, out_small20, status, stream,
                        nullptr, nullptr);
}

extern "C" int nc_pp_both_chains(nc_ctx* ctx, const uint8_t* src, int n, int h, int w, int c, int64_t img_stride,
                                 int channel_order, int compat, uint8_t* out_train_u8, float* out_train_f32,
                                 float* out_predict_f32, uint8_t* out_small20, int* status, int* fused, void* stream) {
    if (fused) *fused = 0;
    if (!ctx) return nc_set_error(nullptr, NC_EINVAL, "ctx is NULL");
    if (n == 0) return NC_OK;
    if (!out_train_u8 && !out_train_f32) return nc_set_error(ctx, NC_EINVAL, "both train-side outputs are NULL");
    FuseArgs fa;
    memset(&fa, 0, sizeof(fa));
    fa.out_u8 = out_train_u8;
    fa.out_f32 = out_train_f32;
    int did = 0;
    int rc = predict_impl(ctx, src, n, h, w, c, img_stride, channel_order, compat, out_predict_f32, out_small20, status,
                          stream, &fa, &did);
    if (rc != NC_OK) return rc;
    if (fused) *fused = did;
    if (did) return NC_OK;
    // shape outside the fused plan: the train-side chain as its own launch (same bytes out)
    return nc_pp_resize_gray_norm(ctx, src, n, h, w, c, img_stride, 0, out_train_u8, out_train_f32, stream);
}

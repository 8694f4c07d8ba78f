// Implicit-GEMM convolution whose A operand is loaded by the TMA unit in im2col mode
// (cp.async.bulk.tensor.4d ... .im2col): the hardware walks 128 consecutive output pixels of the NHWC
// tensor, applies the filter-tap offset, zero-fills the padding and writes the 128 x 32-channel
// tile into shared memory already in the 128-byte-swizzled K-major layout tcgen05.mma reads.  No
// thread gathers anything: warp 0 issues TMA, warp 1 issues MMA (one elect.sync thread each),
// warps 2-9 split operands into hi/lo TF32 parts (3xTF32 only; the A operand goes to tensor
// memory) and run the epilogue.
//
// Covers tf.nn.conv2d(x, W, strides=[1,1,1,1], padding='SAME') of conv2d()
// (apps/construction/util/construct_distribute.py:91-95) and its input gradient, for channel
// counts that are multiples of 32; everything else stays on conv_tcgen05.cu (register gather) or
// the SIMT kernels:
//   forward  y[p][co]  = sum_{tap,ci} x[p + tap][ci] . W[tap][ci][co]      A = im2col(x),  B = W as [K][Cout] (MN-major)
//   dgrad    dx[p][ci] = sum_{tap,co} dy[p - tap][co] . W[tap][ci][co]     A = im2col(dy) with the taps mirrored,
//                                                                           B = W[tap] as [Cin][Cout] (K-major)
//   wgrad    dW[tap][ci][co] = sum_p x[p + tap][ci] . dy[p][co]            A = im2col(x) as an MN-major tile (each pixel a
//                                                                           k-row of 32 ci; 4 taps / channel blocks per tile),
//                                                                           B = dy as [pixels][Cout] (MN-major), split-K
#include "nc_internal.h"
#include "tc_gemm.h"
#include "tc_common.cuh"
#include <stdlib.h>

namespace {

constexpr int BM = 128, BK = 32;
constexpr int kThreads = 320;             // 10 warps: TMA, MMA, 8 x (lo split + epilogue)
constexpr unsigned kConvThreads = 256;
// operand ring: 4 raw stages (1x: 96 KB at BN = 64, two CTAs per SM); 3xTF32 (see gemm_tcgen05.cu: the raw tile is the hi
// part, A's hi / lo live in tensor memory) runs 3 stages + 2 B_lo buffers at BN <= 64 so that two CTAs still share an SM
constexpr int LO_SLOTS = 2;              // two: converting block i+1 overlaps the MMAs of block i
#ifndef NC_TMEM_A
#define NC_TMEM_A 1
#endif
// 3xTF32 at BN <= 64: ring depth and number of lo / TMEM-A buffers.  6 + 4 owns the SM (192 KB, 512 TMEM columns);
// 3 + 2 (88 KB, 256 columns) lets two CTAs share it, which hides the per-CTA prologue and epilogue of short-K tiles.
#ifndef NC_I2C_X3_STAGES
#define NC_I2C_X3_STAGES 3
#endif
#ifndef NC_I2C_X3_LOS
#define NC_I2C_X3_LOS 2
#endif
template <int BN, bool X3> struct StageCount { static constexpr int value = (X3 && BN <= 64) ? NC_I2C_X3_STAGES : 4; };

struct I2cArgs {
    int W, H;              // spatial size (input == output: stride 1, SAME)
    int fw, fh;            // filter
    int low_w, low_h;      // coordinate of the window's first tap relative to the output pixel (<= 0)
    int cpb;               // 32-channel blocks of the reduced channel dimension
    int cin_rows;          // dgrad: rows per tap in the [taps*Cin][Cout] view of W
};

__device__ __forceinline__ void tma_load_im2col(unsigned dst, const CUtensorMap* map, int c, int w, int h, int n,
                                                unsigned short off_w, unsigned short off_h, unsigned bar) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6], {%7, %8};"
        ::"r"(dst), "l"(map), "r"(c), "r"(w), "r"(h), "r"(n), "r"(bar), "h"(off_w), "h"(off_h) : "memory");
}

template <int BN, int MODE, bool X3>
__global__ void __launch_bounds__(kThreads, BN <= 64 ? 2 : 1)
conv_i2c_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const I2cArgs ia,
                float* __restrict__ C, const float* __restrict__ bias, int M, int N, int nkb_total, int kb_per_slice,
                float* __restrict__ partial, int act, const __grid_constant__ CUtensorMap tmC, int tstore) {
    constexpr bool DGRAD = MODE == 1, WGRAD = MODE == 2;
    constexpr bool A_MN = WGRAD;                 // wgrad: the im2col rows are k-rows (pixels) of an MN-major tile
    constexpr bool B_MN = !DGRAD;
    constexpr int STAGES = StageCount<BN, X3>::value;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    constexpr unsigned A_BYTES = BM * BK * 4, B_BYTES = BN * BK * 4;
    constexpr unsigned HI_BYTES = A_BYTES + B_BYTES;
    constexpr unsigned STAGE_BYTES = HI_BYTES;                       // raw fp32 tiles [A | B]
    constexpr unsigned LO_BASE = STAGES * STAGE_BYTES;               // 3xTF32: LO_SLOTS buffers [A_lo | B_lo] behind the ring
    // TA: the K-major A operand of the 3xTF32 product lives in tensor memory (see gemm_tcgen05.cu): hi and lo values of a
    // row are parked in TMEM columns [BN, BN + 256) by the split warps, only B and B_lo are read from shared memory
    constexpr bool TA = X3 && !A_MN && NC_TMEM_A;
    constexpr int LOS = TA ? (BN <= 64 ? NC_I2C_X3_LOS : 4) : LO_SLOTS;   // lo buffers in flight: the split warps may run this far ahead of the MMAs
    constexpr unsigned LO_BYTES = TA ? B_BYTES : HI_BYTES;
    constexpr unsigned TCOLS = TA ? ((unsigned)BN + 64u * LOS <= 256u ? 256u : 512u) : (unsigned)BN;
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) unsigned long long s_full[STAGES], s_empty[STAGES], s_ready[4], s_loempty[4], s_tmem_full;
    __shared__ unsigned s_tmem_base;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int kb0 = blockIdx.z * kb_per_slice;
    const int nkb = min(kb_per_slice, nkb_total - kb0);

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(smem_u32(&s_full[s]), 1);
            mbar_init(smem_u32(&s_empty[s]), 1);
        }
        for (int s = 0; s < 4; ++s) {
            mbar_init(smem_u32(&s_ready[s]), kConvThreads / 32);
            mbar_init(smem_u32(&s_loempty[s]), 1);
        }
        mbar_init(smem_u32(&s_tmem_full), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem_base)), "r"(TCOLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = s_tmem_base;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (elect_one()) {
            if (!WGRAD) {
                // first output pixel of the tile; the unit walks W, then H, then N from here
                const int pw = m0 % ia.W, q = m0 / ia.W, ph0 = q % ia.H, pn = q / ia.H;
                int tap = kb0 / ia.cpb, cb = kb0 - tap * ia.cpb;
                int ti = tap / ia.fw, tj = tap - ti * ia.fw;
                for (int i = 0; i < nkb; ++i) {
                    const int s = i % STAGES;
                    const unsigned ph = (unsigned)(i / STAGES) & 1u;
                    mbar_wait(smem_u32(&s_empty[s]), ph ^ 1u);
                    const unsigned full = smem_u32(&s_full[s]);
                    mbar_expect_tx(full, A_BYTES + B_BYTES);
                    const unsigned sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
                    const unsigned sb = sa + A_BYTES;
                    tma_load_im2col(sa, &tmA, cb * 32, pw + ia.low_w, ph0 + ia.low_h, pn, (unsigned short)tj, (unsigned short)ti, full);
                    if (!DGRAD) {
                        const int k0 = (kb0 + i) * BK;                          // W as [taps*Cin][Cout]: k = tap*Cin + ci
#pragma unroll
                        for (int c = 0; c < BN / 32; ++c) tma_load_2d(sb + c * 4096, &tmB, n0 + c * 32, k0, full);
                    } else {
                        // mirrored tap: the window offset (ti, tj) of dy pairs with W[fh-1-ti][fw-1-tj]
                        const int wt = (ia.fh - 1 - ti) * ia.fw + (ia.fw - 1 - tj);
                        tma_load_2d(sb, &tmB, cb * 32, wt * ia.cin_rows + n0, full);      // rows = ci, 32 contiguous co
                    }
                    if (++cb == ia.cpb) { cb = 0; if (++tj == ia.fw) { tj = 0; ++ti; } }
                }
            } else {
                // wgrad: rows of the tile = 4 chunks of 32 (tap, ci) rows, each with its own window offset; the reduction
                // runs over output pixels, 32 per k block
                int cti[BM / 32], ctj[BM / 32], ccb[BM / 32];
                int nvalid = 0;
#pragma unroll
                for (int c = 0; c < BM / 32; ++c) {
                    const int mc = (m0 >> 5) + c;
                    const int tap = mc / ia.cpb;
                    ccb[c] = mc - tap * ia.cpb; cti[c] = tap / ia.fw; ctj[c] = tap - cti[c] * ia.fw;
                    if (mc * 32 < M) nvalid = c + 1;
                }
                long long p = (long long)kb0 * BK;
                int pw = (int)(p % ia.W); long long q = p / ia.W; int ph0 = (int)(q % ia.H), pn = (int)(q / ia.H);
                for (int i = 0; i < nkb; ++i) {
                    const int s = i % STAGES;
                    const unsigned ph = (unsigned)(i / STAGES) & 1u;
                    mbar_wait(smem_u32(&s_empty[s]), ph ^ 1u);
                    const unsigned full = smem_u32(&s_full[s]);
                    mbar_expect_tx(full, (unsigned)nvalid * 4096u + B_BYTES);
                    const unsigned sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
                    const unsigned sb = sa + A_BYTES;
#pragma unroll
                    for (int c = 0; c < BM / 32; ++c)
                        if (c < nvalid)
                            tma_load_im2col(sa + c * 4096, &tmA, ccb[c] * 32, pw + ia.low_w, ph0 + ia.low_h, pn,
                                            (unsigned short)ctj[c], (unsigned short)cti[c], full);
                    const int k0 = (kb0 + i) * BK;
#pragma unroll
                    for (int c = 0; c < BN / 32; ++c) tma_load_2d(sb + c * 4096, &tmB, n0 + c * 32, k0, full);
                    pw += BK;
                    while (pw >= ia.W) { pw -= ia.W; if (++ph0 == ia.H) { ph0 = 0; ++pn; } }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (elect_one()) {
            const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                                   ((unsigned)(BN >> 3) << 17) | ((unsigned)(BM >> 4) << 24);
            for (int i = 0; i < nkb; ++i) {
                const int s = i % STAGES;
                const unsigned ph = (unsigned)(i / STAGES) & 1u;
                const int l = i % LOS;
                if (X3) mbar_wait(smem_u32(&s_ready[l]), (unsigned)(i / LOS) & 1u);   // raw tile landed AND its lo part is written
                else mbar_wait(smem_u32(&s_full[s]), ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const unsigned sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
                const unsigned sb = sa + A_BYTES;
#pragma unroll
                for (int kk = 0; kk < BK / 8; ++kk) {
                    const unsigned long long da = A_MN ? umma_desc(sa + kk * 1024, 4096, 512, 1) : umma_desc(sa + kk * 32, 16, 1024);
                    const unsigned long long db = B_MN ? umma_desc(sb + kk * 1024, 4096, 512, 1) : umma_desc(sb + kk * 32, 16, 1024);
                    const unsigned accum = (i > 0 || kk > 0) ? 1u : 0u;
                    if (TA) {
                        const unsigned a_hi = tmem + (unsigned)BN + (unsigned)l * 64u + (unsigned)kk * 8u, a_lo = a_hi + 32u;
                        const unsigned long long offb = (unsigned long long)((LO_BASE + (unsigned)l * LO_BYTES - ((unsigned)s * STAGE_BYTES + A_BYTES)) >> 4);
                        umma_tf32_ta(tmem, a_lo, db, idesc, accum);          // lo(A) . hi(B)
                        umma_tf32_ta(tmem, a_hi, db + offb, idesc, 1u);      // hi(A) . lo(B)
                        umma_tf32_ta(tmem, a_hi, db, idesc, 1u);             // hi(A) . hi(B)
                    } else if (X3) {
                        // lo tiles: same layout as the raw ones, LO_BASE + l * HI_BYTES - s * STAGE_BYTES further on
                        const unsigned long long off = (unsigned long long)((LO_BASE + (unsigned)l * HI_BYTES - (unsigned)s * STAGE_BYTES) >> 4);
                        umma_tf32(tmem, da + off, db, idesc, accum);         // lo(A) . hi(B)
                        umma_tf32(tmem, da, db + off, idesc, 1u);            // hi(A) . lo(B)
                        umma_tf32(tmem, da, db, idesc, 1u);                  // hi(A) . hi(B): raw bits, truncated by the tensor core
                    } else {
                        umma_tf32(tmem, da, db, idesc, accum);
                    }
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_empty[s])) : "memory");
                if (X3) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_loempty[l])) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_tmem_full)) : "memory");
        }
    } else {
        // ===================== warps 2..9: hi/lo split (3xTF32), then the epilogue =====================
        const int q = warp & 3;
        const int row = m0 + q * 32 + lane;
        for (int i = 0; X3 && i < nkb; ++i) {
            const int s = i % STAGES, l = i % LOS;
            mbar_wait(smem_u32(&s_full[s]), (unsigned)(i / STAGES) & 1u);               // raw tiles landed
            mbar_wait(smem_u32(&s_loempty[l]), ((unsigned)(i / LOS) & 1u) ^ 1u);   // the MMAs that read this lo buffer are done
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const unsigned praw = smem_u32(smem + (size_t)s * STAGE_BYTES);
            const unsigned plo = smem_u32(smem + LO_BASE + (size_t)l * LO_BYTES) - (TA ? A_BYTES : 0u);   // TA: only B has a lo tile
            if (TA) {
                // this thread's tile row (TMEM lane), 16 of its 32 k values: warps 2-5 take k 0..15, warps 6-9 k 16..31.
                // Row r of the 128-byte-swizzled tile keeps its 16-byte chunk c at r * 128 + ((c ^ (r & 7)) << 4).
                const unsigned r = (unsigned)q * 32u + (unsigned)lane, h = warp >= 6 ? 1u : 0u;
                unsigned v[16], w[16];
#pragma unroll
                for (unsigned c = 0; c < 4; ++c)
                    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                                 : "=r"(v[4 * c]), "=r"(v[4 * c + 1]), "=r"(v[4 * c + 2]), "=r"(v[4 * c + 3])
                                 : "r"(praw + r * 128u + (((4u * h + c) ^ (r & 7u)) << 4)));
#pragma unroll
                for (int e = 0; e < 16; ++e) w[e] = __float_as_uint(__uint_as_float(v[e]) - __uint_as_float(v[e] & 0xFFFFE000u));
                const unsigned t_hi = tmem + (((unsigned)q * 32u) << 16) + (unsigned)BN + (unsigned)l * 64u + 16u * h;
                asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
                             ::"r"(t_hi), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
                               "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
                asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
                             ::"r"(t_hi + 32u), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7]), "r"(w[8]),
                               "r"(w[9]), "r"(w[10]), "r"(w[11]), "r"(w[12]), "r"(w[13]), "r"(w[14]), "r"(w[15]) : "memory");
            }
#pragma unroll 3
            for (unsigned j = (TA ? A_BYTES / 16 : 0u) + (unsigned)(warp - 2) * 32u + (unsigned)lane; j < HI_BYTES / 16; j += kConvThreads) {
                // lo = x - trunc_tf32(x): exact in fp32, at most 13 significant bits; the tensor core keeps its top 11
                unsigned vx, vy, vz, vw;
                asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(vx), "=r"(vy), "=r"(vz), "=r"(vw) : "r"(praw + j * 16u));
                const unsigned lx = __float_as_uint(__uint_as_float(vx) - __uint_as_float(vx & 0xFFFFE000u));
                const unsigned ly = __float_as_uint(__uint_as_float(vy) - __uint_as_float(vy & 0xFFFFE000u));
                const unsigned lz = __float_as_uint(__uint_as_float(vz) - __uint_as_float(vz & 0xFFFFE000u));
                const unsigned lw = __float_as_uint(__uint_as_float(vw) - __uint_as_float(vw & 0xFFFFE000u));
                asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(plo + j * 16u), "r"(lx), "r"(ly), "r"(lz), "r"(lw) : "memory");
            }
            if (TA) {
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> tensor-core (async proxy) reads
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&s_ready[l])) : "memory");
        }
        if (nkb > 0) {
            mbar_wait(smem_u32(&s_tmem_full), 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        const int Mpad = (M + BM - 1) / BM * BM;                    // rows per K slice of the partial buffer: whole tiles
        float* out = partial ? partial + (size_t)blockIdx.z * Mpad * N : C;
        // warps 2-5 take the even 32-column chunks, warps 6-9 the odd ones (a warp reads the TMEM lane quarter warp % 4)
#pragma unroll 1
        for (int c0 = (warp >= 6 ? 32 : 0); c0 < BN; c0 += 64) {
            if (n0 + c0 >= N) break;                     // uniform: tile columns past N hold products with foreign W rows
            unsigned r[32];
            if (nkb > 0) {
                const unsigned taddr = tmem + ((unsigned)(q * 32) << 16) + (unsigned)c0;
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                      "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                      "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                      "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                    : "r"(taddr));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) r[j] = 0u;
            }
            if (tstore) {
                // finished tile -> shared memory (the operand ring is idle by now) -> one TMA tensor store per 32-column chunk
                float v[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int col = n0 + c0 + j;
                    float t = __uint_as_float(r[j]);
                    if (!partial) {
                        if (bias && col < N) t += __ldg(bias + col);
                        if (act == 1) t = fmaxf(t, 0.0f);
                    }
                    v[j] = t;
                }
                stage_row_chunk(smem_u32(smem), c0 >> 5, (unsigned)(q * 32 + lane), v);
            } else if (row < M) {
                float* po = out + (size_t)row * N + n0 + c0;
                if (n0 + c0 + 32 <= N && (N & 3) == 0 && (reinterpret_cast<uintptr_t>(po) & 15) == 0 &&
                    (partial || !bias || (reinterpret_cast<uintptr_t>(bias) & 15) == 0)) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        float4 v = make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
                        if (!partial && bias) {
                            const float4 bv = __ldg(reinterpret_cast<const float4*>(bias + n0 + c0 + j));
                            v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
                        }
                        if (!partial && act == 1) { v.x = fmaxf(v.x, 0.0f); v.y = fmaxf(v.y, 0.0f); v.z = fmaxf(v.z, 0.0f); v.w = fmaxf(v.w, 0.0f); }
                        *reinterpret_cast<float4*>(po + j) = v;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int col = n0 + c0 + j;
                        if (col < N) {
                            float v = __uint_as_float(r[j]);
                            if (!partial && bias) v += __ldg(bias + col);
                            if (!partial && act == 1) v = fmaxf(v, 0.0f);
                            po[j] = v;
                        }
                    }
                }
            }
        }
        if (tstore) store_staged_tile(&tmC, smem_u32(smem), (BN + 31) / 32, n0, (partial ? (int)blockIdx.z * Mpad : 0) + m0, N, warp == 2);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TCOLS));
}

__global__ void i2c_reduce_kernel(float* __restrict__ dst, const float* __restrict__ ws, int splits, long long count,
                                  const float* __restrict__ bias, int ncols, int act, long long slice_stride) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        float s = 0.0f;
        for (int k = 0; k < splits; ++k) s += ws[(long long)k * slice_stride + i];
        if (bias) s += __ldg(bias + (int)(i % ncols));
        if (act == 1) s = fmaxf(s, 0.0f);
        dst[i] = s;
    }
}

static PFN_cuTensorMapEncodeIm2col_v12000 g_encode_i2c = nullptr;

int get_encode_i2c(nc_ctx* ctx) {
    if (g_encode_i2c) return NC_OK;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
        cudaGetLastError();
        return nc_set_error(ctx, NC_EUNSUPPORTED, "cuTensorMapEncodeIm2col is not available from this driver");
    }
    g_encode_i2c = (PFN_cuTensorMapEncodeIm2col_v12000)fn;
    return NC_OK;
}

template <int BN, int MODE, bool X3>
int launch_i2c(nc_ctx* ctx, const CUtensorMap& ma, const CUtensorMap& mb, const I2cArgs& ia, float* C, const float* bias, int M,
               int N, int nkb, int kbps, int splits, float* partial, cudaStream_t st, int act, const CUtensorMap& mc, int tstore) {
    const bool ta = X3 && MODE != 2 && NC_TMEM_A;
    const size_t smem = (size_t)StageCount<BN, X3>::value * (BM * BK * 4 + BN * BK * 4) +
                        (X3 ? (ta ? (size_t)(BN <= 64 ? NC_I2C_X3_LOS : 4) * BN * BK * 4 : (size_t)LO_SLOTS * (BM * BK * 4 + BN * BK * 4)) : 0) + 1024;
    auto kfn = conv_i2c_kernel<BN, MODE, X3>;
    NC_CUDA(ctx, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM, splits);
    kfn<<<grid, kThreads, smem, st>>>(ma, mb, ia, C, bias, M, N, nkb, kbps, partial, act, mc, tstore);
    NC_LAUNCH_CHECK(ctx);
    return NC_OK;
}

}  // namespace

// stride 1, output size == input size (SAME), the channel count of the walked tensor a multiple of 32
bool nc_tc_conv_i2c_supported(int mode, const nck::ConvGeom& g, const float* src, const float* wmat) {
    if (mode < 0 || mode > 2) return false;
    if (getenv("NC_CONV_NO_IM2COL")) return false;
    if (g.sh != 1 || g.sw != 1 || g.oh != g.ih || g.ow != g.iw) return false;
    if ((reinterpret_cast<uintptr_t>(src) & 15) || (reinterpret_cast<uintptr_t>(wmat) & 15)) return false;
    if (g.oc % 4 || g.ic % 4) return false;
    const int walked = mode == 1 ? g.oc : g.ic;          // channels of the tensor the TMA unit walks (x, or dy for dgrad)
    if (walked % 32 || walked > 256 * 32) return false;
    if (g.fh > 16 || g.fw > 16) return false;
    return (mode == 1 ? g.ic : g.oc) >= 8;
}

// mode 0: src = x, wmat = W, out = y;  1: src = dy, wmat = W, out = dx;  2: src = x, wmat = dy, out = dW
int nc_tc_conv_i2c(nc_ctx* ctx, int mode, const nck::ConvGeom& g, int batch, const float* src, const float* wmat, float* out,
                   const float* bias, float* ws, size_t ws_bytes, cudaStream_t st, bool x3, int act) {
    int rc = get_encode(ctx);
    if (rc != NC_OK) return rc;
    rc = get_encode_i2c(ctx);
    if (rc != NC_OK) return rc;
    const bool dgrad = mode == 1, wgrad = mode == 2;
    const int cwalk = dgrad ? g.oc : g.ic;                // channels of the walked tensor
    const int pixels = batch * g.ih * g.iw;
    int M, N, K;
    if (wgrad) { M = g.fh * g.fw * g.ic; N = g.oc; K = pixels; }
    else { M = pixels; N = dgrad ? g.ic : g.oc; K = g.fh * g.fw * cwalk; }
    const int BN = (N <= 32 && !getenv("NC_CONV_NO_BN32")) ? 32 : N <= 64 ? 64 : 128;   // 32: Cout = 32 forward, Cin = 32 dgrad
    int splits, kbps;
    const size_t need = nc_tc_gemm_workspace(ctx->sm_count, M, N, K, &splits, &kbps);
    if (need > ws_bytes) return nc_set_error(ctx, NC_EINVAL, "tcgen05 conv: workspace %zu < %zu", ws_bytes, need);

    I2cArgs ia;
    ia.W = g.iw; ia.H = g.ih; ia.fw = g.fw; ia.fh = g.fh; ia.cpb = cwalk / 32; ia.cin_rows = g.ic;
    // forward / wgrad: the window of output pixel p starts at p - pad_before; dgrad: at p - (f - 1 - pad_before)
    ia.low_w = dgrad ? -(g.fw - 1 - g.pl) : -g.pl;
    ia.low_h = dgrad ? -(g.fh - 1 - g.pt) : -g.pt;

    CUtensorMap ma, mb;
    {
        // NHWC tensor, dims innermost first {C, W, H, N}; the box of base pixels has the size of the output
        cuuint64_t dims[4] = {(cuuint64_t)cwalk, (cuuint64_t)g.iw, (cuuint64_t)g.ih, (cuuint64_t)batch};
        cuuint64_t strides[3] = {(cuuint64_t)cwalk * 4, (cuuint64_t)g.iw * cwalk * 4, (cuuint64_t)g.ih * g.iw * cwalk * 4};
        int lower[2] = {ia.low_w, ia.low_h};
        int upper[2] = {ia.low_w, ia.low_h};              // box size = dim + upper - lower = dim (output == input size)
        cuuint32_t estr[4] = {1u, 1u, 1u, 1u};
        // forward / dgrad: 128 pixels x 32 channels = one K-major A tile; wgrad: 32 pixels x 32 channels = one 32-row
        // chunk of an MN-major A tile (each pixel is a k-row of 32 m), which needs the 32-byte-atom swizzle
        CUresult r = g_encode_i2c(&ma, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(src), dims, strides, lower, upper,
                                  32u /* channels per pixel */, (cuuint32_t)(wgrad ? BK : BM) /* pixels per column */, estr,
                                  CU_TENSOR_MAP_INTERLEAVE_NONE,
                                  wgrad ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return nc_set_error(ctx, NC_ECUDA, "cuTensorMapEncodeIm2col failed (%d)", (int)r);
    }
    if (dgrad) rc = make_map(ctx, &mb, wmat, (long long)g.fh * g.fw * g.ic, g.oc, BN, false);   // [taps*Cin][Cout], K-major boxes
    else rc = make_map(ctx, &mb, wmat, K, N, 32, true);                                          // W or dy as [K][Cout], MN-major boxes
    if (rc != NC_OK) return rc;
    const int nkb = (K + BK - 1) / BK;
    float* partial = splits > 1 ? ws : nullptr;
    CUtensorMap mc = mb;
    const long long Mpad = (long long)(M + BM - 1) / BM * BM;      // K slices of the partial buffer are padded to whole tiles
    float* dstp = partial ? partial : out;
    const int tstore = ((N & 3) == 0 && (reinterpret_cast<uintptr_t>(dstp) & 15) == 0 && !getenv("NC_NO_TMA_STORE")) ? 1 : 0;
    if (tstore) { rc = make_map(ctx, &mc, dstp, partial ? Mpad * splits : M, N, BM, false, N); if (rc != NC_OK) return rc; }
#define NC_I2C(BN_, MODE_)                                                                                                    \
    rc = x3 ? launch_i2c<BN_, MODE_, true>(ctx, ma, mb, ia, out, bias, M, N, nkb, kbps, splits, partial, st, act, mc, tstore)  \
            : launch_i2c<BN_, MODE_, false>(ctx, ma, mb, ia, out, bias, M, N, nkb, kbps, splits, partial, st, act, mc, tstore)
    if (BN == 32) { if (mode == 0) NC_I2C(32, 0); else if (mode == 1) NC_I2C(32, 1); else NC_I2C(32, 2); }
    else if (BN == 64) { if (mode == 0) NC_I2C(64, 0); else if (mode == 1) NC_I2C(64, 1); else NC_I2C(64, 2); }
    else { if (mode == 0) NC_I2C(128, 0); else if (mode == 1) NC_I2C(128, 1); else NC_I2C(128, 2); }
#undef NC_I2C
    if (rc != NC_OK) return rc;
    if (splits > 1) {
        const long long count = (long long)M * N;
        long long blocks = (count + 255) / 256;
        if (blocks > (long long)ctx->sm_count * 8) blocks = (long long)ctx->sm_count * 8;
        i2c_reduce_kernel<<<(int)blocks, 256, 0, st>>>(out, ws, splits, count, bias, N, act, Mpad * N);
        NC_LAUNCH_CHECK(ctx);
    }
    return NC_OK;
}

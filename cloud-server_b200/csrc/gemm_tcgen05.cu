// TF32 tensor-core GEMM for sm_100a: tcgen05.mma with the accumulator in TMEM, operands staged in
// shared memory by TMA (cp.async.bulk.tensor, 128-byte swizzle), mbarrier producer/consumer pipeline.
//
//   C[M][N] = opA(A) . opB(B) (+ bias[N]) (+ activation)           fp32 in, TF32 multiply, fp32 accumulate
//
// It carries the dense contractions of the CNN path -- tf.matmul(x, w) + b of connect_layer() and the
// output layer (apps/construction/util/construct_distribute.py:182, :264) and their two gradients --
// straight from the natural row-major buffers: both operands may be "K-major" (reduction dimension
// contiguous) or "MN-major" (the other dimension contiguous), selected in the UMMA instruction
// descriptor, so no transposed copies of weights or activations are ever made:
//   forward   Y[n][out]  = X[n][in] . W[in][out]          A K-major, B MN-major
//   dgrad     dX[n][in]  = dY[n][out] . W[in][out]^T      A K-major, B K-major
//   wgrad     dW[in][out] = X[n][in]^T . dY[n][out]       A MN-major, B MN-major (reduction over the batch)
//
// Tile: 128 x BN x 32 (BN = 64 or 128), a 4-6 deep TMA ring, one CTA per output tile and K slice (deterministic
// split-K through a workspace, summed in slice order).  Warp roles (320 threads): warp 0 = TMA producer and warp 1 =
// MMA issuer + TMEM allocator, each running its loop in ONE thread chosen by elect.sync (so ptxas emits the UTCHMMAs
// back to back); warps 2-9 = 3xTF32 split (lo = x - trunc(x) into shared memory, or hi / lo of a K-major A into tensor
// memory), then the epilogue (tcgen05.ld 32 lanes x 32 columns per warp -> bias / activation -> global).
#include "nc_internal.h"
#include "tc_gemm.h"
#include "tc_common.cuh"
#include <stdlib.h>

namespace {

constexpr int BM = 128, BK = 32;          // BK * 4 B = 128 B = one swizzle-128B row
constexpr int kThreads = 320;             // 10 warps: TMA, MMA, 8 x (lo split + epilogue)
constexpr unsigned kConvThreads = 256;
// Operand ring.  1x: 3 stages of raw fp32 tiles.  3xTF32: the raw tile doubles as the "hi" part (kind::tf32 reads only the
// top 19 bits of an fp32 operand, i.e. hi = trunc_tf32(x) costs nothing), so only lo = x - trunc(x) is materialised, into
// a separate pair of buffers; the ring itself stays read-only for the converter warps, so the TMA runs 4 blocks ahead.
template <int BN, bool X3> struct StageCount { static constexpr int value = (X3 && BN == 64) ? 6 : 4; };
constexpr int LO_SLOTS = 2;              // two: converting block i+1 overlaps the MMAs of block i
#ifndef NC_TMEM_A
#define NC_TMEM_A 1
#endif

// X3 = true: "3xTF32" -- every fp32 operand is split into hi = rna_tf32(x) and lo = rna_tf32(x - hi) and
// the product is accumulated as hi.hi + lo.hi + hi.lo (three MMAs per k step), which restores fp32-class
// accuracy (error ~2^-21 per product) on the TF32 tensor pipe.  X3 = false: one rounded TF32 product.
template <int BN, bool A_MN, bool B_MN, bool X3>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tf32_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, float* __restrict__ C,
                 const float* __restrict__ bias, int M, int N, int K, int kb_per_slice, int act, float slope,
                 float* __restrict__ partial, const __grid_constant__ CUtensorMap tmC, int tstore) {
    constexpr int STAGES = StageCount<BN, X3>::value;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    constexpr unsigned A_BYTES = BM * BK * 4, B_BYTES = BN * BK * 4;
    constexpr unsigned HI_BYTES = A_BYTES + B_BYTES;                 // TMA destination: raw fp32 tiles [A | B]
    constexpr unsigned STAGE_BYTES = HI_BYTES;
    constexpr unsigned LO_BASE = STAGES * STAGE_BYTES;               // 3xTF32: LO_SLOTS buffers [A_lo | B_lo] behind the ring
    // TA: a K-major A operand of the 3xTF32 product is fed from TENSOR MEMORY instead of shared memory.  Both operands from
    // shared memory make the three MMAs of a k step read 3 x (4 KB + BN x 32 B): at BN = 64 that alone saturates the SM's
    // shared-memory port.  The split warps read each A row once, and park its raw (hi) and lo values in TMEM columns
    // [BN, BN + 256): four buffers of 32 hi + 32 lo columns; lane = tile row, one fp32 per column.
    constexpr bool TA = X3 && !A_MN && NC_TMEM_A;
    constexpr int LOS = TA ? 4 : LO_SLOTS;                           // lo buffers in flight: the split warps may run this far ahead of the MMAs
    constexpr unsigned LO_BYTES = TA ? B_BYTES : HI_BYTES;           // per lo buffer: [B_lo] resp. [A_lo | B_lo]
    constexpr unsigned TCOLS = TA ? 512u : (unsigned)BN;             // TMEM columns (power of two)
    // 1024-byte aligned stage buffers (swizzle atoms are 8 rows x 128 B)
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) unsigned long long s_full[STAGES], s_empty[STAGES], s_ready[4], s_loempty[4], s_tmem_full;
    __shared__ unsigned s_tmem_base;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int nkb_total = (K + BK - 1) / BK;
    const int kb0 = blockIdx.z * kb_per_slice;
    const int nkb = min(kb_per_slice, nkb_total - kb0);

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(smem_u32(&s_full[s]), 1);
            mbar_init(smem_u32(&s_empty[s]), 1);
        }
        for (int s = 0; s < 4; ++s) {
            mbar_init(smem_u32(&s_ready[s]), kConvThreads / 32);       // one arrive per converter warp
            mbar_init(smem_u32(&s_loempty[s]), 1);
        }
        mbar_init(smem_u32(&s_tmem_full), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {   // TMEM: BN fp32 accumulator columns (power of two >= 32)
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
            for (int i = 0; i < nkb; ++i) {
                const int s = i % STAGES;
                const unsigned ph = (unsigned)(i / STAGES) & 1u;
                mbar_wait(smem_u32(&s_empty[s]), ph ^ 1u);
                const unsigned full = smem_u32(&s_full[s]);
                mbar_expect_tx(full, A_BYTES + B_BYTES);
                const unsigned sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
                const unsigned sb = sa + A_BYTES;
                const int k0 = (kb0 + i) * BK;
                if (!A_MN) {
                    tma_load_2d(sa, &tmA, k0, m0, full);                       // [M][K]: box 32 k x 128 rows
                } else {
#pragma unroll
                    for (int c = 0; c < BM / 32; ++c) tma_load_2d(sa + c * 4096, &tmA, m0 + c * 32, k0, full);   // [K][M]: box 32 m x 32 k-rows
                }
                if (!B_MN) {
                    tma_load_2d(sb, &tmB, k0, n0, full);                       // [N][K]: box 32 k x BN rows
                } else {
#pragma unroll
                    for (int c = 0; c < BN / 32; ++c) tma_load_2d(sb + c * 4096, &tmB, n0 + c * 32, k0, full);   // [K][N]
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (elect_one()) {
            // instruction descriptor (cute::UMMA::InstrDescriptor): D = F32, A = B = TF32, majors, N >> 3, M >> 4
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
                for (int kk = 0; kk < BK / 8; ++kk) {       // UMMA_K = 8 for tf32 (32 bytes)
                    // K-major (SWIZZLE_128B): rows of 128 B, 8-row atoms 1024 B apart (SBO); advance 32 B inside the atom.
                    // MN-major 32-bit operands need the 32-byte-granular swizzle (SWIZZLE_128B_BASE32B, type 1;
                    // TMA: SWIZZLE_128B_ATOM_32B): each k is a 128 B row of 32 mn, 4 k-rows = one 512 B atom (SBO),
                    // 32-mn chunks 4096 B apart (LBO); one MMA (K = 8) covers two atoms = 1024 B.
                    const unsigned long long da = A_MN ? umma_desc(sa + kk * 1024, 4096, 512, 1) : umma_desc(sa + kk * 32, 16, 1024);
                    const unsigned long long db = B_MN ? umma_desc(sb + kk * 1024, 4096, 512, 1) : umma_desc(sb + kk * 32, 16, 1024);
                    const unsigned accum = (i > 0 || kk > 0) ? 1u : 0u;
                    if (TA) {
                        // A from tensor memory (8 columns per k step), B and B_lo from shared memory
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
                // frees the smem stage once these MMAs have read it
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_empty[s])) : "memory");
                if (X3) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_loempty[l])) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_tmem_full)) : "memory");
        }
    } else {
        // ===================== warps 2..9: 3xTF32 split, then the epilogue; a warp owns TMEM lanes 32*(warp%4) .. +32 =====================
        const int q = warp & 3;
        const int row = m0 + q * 32 + lane;
        // kind::tf32 TRUNCATES the 13 low mantissa bits of its fp32 operands: alone that is a one-sided error of ~1e-3 per
        // product which does not average out over the dot product (fine for gradients: 1x mode, stated for mode 2).
        // 3xTF32: during the main loop these four warps write lo = x - trunc(x) of every staged operand; the raw tile
        // itself is the hi part.
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
                        else if (act == 2) t = 1.0f / (1.0f + expf(-t));
                        else if (act == 3) t = fmaxf(t, slope * t);
                    }
                    v[j] = t;
                }
                stage_row_chunk(smem_u32(smem), c0 >> 5, (unsigned)(q * 32 + lane), v);
            } else if (row < M) {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int col = n0 + c0 + j;
                    if (col < N) {
                        float v = __uint_as_float(r[j]);
                        if (!partial) {
                            if (bias) v += __ldg(bias + col);
                            if (act == 1) v = fmaxf(v, 0.0f);
                            else if (act == 2) v = 1.0f / (1.0f + expf(-v));
                            else if (act == 3) v = fmaxf(v, slope * v);
                        }
                        out[(size_t)row * N + col] = v;
                    }
                }
            }
        }
        if (tstore) store_staged_tile(&tmC, smem_u32(smem), BN / 32, n0, (partial ? (int)blockIdx.z * Mpad : 0) + m0, N, warp == 2);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TCOLS));
    }
}

__global__ void tc_reduce_kernel(float* __restrict__ dst, const float* __restrict__ ws, int splits, long long count,
                                 const float* __restrict__ bias, int ncols, int act, float slope, long long slice_stride) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        float s = 0.0f;
        for (int k = 0; k < splits; ++k) s += ws[(long long)k * slice_stride + i];
        if (bias) s += __ldg(bias + (int)(i % ncols));
        if (act == 1) s = fmaxf(s, 0.0f);
        else if (act == 2) s = 1.0f / (1.0f + expf(-s));
        else if (act == 3) s = fmaxf(s, slope * s);
        dst[i] = s;
    }
}

template <int BN, bool A_MN, bool B_MN, bool X3>
int launch(nc_ctx* ctx, const CUtensorMap& ma, const CUtensorMap& mb, float* C, const float* bias, int M, int N, int K,
           int kbps, int splits, int act, float slope, float* partial, cudaStream_t st, const CUtensorMap& mc, int tstore) {
    const bool ta = X3 && !A_MN && NC_TMEM_A;
    const size_t smem = (size_t)StageCount<BN, X3>::value * (BM * BK * 4 + BN * BK * 4) +
                        (X3 ? (ta ? (size_t)4 * BN * BK * 4 : (size_t)LO_SLOTS * (BM * BK * 4 + BN * BK * 4)) : 0) + 1024;
    auto kfn = gemm_tf32_kernel<BN, A_MN, B_MN, X3>;
    NC_CUDA(ctx, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM, splits);
    kfn<<<grid, kThreads, smem, st>>>(ma, mb, C, bias, M, N, K, kbps, act, slope, partial, mc, tstore);
    NC_LAUNCH_CHECK(ctx);
    return NC_OK;
}

}  // namespace

bool nc_tc_gemm_supported(const float* A, const float* B, int M, int N, int K, bool a_mn, bool b_mn) {
    // TMA needs 16-byte aligned bases and row pitches
    const long long lda = a_mn ? M : K, ldb = b_mn ? N : K;
    if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15)) return false;
    if ((lda & 3) || (ldb & 3)) return false;
    return M >= 1 && N >= 8 && K >= 8;
}

size_t nc_tc_gemm_workspace(int sm_count, int M, int N, int K, int* splits_out, int* kbps_out) {
    const int BN = N <= 64 ? 64 : 128;
    const long long tiles = (long long)((M + BM - 1) / BM) * ((N + BN - 1) / BN);
    const int nkb = (K + BK - 1) / BK;
    // Split-K by a small cost model in units of one k-block iteration (~0.4 us): whole waves of CTAs over the resident slots
    // (narrow outputs, N <= 64: two CTAs per SM) x (k blocks per CTA + the fixed cost of a CTA: TMEM allocation, ring fill,
    // epilogue) + the reduce pass over the partial slices.  Counting waves is the point: 7 tiles x 43 slices = 301 CTAs on
    // 296 slots, or 16 x 10 = 160 on 148, run as two waves and take twice as long as 294 or 144.
    const long long slots = N <= 64 ? 2LL * sm_count : sm_count;
    const int max_splits = nkb >= 8 ? (nkb / 4 < 64 ? nkb / 4 : 64) : 1;
    int splits = 1;
    double best = 0.0;
    for (int s = 1; s <= max_splits; ++s) {
        const int kb = (nkb + s - 1) / s;
        const int s_eff = (nkb + kb - 1) / kb;
        if (s_eff != s) continue;
        const long long waves = (tiles * s + slots - 1) / slots;
        double cost = (double)waves * (kb + 6);
        if (s > 1) cost += (3.0 + (double)(s + 1) * M * N * 4.0 / 5.0e6) / 0.4;      // launch + slices read once, result written once
        if (s == 1 || cost < best * 0.97) { best = cost; splits = s; }               // a split has to pay for itself
    }
    int kbps = (nkb + splits - 1) / splits;
    splits = (nkb + kbps - 1) / kbps;
    if (splits_out) *splits_out = splits;
    if (kbps_out) *kbps_out = kbps;
    return splits > 1 ? (size_t)splits * ((M + BM - 1) / BM * BM) * N * sizeof(float) : 0;        // slices padded to whole tiles
}

int nc_tc_gemm(nc_ctx* ctx, const float* A, const float* B, float* C, const float* bias, int M, int N, int K, bool a_mn,
               bool b_mn, int act, float slope, float* ws, size_t ws_bytes, cudaStream_t st, bool x3) {
    if (!nc_tc_gemm_supported(A, B, M, N, K, a_mn, b_mn)) return nc_set_error(ctx, NC_EUNSUPPORTED, "tcgen05 GEMM: unaligned operands");
    return nc_tc_gemm_ex(ctx, A, B, C, bias, M, N, K, a_mn, b_mn, 0, 0, K, K, act, slope, ws, ws_bytes, st, x3);
}

// lda / ldb: row pitch in floats (0 = natural); ka / kb: how many of the K reduction rows/columns really exist
// in A / B (TMA zero-fills the rest)
int nc_tc_gemm_ex(nc_ctx* ctx, const float* A, const float* B, float* C, const float* bias, int M, int N, int K, bool a_mn,
                  bool b_mn, int lda, int ldb, int ka, int kb, int act, float slope, float* ws, size_t ws_bytes, cudaStream_t st,
                  bool x3) {
    int rc = get_encode(ctx);
    if (rc != NC_OK) return rc;
    const int BN = N <= 64 ? 64 : 128;
    int splits, kbps;
    const size_t need = nc_tc_gemm_workspace(ctx->sm_count, M, N, K, &splits, &kbps);
    if (need > ws_bytes) return nc_set_error(ctx, NC_EINVAL, "tcgen05 GEMM: workspace %zu < %zu", ws_bytes, need);
    CUtensorMap ma, mb;
    rc = a_mn ? make_map(ctx, &ma, A, ka, M, 32, true, lda) : make_map(ctx, &ma, A, M, ka, BM, false, lda);
    if (rc != NC_OK) return rc;
    rc = b_mn ? make_map(ctx, &mb, B, kb, N, 32, true, ldb) : make_map(ctx, &mb, B, N, kb, BN, false, ldb);
    if (rc != NC_OK) return rc;
    float* partial = splits > 1 ? ws : nullptr;
    // un-split outputs leave through shared memory and a TMA tensor store (row pitch and base 16-byte aligned)
    CUtensorMap mc = ma;
    // (split-K partials too: their slices are padded to whole tiles, so a tile's rows never reach into the next slice)
    const long long Mpad = (long long)(M + BM - 1) / BM * BM;
    float* dstp = partial ? partial : C;
    const int tstore = ((N & 3) == 0 && (reinterpret_cast<uintptr_t>(dstp) & 15) == 0 && !getenv("NC_NO_TMA_STORE")) ? 1 : 0;
    if (tstore) { rc = make_map(ctx, &mc, dstp, partial ? Mpad * splits : M, N, BM, false, N); if (rc != NC_OK) return rc; }
#define NC_TC_LAUNCH(BN_, AM, BMJ)                                                                                          \
    rc = x3 ? launch<BN_, AM, BMJ, true>(ctx, ma, mb, C, bias, M, N, K, kbps, splits, act, slope, partial, st, mc, tstore)    \
            : launch<BN_, AM, BMJ, false>(ctx, ma, mb, C, bias, M, N, K, kbps, splits, act, slope, partial, st, mc, tstore)
    if (BN == 64) {
        if (!a_mn && !b_mn) NC_TC_LAUNCH(64, false, false);
        else if (!a_mn && b_mn) NC_TC_LAUNCH(64, false, true);
        else if (a_mn && !b_mn) NC_TC_LAUNCH(64, true, false);
        else NC_TC_LAUNCH(64, true, true);
    } else {
        if (!a_mn && !b_mn) NC_TC_LAUNCH(128, false, false);
        else if (!a_mn && b_mn) NC_TC_LAUNCH(128, false, true);
        else if (a_mn && !b_mn) NC_TC_LAUNCH(128, true, false);
        else NC_TC_LAUNCH(128, true, true);
    }
#undef NC_TC_LAUNCH
    if (rc != NC_OK) return rc;
    if (splits > 1) {
        const long long count = (long long)M * N;
        long long blocks = (count + 255) / 256;
        if (blocks > (long long)ctx->sm_count * 8) blocks = (long long)ctx->sm_count * 8;
        tc_reduce_kernel<<<(int)blocks, 256, 0, st>>>(C, ws, splits, count, bias, N, act, slope, Mpad * N);
        NC_LAUNCH_CHECK(ctx);
    }
    return NC_OK;
}

// tf.matmul(a, b) (+ bias): construct_distribute.py:182, :264
extern "C" int nc_matmul(nc_ctx* ctx, const float* a, const float* b, const float* bias, int m, int k, int n, int trans_a,
                         int trans_b, int precision, float* out, void* stream) {
    if (!ctx) return nc_set_error(nullptr, NC_EINVAL, "ctx is NULL");
    if (!a || !b || !out || m <= 0 || k <= 0 || n <= 0) return nc_set_error(ctx, NC_EINVAL, "bad arguments");
    NC_CUDA(ctx, cudaSetDevice(ctx->device));
    // row-major a: [m][k] (or [k][m] if trans_a), b: [k][n] (or [n][k] if trans_b)
    const bool a_mn = trans_a != 0, b_mn = trans_b == 0;
    size_t need = nc_tc_gemm_workspace(ctx->sm_count, m, n, k, nullptr, nullptr);
    if (need) { int rc = nc_ws_reserve(ctx, need); if (rc != NC_OK) return rc; }
    return nc_tc_gemm(ctx, a, b, out, bias, m, n, k, a_mn, b_mn, 0, 0.0f, (float*)ctx->ws, ctx->ws_bytes, (cudaStream_t)stream,
                      precision != 1);
}

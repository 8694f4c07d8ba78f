// Implicit-GEMM convolution on the tcgen05 tensor cores (TF32 / 3xTF32, fp32 accumulate in TMEM).
//
// Carries tf.nn.conv2d of conv2d() (apps/construction/util/construct_distribute.py:91-118) and its two
// gradients (Conv2DBackpropInput / Conv2DBackpropFilter) for layers whose reduced channel count is a
// multiple of 32: a 32-channel slice of one filter tap is exactly one 128-byte row of the NHWC tensor,
// i.e. one swizzle-128B row of the UMMA operand tile, so the im2col matrix is never materialised:
//   forward  Y[(n,oy,ox)][co] = sum_(i,j,c)  X[n,oy*s-p+i,ox*s-p+j,c] . W[i][j][c][co]
//            A rows gathered (K-major), B = W as [K][Cout] by TMA (MN-major)
//   dgrad    dX[(n,y,x)][c]   = sum_(i,j,co) dY[n,(y+p-i)/s,(x+p-j)/s,co] . W[i][j][c][co]
//            A rows gathered from dY (K-major), B = W[tap] as [Cin][Cout] by TMA (K-major)
//   wgrad    dW[(i,j,c)][co]  = sum_(n,oy,ox) X[n,oy*s-p+i,ox*s-p+j,c] . dY[(n,oy,ox)][co]
//            A = gathered rows placed as k-rows of an MN-major tile, B = dY by TMA (MN-major);
//            the reduction runs over all output positions: deterministic split-K through a workspace
// Warps 2-5 (128 threads) gather: each thread owns ONE 128-byte row of the A tile per k block, loads it
// with 8 x LDG.128 (the next block's row is prefetched into registers), rounds it to TF32 (hi, and lo
// for 3xTF32) and stores it into the swizzled UMMA layout; the same warps round the TMA-loaded B tile.
#include "nc_internal.h"
#include "tc_gemm.h"
#include "tc_common.cuh"

namespace {

constexpr int BM = 128, BK = 32, STAGES = 3, kThreads = 192;
enum { CM_FWD = 0, CM_DGRAD = 1, CM_WGRAD = 2 };

struct ConvArgs {
    nck::ConvGeom g;
    const float* src;     // gathered tensor: x (fwd, wgrad) or dy (dgrad)
    int batch;
};

__device__ __forceinline__ unsigned rna(float v) {
    unsigned r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return r;
}

// address of the 128-byte source row for A-tile row `t` of k block `kb` (nullptr = zero row)
template <int MODE>
__device__ __forceinline__ const float* gather_row(const ConvArgs& a, int t, int m0, int kb, int M, int K) {
    const nck::ConvGeom& g = a.g;
    if (MODE == CM_FWD) {
        const int m = m0 + t;
        if (m >= M) return nullptr;
        const int cpb = g.ic >> 5;                        // 32-channel blocks per tap
        const int tap = kb / cpb, cb = kb - tap * cpb;
        const int i = tap / g.fw, j = tap - i * g.fw;
        const int ox = m % g.ow; const int q = m / g.ow; const int oy = q % g.oh, n = q / g.oh;
        const int y = oy * g.sh - g.pt + i, x = ox * g.sw - g.pl + j;
        if (y < 0 || y >= g.ih || x < 0 || x >= g.iw) return nullptr;
        return a.src + (((long long)n * g.ih + y) * g.iw + x) * g.ic + cb * 32;
    } else if (MODE == CM_DGRAD) {
        const int m = m0 + t;
        if (m >= M) return nullptr;
        const int cpb = g.oc >> 5;
        const int tap = kb / cpb, cb = kb - tap * cpb;
        const int i = tap / g.fw, j = tap - i * g.fw;
        const int x = m % g.iw; const int q = m / g.iw; const int y = q % g.ih, n = q / g.ih;
        const int ny = y + g.pt - i, nx = x + g.pl - j;
        if (ny < 0 || nx < 0 || ny % g.sh || nx % g.sw) return nullptr;
        const int oy = ny / g.sh, ox = nx / g.sw;
        if (oy >= g.oh || ox >= g.ow) return nullptr;
        return a.src + (((long long)n * g.oh + oy) * g.ow + ox) * g.oc + cb * 32;
    } else {
        // wgrad: tile row t = (32-row chunk of dW rows, k-row); k = output position
        const int chunk = t >> 5, krow = t & 31;
        const int mc = (m0 >> 5) + chunk;                 // which 32-channel block of which tap
        if (mc * 32 >= M) return nullptr;
        const int p = kb * 32 + krow;                     // output position (n, oy, ox)
        if (p >= K) return nullptr;
        const int cpb = g.ic >> 5;
        const int tap = mc / cpb, cb = mc - tap * cpb;
        const int i = tap / g.fw, j = tap - i * g.fw;
        const int ox = p % g.ow; const int q = p / g.ow; const int oy = q % g.oh, n = q / g.oh;
        const int y = oy * g.sh - g.pt + i, x = ox * g.sw - g.pl + j;
        if (y < 0 || y >= g.ih || x < 0 || x >= g.iw) return nullptr;
        return a.src + (((long long)n * g.ih + y) * g.iw + x) * g.ic + cb * 32;
    }
}

template <int BN, int MODE, bool X3>
__global__ void __launch_bounds__(kThreads, 1)
conv_tf32_kernel(const __grid_constant__ ConvArgs args, const __grid_constant__ CUtensorMap tmB, float* __restrict__ C,
                 const float* __restrict__ bias, int M, int N, int K, int kb_per_slice, float* __restrict__ partial) {
    constexpr bool A_MN = MODE == CM_WGRAD;      // gathered rows are k-rows of an MN-major tile
    constexpr bool B_MN = MODE != CM_DGRAD;      // W as [K][Cout] / dY as [K][Cout]; dgrad reads W[tap] as [Cin][Cout] K-major
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    constexpr unsigned A_BYTES = BM * BK * 4, B_BYTES = BN * BK * 4;
    constexpr unsigned HI_BYTES = A_BYTES + B_BYTES;
    constexpr unsigned STAGE_BYTES = X3 ? 2 * HI_BYTES : HI_BYTES;
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) unsigned long long s_full[STAGES], s_empty[STAGES], s_ready[STAGES], s_tmem_full;
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
            mbar_init(smem_u32(&s_ready[s]), 4);
        }
        mbar_init(smem_u32(&s_tmem_full), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem_base)), "r"((unsigned)BN));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = s_tmem_base;

    if (warp == 0) {
        // ===================== TMA producer: B tiles only =====================
        if (lane == 0) {
            for (int i = 0; i < nkb; ++i) {
                const int s = i % STAGES;
                const unsigned ph = (unsigned)(i / STAGES) & 1u;
                mbar_wait(smem_u32(&s_empty[s]), ph ^ 1u);
                const unsigned full = smem_u32(&s_full[s]);
                mbar_expect_tx(full, B_BYTES);
                const unsigned sb = smem_u32(smem + (size_t)s * STAGE_BYTES) + A_BYTES;
                const int kb = kb0 + i;
                if (MODE == CM_DGRAD) {
                    // W[tap][c][co]: rows tap*Cin + n0 .., columns cb*32 ..   (K-major: co is the reduction index)
                    const int cpb = args.g.oc >> 5;
                    const int tap = kb / cpb, cb = kb - tap * cpb;
                    tma_load_2d(sb, &tmB, cb * 32, tap * args.g.ic + n0, full);
                } else {
#pragma unroll
                    for (int c = 0; c < BN / 32; ++c) tma_load_2d(sb + c * 4096, &tmB, n0 + c * 32, kb * BK, full);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                                   ((unsigned)(BN >> 3) << 17) | ((unsigned)(BM >> 4) << 24);
            for (int i = 0; i < nkb; ++i) {
                const int s = i % STAGES;
                const unsigned ph = (unsigned)(i / STAGES) & 1u;
                mbar_wait(smem_u32(&s_ready[s]), ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const unsigned sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
                const unsigned sb = sa + A_BYTES;
#pragma unroll
                for (int kk = 0; kk < BK / 8; ++kk) {
                    const unsigned long long da = A_MN ? umma_desc(sa + kk * 1024, 4096, 512, 1) : umma_desc(sa + kk * 32, 16, 1024);
                    const unsigned long long db = B_MN ? umma_desc(sb + kk * 1024, 4096, 512, 1) : umma_desc(sb + kk * 32, 16, 1024);
                    const unsigned accum = (i > 0 || kk > 0) ? 1u : 0u;
                    if (X3) {
                        const unsigned long long off = (unsigned long long)(HI_BYTES >> 4);
                        umma_tf32(tmem, da + off, db, idesc, accum);
                        umma_tf32(tmem, da, db + off, idesc, 1u);
                        umma_tf32(tmem, da, db, idesc, 1u);
                    } else {
                        umma_tf32(tmem, da, db, idesc, accum);
                    }
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_empty[s])) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_tmem_full)) : "memory");
        }
    } else {
        // ===================== gather + round (main loop), then epilogue =====================
        const int t = threadIdx.x - 64;              // 0..127: the A-tile row this thread owns
        const int q = warp & 3;
        uint4 cur[8];
        {
            const float* p = nkb > 0 ? gather_row<MODE>(args, t, m0, kb0, M, K) : nullptr;
#pragma unroll
            for (int c = 0; c < 8; ++c) cur[c] = p ? __ldg(reinterpret_cast<const uint4*>(p) + c) : make_uint4(0, 0, 0, 0);
        }
        for (int i = 0; i < nkb; ++i) {
            const int s = i % STAGES;
            const unsigned ph = (unsigned)(i / STAGES) & 1u;
            // prefetch the next k block's row while this one is rounded and stored
            uint4 nxt[8];
            {
                const float* p = (i + 1 < nkb) ? gather_row<MODE>(args, t, m0, kb0 + i + 1, M, K) : nullptr;
#pragma unroll
                for (int c = 0; c < 8; ++c) nxt[c] = p ? __ldg(reinterpret_cast<const uint4*>(p) + c) : make_uint4(0, 0, 0, 0);
            }
            mbar_wait(smem_u32(&s_full[s]), ph);       // B landed => the stage is free and B can be rounded
            unsigned char* st = smem + (size_t)s * STAGE_BYTES;
            // ---- A row -> swizzled UMMA layout
            unsigned char* arow;
            if (!A_MN) arow = st + t * 128;                                  // K-major: row t, 16-byte chunks XOR (t & 7)
            else arow = st + (t >> 5) * 4096 + (t & 31) * 128;               // MN-major: chunk t/32, k-row t%32, 32-byte units XOR (krow & 3)
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const unsigned off = !A_MN ? ((unsigned)(c ^ (t & 7)) << 4)
                                           : ((((unsigned)(c >> 1) ^ (unsigned)(t & 3)) << 5) | ((unsigned)(c & 1) << 4));
                const uint4 v = cur[c];
                uint4 hi = make_uint4(rna(__uint_as_float(v.x)), rna(__uint_as_float(v.y)), rna(__uint_as_float(v.z)), rna(__uint_as_float(v.w)));
                *reinterpret_cast<uint4*>(arow + off) = hi;
                if (X3) {
                    uint4 lo = make_uint4(rna(__uint_as_float(v.x) - __uint_as_float(hi.x)), rna(__uint_as_float(v.y) - __uint_as_float(hi.y)),
                                          rna(__uint_as_float(v.z) - __uint_as_float(hi.z)), rna(__uint_as_float(v.w) - __uint_as_float(hi.w)));
                    *reinterpret_cast<uint4*>(arow + HI_BYTES + off) = lo;
                }
            }
            // ---- B tile (TMA, raw fp32) -> rounded in place (+ lo copy)
            uint4* pb = reinterpret_cast<uint4*>(st + A_BYTES);
#pragma unroll 2
            for (int j = t; j < (int)(B_BYTES / 16); j += 128) {
                const uint4 v = pb[j];
                uint4 hi = make_uint4(rna(__uint_as_float(v.x)), rna(__uint_as_float(v.y)), rna(__uint_as_float(v.z)), rna(__uint_as_float(v.w)));
                pb[j] = hi;
                if (X3) {
                    uint4 lo = make_uint4(rna(__uint_as_float(v.x) - __uint_as_float(hi.x)), rna(__uint_as_float(v.y) - __uint_as_float(hi.y)),
                                          rna(__uint_as_float(v.z) - __uint_as_float(hi.z)), rna(__uint_as_float(v.w) - __uint_as_float(hi.w)));
                    pb[j + HI_BYTES / 16] = lo;
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&s_ready[s])) : "memory");
#pragma unroll
            for (int c = 0; c < 8; ++c) cur[c] = nxt[c];
        }
        // ---- epilogue
        const int row = m0 + q * 32 + lane;
        if (nkb > 0) {
            mbar_wait(smem_u32(&s_tmem_full), 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        float* out = partial ? partial + (size_t)blockIdx.z * M * N : C;
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
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
            if (row < M) {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int col = n0 + c0 + j;
                    if (col < N) {
                        float v = __uint_as_float(r[j]);
                        if (!partial && bias) v += __ldg(bias + col);
                        out[(size_t)row * N + col] = v;
                    }
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((unsigned)BN));
}

__global__ void conv_reduce_kernel(float* __restrict__ dst, const float* __restrict__ ws, int splits, long long count,
                                   const float* __restrict__ bias, int ncols) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        float s = 0.0f;
        for (int k = 0; k < splits; ++k) s += ws[(long long)k * count + i];
        if (bias) s += __ldg(bias + (int)(i % ncols));
        dst[i] = s;
    }
}

template <int BN, int MODE, bool X3>
int launch_conv(nc_ctx* ctx, const ConvArgs& a, const CUtensorMap& mb, float* C, const float* bias, int M, int N, int K,
                int kbps, int splits, float* partial, cudaStream_t st) {
    const size_t smem = (size_t)STAGES * (BM * BK * 4 + BN * BK * 4) * (X3 ? 2 : 1) + 1024;
    auto kfn = conv_tf32_kernel<BN, MODE, X3>;
    NC_CUDA(ctx, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM, splits);
    kfn<<<grid, kThreads, smem, st>>>(a, mb, C, bias, M, N, K, kbps, partial);
    NC_LAUNCH_CHECK(ctx);
    return NC_OK;
}

}  // namespace

// mode: 0 forward (src = x, wmat = W, out = y), 1 dgrad (src = dy, wmat = W, out = dx),
//       2 wgrad (src = x, wmat = dy, out = dW)
bool nc_tc_conv_supported(int mode, const nck::ConvGeom& g, const float* src, const float* wmat) {
    if ((reinterpret_cast<uintptr_t>(src) & 15) || (reinterpret_cast<uintptr_t>(wmat) & 15)) return false;
    if (g.oc % 4) return false;                                   // TMA row pitch of W / dY
    if (mode == CM_DGRAD) return g.oc % 32 == 0 && g.ic >= 8;     // reduction over (tap, co) in 32-channel slices
    return g.ic % 32 == 0 && g.oc >= 8;                           // reduction / rows in 32-channel slices of Cin
}

void nc_tc_conv_dims(int mode, const nck::ConvGeom& g, int batch, int* M, int* N, int* K) {
    if (mode == CM_FWD) { *M = batch * g.oh * g.ow; *N = g.oc; *K = g.fh * g.fw * g.ic; }
    else if (mode == CM_DGRAD) { *M = batch * g.ih * g.iw; *N = g.ic; *K = g.fh * g.fw * g.oc; }
    else { *M = g.fh * g.fw * g.ic; *N = g.oc; *K = batch * g.oh * g.ow; }
}

int nc_tc_conv(nc_ctx* ctx, int mode, const nck::ConvGeom& g, int batch, const float* src, const float* wmat, float* out,
               const float* bias, float* ws, size_t ws_bytes, cudaStream_t st, bool x3) {
    int rc = get_encode(ctx);
    if (rc != NC_OK) return rc;
    int M, N, K;
    nc_tc_conv_dims(mode, g, batch, &M, &N, &K);
    const int BN = N <= 64 ? 64 : 128;
    int splits, kbps;
    const size_t need = nc_tc_gemm_workspace(ctx->sm_count, M, N, K, &splits, &kbps);
    if (need > ws_bytes) return nc_set_error(ctx, NC_EINVAL, "tcgen05 conv: workspace %zu < %zu", ws_bytes, need);
    CUtensorMap mb;
    if (mode == CM_DGRAD) rc = make_map(ctx, &mb, wmat, (long long)g.fh * g.fw * g.ic, g.oc, BN, false);   // [taps*Cin][Cout], K-major boxes
    else rc = make_map(ctx, &mb, wmat, K, N, 32, true);                                                       // [K][Cout], MN-major boxes
    if (rc != NC_OK) return rc;
    ConvArgs a;
    a.g = g; a.src = src; a.batch = batch;
    float* partial = splits > 1 ? ws : nullptr;
#define NC_CONV_LAUNCH(BN_, MODE_)                                                                              \
    rc = x3 ? launch_conv<BN_, MODE_, true>(ctx, a, mb, out, bias, M, N, K, kbps, splits, partial, st)            \
            : launch_conv<BN_, MODE_, false>(ctx, a, mb, out, bias, M, N, K, kbps, splits, partial, st)
    if (BN == 64) {
        if (mode == CM_FWD) NC_CONV_LAUNCH(64, CM_FWD);
        else if (mode == CM_DGRAD) NC_CONV_LAUNCH(64, CM_DGRAD);
        else NC_CONV_LAUNCH(64, CM_WGRAD);
    } else {
        if (mode == CM_FWD) NC_CONV_LAUNCH(128, CM_FWD);
        else if (mode == CM_DGRAD) NC_CONV_LAUNCH(128, CM_DGRAD);
        else NC_CONV_LAUNCH(128, CM_WGRAD);
    }
#undef NC_CONV_LAUNCH
    if (rc != NC_OK) return rc;
    if (splits > 1) {
        const long long count = (long long)M * N;
        long long blocks = (count + 255) / 256;
        if (blocks > (long long)ctx->sm_count * 8) blocks = (long long)ctx->sm_count * 8;
        conv_reduce_kernel<<<(int)blocks, 256, 0, st>>>(out, ws, splits, count, bias, N);
        NC_LAUNCH_CHECK(ctx);
    }
    return NC_OK;
}

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
// Warps 2-9 (256 threads) gather: two threads share ONE 128-byte row of the A tile per k block (4 x LDG.128
// each, a register ring keeps the rows of the next two k blocks in flight), and store it into the swizzled UMMA
// layout.  3xTF32 (forward): the row is split into rounded hi + lo parts, and the same warps split the TMA-loaded
// B tile from a raw ring into the operand stage.  1x (gradients): rows are stored as loaded and the MMA reads B
// straight from the TMA ring -- kind::tf32 truncates the operands itself.
#include "nc_internal.h"
#include "tc_gemm.h"
#include "tc_common.cuh"

namespace {

constexpr int BM = 128, BK = 32, STAGES = 2, BRING = 2, kThreads = 320;   // warps: 0 TMA, 1 MMA, 2..9 gather/round/epilogue
// STAGES = operand stages the MMA reads (rounded hi/lo tiles); BRING = raw fp32 B tiles in flight from TMA.  Keeping the TMA
// ring separate takes the B load latency out of the stage-reuse chain (MMA done -> gather/round -> MMA).
constexpr int kGatherThreads = 256;
enum { CM_FWD = 0, CM_DGRAD = 1, CM_WGRAD = 2 };

struct ConvArgs {
    nck::ConvGeom g;
    const float* src;     // gathered tensor: x (fwd, wgrad) or dy (dgrad)
    int batch;
};

// round-to-nearest (ties away) to TF32 in two integer ops: add half an ulp of the 10-bit mantissa and clear
// the 13 low bits (cvt.rna.tf32.f32 expands to a ~5 instruction inf/nan-safe sequence; operands here are finite).
__device__ __forceinline__ unsigned rna(float v) { return (__float_as_uint(v) + 0x1000u) & 0xFFFFE000u; }
// the lo part is only TRUNCATED by the tensor core (error 2^-21 relative to x): no rounding needed
__device__ __forceinline__ unsigned lo_part(float v, unsigned hi) { return __float_as_uint(v - __uint_as_float(hi)); }

// Per-thread iterator over the 128-byte source rows of ONE A-tile row across consecutive k blocks.
// All divisions happen once in init(); advancing to the next k block is a few adds and compares.
template <int MODE>
struct RowIter {
    int n, y, x;          // fwd: (n, oy, ox) of the output row; dgrad: (n, y, x) of the input row; wgrad: running output position
    int i, j, cb, cpb;    // filter tap and 32-channel block (fwd/dgrad: advance with kb; wgrad: fixed per tile row)
    int p;                // wgrad: running output position index
    bool valid;           // the tile row exists at all

    __device__ __forceinline__ void init(const ConvArgs& a, int row, int m0, int kb, int M) {
        const nck::ConvGeom& g = a.g;
        if (MODE == CM_WGRAD) {
            const int chunk = row >> 5, krow = row & 31;
            const int mc = (m0 >> 5) + chunk;
            valid = mc * 32 < M;
            cpb = g.ic >> 5;
            const int tap = mc / cpb;
            cb = mc - tap * cpb; i = tap / g.fw; j = tap - i * g.fw;
            p = kb * 32 + krow;
            x = p % g.ow; const int q = p / g.ow; y = q % g.oh; n = q / g.oh;
        } else {
            const int m = m0 + row;
            valid = m < M;
            const int w = MODE == CM_FWD ? g.ow : g.iw, h = MODE == CM_FWD ? g.oh : g.ih;
            x = m % w; const int q = m / w; y = q % h; n = q / h;
            cpb = (MODE == CM_FWD ? g.ic : g.oc) >> 5;
            const int tap = kb / cpb;
            cb = kb - tap * cpb; i = tap / g.fw; j = tap - i * g.fw;
            p = 0;
        }
    }
    __device__ __forceinline__ const float* ptr(const ConvArgs& a, int K) const {
        const nck::ConvGeom& g = a.g;
        if (!valid) return nullptr;
        if (MODE == CM_DGRAD) {
            const int ny = y + g.pt - i, nx = x + g.pl - j;
            if (ny < 0 || nx < 0 || ny % g.sh || nx % g.sw) return nullptr;
            const int oy = ny / g.sh, ox = nx / g.sw;
            if (oy >= g.oh || ox >= g.ow) return nullptr;
            return a.src + (((long long)n * g.oh + oy) * g.ow + ox) * g.oc + cb * 32;
        }
        if (MODE == CM_WGRAD && p >= K) return nullptr;
        const int yy = y * g.sh - g.pt + i, xx = x * g.sw - g.pl + j;
        if (yy < 0 || yy >= g.ih || xx < 0 || xx >= g.iw) return nullptr;
        return a.src + (((long long)n * g.ih + yy) * g.iw + xx) * g.ic + cb * 32;
    }
    __device__ __forceinline__ void next(const ConvArgs& a) {
        const nck::ConvGeom& g = a.g;
        if (MODE == CM_WGRAD) {
            p += 32; x += 32;
            while (x >= g.ow) { x -= g.ow; if (++y == g.oh) { y = 0; ++n; } }
        } else {
            if (++cb == cpb) { cb = 0; if (++j == g.fw) { j = 0; ++i; } }
        }
    }
};

template <int BN, int MODE, bool X3>
__global__ void __launch_bounds__(kThreads, 2)
conv_tf32_kernel(const __grid_constant__ ConvArgs args, const __grid_constant__ CUtensorMap tmB, float* __restrict__ C,
                 const float* __restrict__ bias, int M, int N, int K, int kb_per_slice, float* __restrict__ partial) {
    constexpr bool A_MN = MODE == CM_WGRAD;      // gathered rows are k-rows of an MN-major tile
    constexpr bool B_MN = MODE != CM_DGRAD;      // W as [K][Cout] / dY as [K][Cout]; dgrad reads W[tap] as [Cin][Cout] K-major
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    constexpr unsigned A_BYTES = BM * BK * 4, B_BYTES = BN * BK * 4;
    constexpr unsigned HI_BYTES = A_BYTES + B_BYTES;
    constexpr unsigned STAGE_BYTES = X3 ? 2 * HI_BYTES : HI_BYTES;
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) unsigned long long s_full[BRING], s_bempty[BRING], s_empty[STAGES], s_ready[STAGES], s_tmem_full;
    unsigned char* braw = smem + (size_t)STAGES * STAGE_BYTES;       // raw B ring
    __shared__ unsigned s_tmem_base;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int nkb_total = (K + BK - 1) / BK;
    const int kb0 = blockIdx.z * kb_per_slice;
    const int nkb = min(kb_per_slice, nkb_total - kb0);

    if (threadIdx.x == 0) {
        for (int s = 0; s < BRING; ++s) {
            mbar_init(smem_u32(&s_full[s]), 1);
            mbar_init(smem_u32(&s_bempty[s]), X3 ? kGatherThreads / 32 : 1);   // 1x: released by the MMA commit
        }
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(smem_u32(&s_empty[s]), 1);
            mbar_init(smem_u32(&s_ready[s]), kGatherThreads / 32);
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
        if (elect_one()) {
            for (int i = 0; i < nkb; ++i) {
                const int s = i % BRING;
                const unsigned ph = (unsigned)(i / BRING) & 1u;
                mbar_wait(smem_u32(&s_bempty[s]), ph ^ 1u);
                const unsigned full = smem_u32(&s_full[s]);
                mbar_expect_tx(full, B_BYTES);
                const unsigned sb = smem_u32(braw + (size_t)s * B_BYTES);
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
        if (elect_one()) {
            const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                                   ((unsigned)(BN >> 3) << 17) | ((unsigned)(BM >> 4) << 24);
            for (int i = 0; i < nkb; ++i) {
                const int s = i % STAGES;
                const unsigned ph = (unsigned)(i / STAGES) & 1u;
                mbar_wait(smem_u32(&s_ready[s]), ph);
                const int sbr = i % BRING;
                if (!X3) mbar_wait(smem_u32(&s_full[sbr]), (unsigned)(i / BRING) & 1u);   // 1x: B is consumed raw, straight from the TMA ring
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const unsigned sa = smem_u32(smem + (size_t)s * STAGE_BYTES);
                const unsigned sb = X3 ? sa + A_BYTES : smem_u32(braw + (size_t)sbr * B_BYTES);
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
                if (!X3) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_bempty[sbr])) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_tmem_full)) : "memory");
        }
    } else {
        // ===================== gather + round (main loop), then epilogue =====================
        const int t2 = threadIdx.x - 64;             // 0..255: two threads per A-tile row
        const int t = t2 >> 1, half = t2 & 1;        // row, and which four 16-byte chunks of it
        const int q = warp & 3;
        RowIter<MODE> it;
        it.init(args, t, m0, kb0, M);
        // register ring: the half rows of k blocks i, i+1, i+2 are in flight / in registers (global latency
        // is ~2 k-block periods; one block of look-ahead left long_scoreboard as the top stall)
        uint4 cur[4], nx1[4], nx2[4];
        auto load_row = [&](uint4 (&dst)[4], bool live) {
            const float* p = live ? it.ptr(args, K) : nullptr;
#pragma unroll
            for (int c = 0; c < 4; ++c) dst[c] = p ? __ldg(reinterpret_cast<const uint4*>(p) + half * 4 + c) : make_uint4(0, 0, 0, 0);
        };
        load_row(cur, nkb > 0);
        it.next(args);
        load_row(nx1, nkb > 1);
        for (int i = 0; i < nkb; ++i) {
            const int s = i % STAGES;
            const unsigned ph = (unsigned)(i / STAGES) & 1u;
            it.next(args);
            load_row(nx2, i + 2 < nkb);
            mbar_wait(smem_u32(&s_empty[s]), ph ^ 1u);  // the MMAs that read this operand stage two blocks ago are done
            unsigned char* st = smem + (size_t)s * STAGE_BYTES;
            // ---- A half row -> swizzled UMMA layout
            unsigned char* arow;
            if (!A_MN) arow = st + t * 128;                                  // K-major: row t, 16-byte chunks XOR (t & 7)
            else arow = st + (t >> 5) * 4096 + (t & 31) * 128;               // MN-major: chunk t/32, k-row t%32, 32-byte units XOR (krow & 3)
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) {
                const int c = half * 4 + cc;
                const unsigned off = !A_MN ? ((unsigned)(c ^ (t & 7)) << 4)
                                           : ((((unsigned)(c >> 1) ^ (unsigned)(t & 3)) << 5) | ((unsigned)(c & 1) << 4));
                const uint4 v = cur[cc];
                // 1x mode stores the fp32 bits as they are: kind::tf32 truncates them, a one-sided ~5e-4 relative shrink
                // that is harmless for gradients (mode 3) and is what the lossy mode 2 states
                uint4 hi = X3 ? make_uint4(rna(__uint_as_float(v.x)), rna(__uint_as_float(v.y)), rna(__uint_as_float(v.z)), rna(__uint_as_float(v.w))) : v;
                *reinterpret_cast<uint4*>(arow + off) = hi;
                if (X3) {
                    uint4 lo = make_uint4(lo_part(__uint_as_float(v.x), hi.x), lo_part(__uint_as_float(v.y), hi.y),
                                          lo_part(__uint_as_float(v.z), hi.z), lo_part(__uint_as_float(v.w), hi.w));
                    *reinterpret_cast<uint4*>(arow + HI_BYTES + off) = lo;
                }
            }
            // ---- B tile (3x only): raw fp32 from the TMA ring -> rounded hi + lo in the operand stage
            const int sbr = i % BRING;
            if (X3) mbar_wait(smem_u32(&s_full[sbr]), (unsigned)(i / BRING) & 1u);
            const uint4* praw = reinterpret_cast<const uint4*>(braw + (size_t)sbr * B_BYTES);
            uint4* pb = reinterpret_cast<uint4*>(st + A_BYTES);
#pragma unroll 2
            for (int jj = t2; X3 && jj < (int)(B_BYTES / 16); jj += kGatherThreads) {
                const uint4 v = praw[jj];
                uint4 hi = make_uint4(rna(__uint_as_float(v.x)), rna(__uint_as_float(v.y)), rna(__uint_as_float(v.z)), rna(__uint_as_float(v.w)));
                pb[jj] = hi;
                if (X3) {
                    uint4 lo = make_uint4(lo_part(__uint_as_float(v.x), hi.x), lo_part(__uint_as_float(v.y), hi.y),
                                          lo_part(__uint_as_float(v.z), hi.z), lo_part(__uint_as_float(v.w), hi.w));
                    pb[jj + HI_BYTES / 16] = lo;
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&s_ready[s])) : "memory");
                if (X3) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&s_bempty[sbr])) : "memory");
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) { cur[c] = nx1[c]; nx1[c] = nx2[c]; }
        }
        // ---- epilogue
        const int row = m0 + q * 32 + lane;
        if (nkb > 0) {
            mbar_wait(smem_u32(&s_tmem_full), 0);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        float* out = partial ? partial + (size_t)blockIdx.z * M * N : C;
        // warps 2-5 take the even 32-column chunks, warps 6-9 the odd ones (each warp reads its own TMEM lane quarter)
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
            if (row < M) {
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
                        *reinterpret_cast<float4*>(po + j) = v;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int col = n0 + c0 + j;
                        if (col < N) {
                            float v = __uint_as_float(r[j]);
                            if (!partial && bias) v += __ldg(bias + col);
                            po[j] = v;
                        }
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
    const size_t smem = (size_t)STAGES * (BM * BK * 4 + BN * BK * 4) * (X3 ? 2 : 1) + (size_t)BRING * BN * BK * 4 + 1024;
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
    // input gradient of a small picture: the padded picture stays in shared memory, one fetch per pixel instead of one per tap
    if (!x3 && !bias && batch >= ctx->sm_count && nc_tc_conv_strip_supported(mode, g, src, wmat, out))     // one picture per SM at least
        return nc_tc_conv_strip_dgrad(ctx, g, batch, src, wmat, out, st);
    // forward / dgrad of a stride-1 SAME convolution: let the TMA unit do the im2col walk (no gather threads)
    if (nc_tc_conv_i2c_supported(mode, g, src, wmat))
        return nc_tc_conv_i2c(ctx, mode, g, batch, src, wmat, out, bias, ws, ws_bytes, st, x3);
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

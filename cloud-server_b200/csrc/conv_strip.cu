// Input gradient of a stride-1 SAME convolution on small pictures with the WHOLE zero-padded picture resident in shared
// memory (conv2d() of apps/construction/util/construct_distribute.py:91-95, backward; LeNet's 14 x 14 x 64 -> 32).
//
// Why: the im2col kernel (conv_im2col.cu) fetches every pixel of dy once per filter tap -- 25 times for 5 x 5 -- and is bound
// by L2 -> shared-memory delivery (profiles/r02_ncu_cnn.md: lts 67 %, tensor pipe 23 %).  Here the TMA unit loads a picture
// once, as a (H + fh - 1) x (W + fw - 1) box of 32-channel pixels with the padding zero-filled, i.e. PH * PW rows of 128 bytes
// in the 128-byte-swizzled K-major layout.  Output "pixels" are numbered in the PADDED width, m = y * PW + x, so the A
// operand of filter tap (ti, tj) is simply the rows m + ti * PW + tj of that box: the same shared memory read through a
// matrix descriptor whose start address is moved by whole rows.  Columns x >= W of every row of m are junk and are not
// stored (22 % of the MMA work at 14 + 4), in exchange for ~8x less operand traffic.
//
// Warps: 0 streams the weights (one filter row of fw 32 x 32 tiles per TMA load and channel panel, a 4-deep ring), 6 loads pictures (one
// buffer per 32-channel panel; the loop runs panel-outer so the next picture's panel p is fetched while panel p + 1 is
// still in use), 1 issues tcgen05.mma (accumulators double-buffered in TMEM), 2-5 store finished pictures.
#include "nc_internal.h"
#include "tc_gemm.h"
#include "tc_common.cuh"
#include <stdlib.h>

namespace {

constexpr int kStripThreads = 224;
constexpr int WSTAGES = 4;              // ring of weight groups: one filter row (fw taps x 32 x 32 floats) per TMA load
constexpr unsigned W_TILE_BYTES = 32 * 32 * 4;

struct StripArgs {
    int W, H, PW, PH;
    int fw, fh;
    int low_w, low_h;      // coordinate of the box origin relative to the picture (<= 0)
    int panels;            // 32-channel panels of dy
    int panel_bytes;       // bytes reserved per panel (multiple of 1024; >= rows addressed * 128)
    int cin_rows;          // rows per tap in the [taps * Cin][Cout] view of W
    int batch;
};

__device__ __forceinline__ void tma_load_4d(unsigned dst, const CUtensorMap* map, int c0, int c1, int c2, int c3, unsigned bar) {
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(bar) : "memory");
}

template <int MB>
__global__ void __launch_bounds__(kStripThreads, 1)
conv_strip_dgrad_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW, const StripArgs a,
                        float* __restrict__ out) {
    constexpr unsigned ACC_COLS = MB * 32;                          // one picture's accumulators: MB blocks x 32 channels
    constexpr unsigned TCOLS = 2 * ACC_COLS <= 64 ? 64u : (2 * ACC_COLS <= 128 ? 128u : 256u);
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) unsigned long long s_img_full[2], s_img_empty[2], s_w_full[WSTAGES], s_w_empty[WSTAGES], s_acc_full[2], s_acc_empty[2];
    __shared__ unsigned s_tmem_base;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned panel0 = smem_u32(smem);
    const unsigned wring = panel0 + (unsigned)a.panels * (unsigned)a.panel_bytes;
    const int taps = a.fh * a.fw;
    const unsigned wgroup = (unsigned)a.fw * W_TILE_BYTES;

    if (threadIdx.x == 0) {
        for (int s = 0; s < 2; ++s) {
            mbar_init(smem_u32(&s_img_full[s]), 1);
            mbar_init(smem_u32(&s_img_empty[s]), 1);
            mbar_init(smem_u32(&s_acc_full[s]), 1);
            mbar_init(smem_u32(&s_acc_empty[s]), 4);
        }
        for (int s = 0; s < WSTAGES; ++s) {
            mbar_init(smem_u32(&s_w_full[s]), 1);
            mbar_init(smem_u32(&s_w_empty[s]), 1);
        }
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

    if (warp == 6) {
        // ===================== picture producer =====================
        if (elect_one()) {
            const unsigned box_bytes = (unsigned)a.PW * (unsigned)a.PH * 128u;
            int k = 0;
            for (int img = blockIdx.x; img < a.batch; img += gridDim.x, ++k) {
                for (int p = 0; p < a.panels; ++p) {
                    mbar_wait(smem_u32(&s_img_empty[p]), ((unsigned)k & 1u) ^ 1u);
                    const unsigned full = smem_u32(&s_img_full[p]);
                    mbar_expect_tx(full, box_bytes);
                    tma_load_4d(panel0 + (unsigned)p * (unsigned)a.panel_bytes, &tmX, p * 32, a.low_w, a.low_h, img, full);
                }
            }
        }
    } else if (warp == 0) {
        // ===================== weight producer =====================
        if (elect_one()) {
            unsigned cnt = 0;
            for (int img = blockIdx.x; img < a.batch; img += gridDim.x) {
                for (int p = 0; p < a.panels; ++p) {
                    for (int ti = 0; ti < a.fh; ++ti, ++cnt) {
                        const unsigned s = cnt % WSTAGES, ph = (cnt / WSTAGES) & 1u;
                        mbar_wait(smem_u32(&s_w_empty[s]), ph ^ 1u);
                        const unsigned full = smem_u32(&s_w_full[s]);
                        mbar_expect_tx(full, wgroup);
                        // mirrored taps: window offset (ti, tj) of dy pairs with W[fh-1-ti][fw-1-tj]; the fw taps of window row ti are
                        // the consecutive row blocks [taps - (ti+1)*fw, taps - ti*fw) of the [taps*Cin][Cout] view, tap tj at block fw-1-tj
                        tma_load_2d(wring + s * wgroup, &tmW, p * 32, (taps - (ti + 1) * a.fw) * a.cin_rows, full);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (elect_one()) {
            const unsigned idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((32u >> 3) << 17) | ((128u >> 4) << 24);
            unsigned cnt = 0;
            int k = 0;
            for (int img = blockIdx.x; img < a.batch; img += gridDim.x, ++k) {
                const unsigned b = (unsigned)k & 1u;
                mbar_wait(smem_u32(&s_acc_empty[b]), (((unsigned)k >> 1) & 1u) ^ 1u);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                for (int p = 0; p < a.panels; ++p) {
                    mbar_wait(smem_u32(&s_img_full[p]), (unsigned)k & 1u);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const unsigned pbase = panel0 + (unsigned)p * (unsigned)a.panel_bytes;
                    for (int ti = 0; ti < a.fh; ++ti, ++cnt) {
                        const unsigned s = cnt % WSTAGES, ph = (cnt / WSTAGES) & 1u;
                        mbar_wait(smem_u32(&s_w_full[s]), ph);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        for (int tj = 0; tj < a.fw; ++tj) {
                            const unsigned sb = wring + s * wgroup + (unsigned)(a.fw - 1 - tj) * W_TILE_BYTES;
                            const unsigned row0 = (unsigned)(ti * a.PW + tj);
#pragma unroll
                            for (int mb = 0; mb < MB; ++mb) {
                                const unsigned sa = pbase + ((unsigned)mb * 128u + row0) * 128u;
#pragma unroll
                                for (int kk = 0; kk < 4; ++kk) {
                                    const unsigned long long da = umma_desc(sa + kk * 32, 16, 1024);
                                    const unsigned long long db = umma_desc(sb + kk * 32, 16, 1024);
                                    umma_tf32(tmem + b * ACC_COLS + (unsigned)mb * 32u, da, db, idesc, (p > 0 || ti > 0 || tj > 0 || kk > 0) ? 1u : 0u);
                                }
                            }
                        }
                        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_w_empty[s])) : "memory");
                    }
                    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_img_empty[p])) : "memory");
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&s_acc_full[b])) : "memory");
            }
        }
    } else if (warp >= 2 && warp <= 5) {
        // ===================== epilogue: TMEM -> global, junk columns skipped =====================
        const int q = warp & 3;
        int k = 0;
        for (int img = blockIdx.x; img < a.batch; img += gridDim.x, ++k) {
            const unsigned b = (unsigned)k & 1u;
            mbar_wait(smem_u32(&s_acc_full[b]), ((unsigned)k >> 1) & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int mb = 0; mb < MB; ++mb) {
                unsigned r[32];
                const unsigned taddr = tmem + ((unsigned)(q * 32) << 16) + b * ACC_COLS + (unsigned)mb * 32u;
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
                const int m = mb * 128 + q * 32 + lane;
                const int y = m / a.PW, x = m - y * a.PW;
                if (y < a.H && x < a.W) {
                    float4* po = reinterpret_cast<float4*>(out + (((size_t)img * a.H + y) * a.W + x) * 32);
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        po[j] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&s_acc_empty[b])) : "memory");
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TCOLS));
}

}  // namespace

// dgrad of a stride-1 SAME convolution with Cin == 32, Cout a multiple of 32 (at most 2 panels), pictures small enough that the
// padded box and 2 x MB x 32 accumulator columns fit: MB = ceil(H * PW / 128) <= 3
bool nc_tc_conv_strip_supported(int mode, const nck::ConvGeom& g, const float* src, const float* wmat, float* out) {
    if (mode != 1) return false;
    const char* e = getenv("NC_CONV_STRIP");
    if (e && atoi(e) == 0) return false;
    if (g.sh != 1 || g.sw != 1 || g.oh != g.ih || g.ow != g.iw) return false;
    if ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(wmat) | reinterpret_cast<uintptr_t>(out)) & 15) return false;
    if (g.ic != 32 || g.oc % 32 || g.oc / 32 > 2) return false;
    const int PW = g.iw + g.fw - 1, PH = g.ih + g.fh - 1;
    if (PW > 256 || PH > 256) return false;
    const int mb = (g.ih * PW + 127) / 128;
    if (mb < 1 || mb > 3) return false;
    const int rows = mb * 128 + (g.fh - 1) * PW + g.fw - 1;
    const size_t panel_bytes = (size_t)((rows > PW * PH ? rows : PW * PH) * 128 + 1023) / 1024 * 1024;
    return g.fw * 32 <= 256 && (size_t)(g.oc / 32) * panel_bytes + (size_t)WSTAGES * g.fw * W_TILE_BYTES + 1024 <= 220 * 1024;
}

int nc_tc_conv_strip_dgrad(nc_ctx* ctx, const nck::ConvGeom& g, int batch, const float* dy, const float* wmat, float* dx, cudaStream_t st) {
    int rc = get_encode(ctx);
    if (rc != NC_OK) return rc;
    StripArgs a;
    a.W = g.iw; a.H = g.ih; a.fw = g.fw; a.fh = g.fh;
    a.PW = g.iw + g.fw - 1; a.PH = g.ih + g.fh - 1;
    a.low_w = -(g.fw - 1 - g.pl); a.low_h = -(g.fh - 1 - g.pt);          // dgrad: the window of input pixel p starts at p - (f - 1 - pad_before)
    a.panels = g.oc / 32;
    const int mb = (g.ih * a.PW + 127) / 128;
    const int rows = mb * 128 + (g.fh - 1) * a.PW + g.fw - 1;
    a.panel_bytes = ((rows > a.PW * a.PH ? rows : a.PW * a.PH) * 128 + 1023) / 1024 * 1024;
    a.cin_rows = g.ic; a.batch = batch;

    CUtensorMap mx, mw;
    {
        // dy as NHWC {C, W, H, N}; one box = a whole padded picture of one 32-channel panel (out-of-range pixels read as zero)
        cuuint64_t dims[4] = {(cuuint64_t)g.oc, (cuuint64_t)g.iw, (cuuint64_t)g.ih, (cuuint64_t)batch};
        cuuint64_t strides[3] = {(cuuint64_t)g.oc * 4, (cuuint64_t)g.iw * g.oc * 4, (cuuint64_t)g.ih * g.iw * g.oc * 4};
        cuuint32_t box[4] = {32u, (cuuint32_t)a.PW, (cuuint32_t)a.PH, 1u};
        cuuint32_t estr[4] = {1u, 1u, 1u, 1u};
        CUresult r = g_encode(&mx, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(dy), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return nc_set_error(ctx, NC_ECUDA, "cuTensorMapEncodeTiled (padded picture box) failed (%d)", (int)r);
    }
    rc = make_map(ctx, &mw, wmat, (long long)g.fh * g.fw * g.ic, g.oc, 32 * g.fw, false);      // [taps*Cin][Cout], K-major boxes of fw taps x 32 rows
    if (rc != NC_OK) return rc;
    const size_t smem = (size_t)a.panels * a.panel_bytes + (size_t)WSTAGES * g.fw * W_TILE_BYTES + 1024;
    const int grid = batch < ctx->sm_count ? batch : ctx->sm_count;
#define NC_STRIP(MB_)                                                                                                  \
    do {                                                                                                               \
        NC_CUDA(ctx, cudaFuncSetAttribute(conv_strip_dgrad_kernel<MB_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        conv_strip_dgrad_kernel<MB_><<<grid, kStripThreads, smem, st>>>(mx, mw, a, dx);                                 \
    } while (0)
    if (mb == 1) NC_STRIP(1); else if (mb == 2) NC_STRIP(2); else NC_STRIP(3);
#undef NC_STRIP
    NC_LAUNCH_CHECK(ctx);
    return NC_OK;
}

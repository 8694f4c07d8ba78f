// Shared device / host helpers of the tcgen05 kernels (gemm_tcgen05.cu, conv_tcgen05.cu).
#pragma once
#include "nc_internal.h"
#include <cuda.h>
#include <cudaTypedefs.h>

namespace {

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

// one lane of a fully active warp, chosen by the hardware.  Code under `if (elect_one())` is known to ptxas to run in a
// single thread, so tcgen05.mma / commit are emitted back to back; under `if (lane == 0)` every one of them is wrapped in
// an elect-and-branch loop.
__device__ __forceinline__ bool elect_one() {
    unsigned pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(pred));
    return pred != 0;
}

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(unsigned dst, const CUtensorMap* map, int c0, int c1, unsigned bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}

// TMA tensor store of one 128-row x 32-float box (128-byte-swizzled rows in shared memory) to [c1 .. c1+128) x [c0 .. c0+32) of
// the mapped matrix; rows / columns outside the matrix are clipped by the unit
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, unsigned src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(map), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_commit_and_wait_read() {
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");       // shared memory may be reused / the CTA may exit
}

// Epilogue through shared memory + TMA store.  Called by the 8 epilogue warps (256 threads) for each 32-column chunk they
// read from TMEM: v[32] = this thread's row (tile row r) of chunk `chunk`, already biased / activated.  The chunk is staged as a
// [128][32] fp32 box whose 16-byte pieces are XOR-swizzled by (row & 7) -- conflict-free 128-bit shared stores, and the layout
// CU_TENSOR_MAP_SWIZZLE_128B expects.
__device__ __forceinline__ void stage_row_chunk(unsigned stage_base, int chunk, unsigned r, const float* v) {
    const unsigned row = stage_base + (unsigned)chunk * 16384u + r * 128u;
#pragma unroll
    for (unsigned j = 0; j < 8; ++j)
        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(row + ((j ^ (r & 7u)) << 4)), "f"(v[4 * j]), "f"(v[4 * j + 1]),
                     "f"(v[4 * j + 2]), "f"(v[4 * j + 3]) : "memory");
}
// after the last chunk: make the staged tile visible to the TMA unit, meet the other epilogue warps, one thread stores
__device__ __forceinline__ void store_staged_tile(const CUtensorMap* map, unsigned stage_base, int chunks, int n0, int m0, int N, bool leader_warp) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (leader_warp && elect_one()) {
        for (int c = 0; c < chunks; ++c)
            if (n0 + c * 32 < N) tma_store_2d(map, stage_base + (unsigned)c * 16384u, n0 + c * 32, m0);
        tma_store_commit_and_wait_read();
    }
}

// shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start>>4 [0,14), LBO>>4 [16,30),
// SBO>>4 [32,46), version=1 [46,48), layout SWIZZLE_128B = 2 [61,64)
__device__ __forceinline__ unsigned long long umma_desc(unsigned saddr, unsigned lbo_bytes, unsigned sbo_bytes,
                                                        unsigned long long layout_type = 2) {
    unsigned long long d = 0;
    d |= (unsigned long long)((saddr >> 4) & 0x3FFF);
    d |= (unsigned long long)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (unsigned long long)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= 1ull << 46;
    d |= layout_type << 61;
    return d;
}

// X3 = true: "3xTF32" -- every fp32 operand is split into hi = rna_tf32(x) and lo = rna_tf32(x - hi) and
// the product is accumulated as hi.hi + lo.hi + hi.lo (three MMAs per k step), which restores fp32-class
// accuracy (error ~2^-21 per product) on the TF32 tensor pipe.  X3 = false: one rounded TF32 product.
__device__ __forceinline__ void umma_tf32(unsigned tmem, unsigned long long da, unsigned long long db, unsigned idesc,
                                          unsigned accum) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
        ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(accum) : "memory");
}

// the same product with the A operand in tensor memory: [a_tmem] addresses 128 lanes (rows) x 8 columns (k)
__device__ __forceinline__ void umma_tf32_ta(unsigned tmem, unsigned a_tmem, unsigned long long db, unsigned idesc, unsigned accum) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}\n"
        ::"r"(tmem), "r"(a_tmem), "l"(db), "r"(idesc), "r"(accum) : "memory");
}

static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;

int get_encode(nc_ctx* ctx) {
    if (g_encode) return NC_OK;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !fn) {
        cudaGetLastError();
        return nc_set_error(ctx, NC_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    }
    g_encode = (PFN_cuTensorMapEncodeTiled_v12000)fn;
    return NC_OK;
}

// row-major fp32 matrix [rows][cols] (ld = cols), box = 32 columns x box_rows rows, 128-byte swizzle, zero OOB fill
int make_map(nc_ctx* ctx, CUtensorMap* map, const float* base, long long rows, long long cols, int box_rows,
             bool mn_major = false, long long ld = 0) {
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)(ld > 0 ? ld : cols) * sizeof(float)};
    cuuint32_t box[2] = {32u, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1u, 1u};
    CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return nc_set_error(ctx, NC_ECUDA, "cuTensorMapEncodeTiled failed (%d) for %lld x %lld", (int)r, rows, cols);
    return NC_OK;
}


}  // namespace

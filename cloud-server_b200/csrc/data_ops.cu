// Device-side batch assembly for DataSet.next_batch (apps/construction/util/construct_distribute.py:526-569):
// the pictures stay resident in HBM as float32 rows; a batch is a gather of rows by index.
#include "nc_internal.h"

namespace {
// one warp per output row, 16-byte vector copies when the row pitch allows
__global__ void gather_rows_kernel(const float* __restrict__ src, const long long* __restrict__ idx, int n, int width,
                                   long long n_src, float* __restrict__ out, int* __restrict__ bad) {
    const int warps_per_block = blockDim.x >> 5;
    const int lane = threadIdx.x & 31;
    for (int r = blockIdx.x * warps_per_block + (threadIdx.x >> 5); r < n; r += gridDim.x * warps_per_block) {
        const long long s = idx[r];
        float* po = out + (long long)r * width;
        if (s < 0 || s >= n_src) {        // numpy raises IndexError here: the row is zero-filled and the sticky flag set
            if (lane == 0) atomicExch(bad, 1);
            for (int i = lane; i < width; i += 32) po[i] = 0.0f;
            continue;
        }
        const float* ps = src + s * width;
        if ((width & 3) == 0 && ((reinterpret_cast<uintptr_t>(ps) | reinterpret_cast<uintptr_t>(po)) & 15) == 0) {
            const float4* a = reinterpret_cast<const float4*>(ps);
            float4* b = reinterpret_cast<float4*>(po);
            for (int i = lane; i < width / 4; i += 32) b[i] = __ldg(a + i);
        } else {
            for (int i = lane; i < width; i += 32) po[i] = __ldg(ps + i);
        }
    }
}
}  // namespace

extern "C" int nc_gather_rows(nc_ctx* ctx, const float* src, int64_t n_src, int width, const int64_t* idx, int n, float* out,
                              void* stream) {
    if (!ctx) return nc_set_error(nullptr, NC_EINVAL, "ctx is NULL");
    if (n < 0 || width <= 0 || n_src < 0 || (n > 0 && (!src || !idx || !out))) return nc_set_error(ctx, NC_EINVAL, "bad arguments");
    if (n == 0) return NC_OK;
    NC_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    if (!ctx->gather_bad) {           // sticky out-of-range flag, read (and cleared) by nc_gather_rows_status
        NC_CUDA(ctx, cudaMalloc(&ctx->gather_bad, sizeof(int)));
        NC_CUDA(ctx, cudaMemset(ctx->gather_bad, 0, sizeof(int)));
    }
    int* bad = ctx->gather_bad;
    int blocks = (n + 7) / 8;
    if (blocks > ctx->sm_count * 8) blocks = ctx->sm_count * 8;
    gather_rows_kernel<<<blocks, 256, 0, st>>>(src, reinterpret_cast<const long long*>(idx), n, width, n_src, out, bad);
    NC_LAUNCH_CHECK(ctx);
    return NC_OK;
}

// Did any nc_gather_rows on this ctx since the last status call see an index outside [0, n_src)?  The reference's
// `self._images[perm]` (construct_distribute.py:541, :556) raises IndexError for such an index; the gather kernel cannot
// raise, so it zero-fills the row and sets a sticky device flag that this call reads back (synchronises `stream`).
extern "C" int nc_gather_rows_status(nc_ctx* ctx, void* stream) {
    if (!ctx) return nc_set_error(nullptr, NC_EINVAL, "ctx is NULL");
    if (!ctx->gather_bad) return NC_OK;
    NC_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    int host = 0;
    NC_CUDA(ctx, cudaMemcpyAsync(&host, ctx->gather_bad, sizeof(int), cudaMemcpyDeviceToHost, st));
    NC_CUDA(ctx, cudaMemsetAsync(ctx->gather_bad, 0, sizeof(int), st));
    NC_CUDA(ctx, cudaStreamSynchronize(st));
    if (host) return nc_set_error(ctx, NC_EINVAL, "IndexError: nc_gather_rows was given a row index outside [0, n_src)");
    return NC_OK;
}

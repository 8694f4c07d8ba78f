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
        if (s < 0 || s >= n_src) { if (lane == 0) atomicExch(bad, 1); continue; }
        const float* ps = src + s * width;
        float* po = out + (long long)r * width;
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
    int rc = nc_ws_reserve(ctx, 256);
    if (rc != NC_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    int* bad = reinterpret_cast<int*>(ctx->ws);
    NC_CUDA(ctx, cudaMemsetAsync(bad, 0, sizeof(int), st));
    int blocks = (n + 7) / 8;
    if (blocks > ctx->sm_count * 8) blocks = ctx->sm_count * 8;
    gather_rows_kernel<<<blocks, 256, 0, st>>>(src, reinterpret_cast<const long long*>(idx), n, width, n_src, out, bad);
    NC_LAUNCH_CHECK(ctx);
    return NC_OK;
}

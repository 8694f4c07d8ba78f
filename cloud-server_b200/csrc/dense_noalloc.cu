// nc_dense_forward_noalloc: y = x . W + b for one dense layer, with NO allocation of any kind.
//
// Who this is for.  nc_net_forward keeps its split-K partial sums in a workspace that the library
// grows with cudaMalloc on demand.  Some hosts of this library cannot tolerate that call at
// request time:
//   * a serving process that has handed all device memory to its own pool at start-up (a
//     cudaMalloc from a plug-in then fails, or fragments the pool it was promised);
//   * a latency-critical request thread: cudaMalloc takes a device-wide lock and may synchronise
//     the device, stalling every other stream of the process;
//   * code running under a memory budget audited per tenant, where hidden allocations made by a
//     library are forbidden.
// For them this entry point computes the dense (fully connected / logits) layer
// -- connect_layer(), apps/construction/util/construct_distribute.py:168-182, and the output
// layer :262-264 -- using only the buffers the caller passes plus the fixed scratch array below.
//
// The scratch array lives in the library image: every GPU that loads the library carries
// NC_NOALLOC_SCRATCH_BYTES of it from load until the process exits, whether or not this function
// is ever called.  All calls on a device share it.  The kernels write every partial sum before
// they read it, so nothing in it has to start at zero and the previous call's leftovers are never
// observed; but two calls whose kernels overlap on the GPU (different streams, or different host
// threads) would overwrite each other's partial sums -- callers must serialise calls per device.
//
// Set-up: the first call on each device resolves the device address of the array (one
// cudaGetSymbolAddress, no allocation) and remembers it.  Make that first call before starting
// any stream capture on that device.
#include "nc_internal.h"
#include <atomic>

#define NC_NOALLOC_SCRATCH_FLOATS (1u << 20)                       // 4 MiB
#define NC_NOALLOC_SCRATCH_BYTES (NC_NOALLOC_SCRATCH_FLOATS * 4u)

// In host code this name is only a handle for cudaGetSymbolAddress; it is not a device address.
static __device__ __align__(256) float g_noalloc_scratch[NC_NOALLOC_SCRATCH_FLOATS];

namespace {

constexpr int TM = 64, TN = 64, TK = 16;

// split-K tile kernel: partial[z][m][n] = sum over this slice of k of x[m][k] * w[k][n]
__global__ void __launch_bounds__(256)
dense_partial_kernel(const float* __restrict__ x, const float* __restrict__ w, int M, int N, int K, int kslice,
                     float* __restrict__ partial) {
    __shared__ __align__(16) float As[TK][TM + 4];     // read back through float4: rows are (TM + 4) * 4 bytes, a multiple of 16
    __shared__ __align__(16) float Bs[TK][TN + 4];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
    const int kbeg = blockIdx.z * kslice, kend = min(K, kbeg + kslice);
    const int tx = tid % 16, ty = tid / 16;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
    for (int k0 = kbeg; k0 < kend; k0 += TK) {
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            const int e = tid + l * 256;
            const int kk = e % TK, mm = e / TK;
            const int gm = m0 + mm, gk = k0 + kk;
            As[kk][mm] = (gm < M && gk < kend) ? __ldg(x + (long long)gm * K + gk) : 0.0f;
            const int nn = e % TN, kb = e / TN;
            const int gn = n0 + nn, gkb = k0 + kb;
            Bs[kb][nn] = (gn < N && gkb < kend) ? __ldg(w + (long long)gkb * N + gn) : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < TK; ++kk) {
            const float4 a4 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
            const float4 b4 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
            const float a[4] = {a4.x, a4.y, a4.z, a4.w};
            const float b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
    float* plane = partial + (long long)blockIdx.z * M * N;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gm = m0 + ty * 4 + i;
        if (gm >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gn = n0 + tx * 4 + j;
            if (gn < N) plane[(long long)gm * N + gn] = acc[i][j];
        }
    }
}

// y[i] = bias[i % N] + sum_z partial[z][i], summed in slice order (deterministic)
__global__ void dense_reduce_kernel(float* __restrict__ y, const float* __restrict__ partial, int splits, long long count,
                                    const float* __restrict__ bias, int N) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        float s = 0.0f;
        for (int z = 0; z < splits; ++z) s += partial[(long long)z * count + i];
        if (bias) s += __ldg(bias + (int)(i % N));
        y[i] = s;
    }
}

constexpr int kMaxDevices = 64;
// device address of g_noalloc_scratch, per device ordinal.  Host threads of different ctxs may look the same device up at
// the same time: both get the same address from the runtime, so the only requirement is an untorn pointer (atomic).
std::atomic<float*> g_scratch_addr[kMaxDevices];

}  // namespace

extern "C" int nc_dense_forward_noalloc(nc_ctx* ctx, const float* x, const float* w, const float* bias, int m, int k,
                                        int n, float* y, void* stream) {
    if (!ctx) return nc_set_error(nullptr, NC_EINVAL, "ctx is NULL");
    if (!x || !w || !y) return nc_set_error(ctx, NC_EINVAL, "x, w and y must not be NULL");
    if (m <= 0 || k <= 0 || n <= 0) return nc_set_error(ctx, NC_EINVAL, "bad shape m=%d k=%d n=%d", m, k, n);
    if (ctx->device < 0 || ctx->device >= kMaxDevices) return nc_set_error(ctx, NC_EUNSUPPORTED, "device ordinal %d >= %d", ctx->device, kMaxDevices);
    NC_CUDA(ctx, cudaSetDevice(ctx->device));

    // one-time set-up per device (no allocation): where does the scratch array live on THIS device?
    float* scratch = g_scratch_addr[ctx->device].load(std::memory_order_acquire);
    if (!scratch) {
        NC_CUDA(ctx, cudaGetSymbolAddress((void**)&scratch, g_noalloc_scratch));
        g_scratch_addr[ctx->device].store(scratch, std::memory_order_release);
    }

    // same slicing rule as the library's workspace path, then shrink until the partials fit
    const int tm = (m + TM - 1) / TM, tn = (n + TN - 1) / TN;
    const long long tiles = (long long)tm * tn, plane = (long long)m * n;
    if (plane > (long long)NC_NOALLOC_SCRATCH_FLOATS)
        return nc_set_error(ctx, NC_EINVAL,
                            "nc_dense_forward_noalloc: m*n = %lld outputs exceed the fixed scratch of %u floats (%u bytes); "
                            "split the batch or use nc_net_forward", plane, NC_NOALLOC_SCRATCH_FLOATS, NC_NOALLOC_SCRATCH_BYTES);
    int splits = 1;
    if (tiles < 2LL * ctx->sm_count && k >= 512) {
        splits = (int)((2LL * ctx->sm_count + tiles - 1) / tiles);
        const int maxs = k / 128;
        if (splits > maxs) splits = maxs;
        if (splits < 1) splits = 1;
    }
    const long long fit = (long long)NC_NOALLOC_SCRATCH_FLOATS / plane;
    if (splits > fit) splits = (int)fit;
    int kslice = (k + splits - 1) / splits;
    kslice = (kslice + TK - 1) / TK * TK;
    splits = (k + kslice - 1) / kslice;

    cudaStream_t st = (cudaStream_t)stream;
    dim3 grid(tn, tm, splits);
    dense_partial_kernel<<<grid, 256, 0, st>>>(x, w, m, n, k, kslice, scratch);
    NC_LAUNCH_CHECK(ctx);
    long long blocks = (plane + 255) / 256;
    if (blocks > (long long)ctx->sm_count * 16) blocks = (long long)ctx->sm_count * 16;
    dense_reduce_kernel<<<(int)blocks, 256, 0, st>>>(y, scratch, splits, plane, bias, n);
    NC_LAUNCH_CHECK(ctx);
    return NC_OK;
}

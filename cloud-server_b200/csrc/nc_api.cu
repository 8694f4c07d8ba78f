// Context, error reporting and workspace management of libnjucloud_b200 (C ABI, see
// include/njucloud_b200.h).
#include "nc_internal.h"
#include <string.h>

static thread_local std::string g_noctx_err;

int nc_set_error(nc_ctx* ctx, int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf; else g_noctx_err = buf;
    return code;
}

int nc_ws_reserve(nc_ctx* ctx, size_t bytes) {
    if (ctx->ws_bytes >= bytes) return NC_OK;
    if (ctx->ws) { cudaFree(ctx->ws); ctx->ws = nullptr; ctx->ws_bytes = 0; }
    size_t want = (bytes + ((size_t)1 << 20) - 1) & ~(((size_t)1 << 20) - 1);
    cudaError_t e = cudaMalloc(&ctx->ws, want);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return nc_set_error(ctx, NC_ENOMEM, "workspace cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
    }
    ctx->ws_bytes = want;
    return NC_OK;
}

int nc_pinned_reserve(nc_ctx* ctx, size_t bytes) {
    if (ctx->pinned_bytes >= bytes) return NC_OK;
    if (ctx->pinned) { cudaFreeHost(ctx->pinned); ctx->pinned = nullptr; ctx->pinned_bytes = 0; }
    size_t want = (bytes + 4095) & ~(size_t)4095;
    cudaError_t e = cudaHostAlloc(&ctx->pinned, want, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return nc_set_error(ctx, NC_ENOMEM, "cudaHostAlloc(%zu) failed: %s", want, cudaGetErrorString(e));
    }
    ctx->pinned_bytes = want;
    return NC_OK;
}

extern "C" int nc_version(void) { return 100; }

extern "C" int nc_init(int device, nc_ctx** out) {
    if (!out) return nc_set_error(nullptr, NC_EINVAL, "out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        cudaGetLastError();
        return nc_set_error(nullptr, NC_ECUDA, "no CUDA device available (%s): this library has no CPU fallback",
                            e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    }
    if (device < 0 || device >= count) return nc_set_error(nullptr, NC_EINVAL, "device %d out of range [0,%d)", device, count);
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return nc_set_error(nullptr, NC_ECUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) return nc_set_error(nullptr, NC_ECUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
    if (prop.major != 10)
        return nc_set_error(nullptr, NC_EUNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a (B200) only",
                            device, prop.major, prop.minor);
    nc_ctx* c = new (std::nothrow) nc_ctx();
    if (!c) return nc_set_error(nullptr, NC_ENOMEM, "out of host memory");
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    *out = c;
    return NC_OK;
}

extern "C" int nc_destroy(nc_ctx* ctx) {
    if (!ctx) return NC_OK;
    cudaSetDevice(ctx->device);
    for (auto& kv : ctx->pil_tables) cudaFree(kv.second.dptr);
    if (ctx->ws) cudaFree(ctx->ws);
    if (ctx->gather_bad) cudaFree(ctx->gather_bad);
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    delete ctx;
    return NC_OK;
}

extern "C" const char* nc_last_error(nc_ctx* ctx) { return ctx ? ctx->err.c_str() : g_noctx_err.c_str(); }

extern "C" int64_t nc_launch_count(nc_ctx* ctx) { return ctx ? ctx->launches : 0; }

namespace {
__global__ void normalise_kernel(const uint8_t* __restrict__ in, long long count, float* __restrict__ out) {
    const float inv255 = (float)(1.0 / 255.0);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        out[i] = __fmul_rn((float)__ldg(in + i), inv255);
}
}  // namespace

extern "C" int nc_pp_normalise(nc_ctx* ctx, const uint8_t* in, int64_t count, float* out, void* stream) {
    if (!ctx) return nc_set_error(nullptr, NC_EINVAL, "ctx is NULL");
    if (count < 0 || (count > 0 && (!in || !out))) return nc_set_error(ctx, NC_EINVAL, "bad arguments");
    if (count == 0) return NC_OK;
    NC_CUDA(ctx, cudaSetDevice(ctx->device));
    long long blocks = (count + 255) / 256;
    const long long maxb = (long long)ctx->sm_count * 8;
    if (blocks > maxb) blocks = maxb;
    normalise_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(in, count, out);
    NC_LAUNCH_CHECK(ctx);
    return NC_OK;
}

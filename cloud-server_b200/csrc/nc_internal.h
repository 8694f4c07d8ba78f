// Internal declarations shared by the .cu files of libnjucloud_b200 (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include <string>
#include <vector>
#include <map>
#include <utility>

#include "../../include/njucloud_b200.h"

struct nc_dev_table {  // a small coefficient table resident in device memory, cached per shape
    void* dptr = nullptr;
    size_t bytes = 0;
    std::vector<unsigned char> host;  // host copy of the table header
};

struct nc_ctx {
    int device = 0;
    int sm_count = 148;
    int64_t launches = 0;
    std::string err;
    // PIL antialias tables keyed by (in_size, out_size)
    std::map<std::pair<int, int>, nc_dev_table> pil_tables;
    // reusable device workspace (grown on demand, never shrunk)
    void* ws = nullptr;
    size_t ws_bytes = 0;
    // small pinned host staging buffer for host_draws etc.
    void* pinned = nullptr;
    size_t pinned_bytes = 0;
    int* gather_bad = nullptr;      // device flag: nc_gather_rows saw an out-of-range index (nc_gather_rows_status)
};

int nc_set_error(nc_ctx* ctx, int code, const char* fmt, ...);
int nc_ws_reserve(nc_ctx* ctx, size_t bytes);       // ctx->ws valid for >= bytes afterwards
int nc_pinned_reserve(nc_ctx* ctx, size_t bytes);

// ---- NCCL, bound with dlopen at run time (nccl_glue.cpp) ------------------------------------
int nc_nccl_load(nc_ctx* ctx, const char* path);    // idempotent; NULL / "" = "libnccl.so.2" on the loader path
int nc_nccl_version();
int nc_nccl_unique_id_impl(nc_ctx* ctx, unsigned char out[128]);
int nc_nccl_comm_init(nc_ctx* ctx, void** comm, const unsigned char id[128], int rank, int world);
int nc_nccl_comm_dup(nc_ctx* ctx, void* comm, int rank, void** out);
void nc_nccl_comm_destroy(void* comm);
int nc_nccl_allreduce_sum_f32(nc_ctx* ctx, void* comm, float* buf, size_t count, cudaStream_t st);

// ---- gradient all-reduce over NVLink peer memory (p2p_allreduce.cu) ---------------------------
struct nc_p2p;
int nc_p2p_export(nc_ctx* ctx, nc_p2p** pp, float* grads, unsigned char out[128]);
int nc_p2p_attach(nc_ctx* ctx, nc_p2p* p, float* grads, const unsigned char* all_handles, int rank, int world);
bool nc_p2p_ready(const nc_p2p* p);
int nc_p2p_allreduce(nc_ctx* ctx, nc_p2p* p, long long off, long long count, cudaStream_t st);
int nc_p2p_error(nc_ctx* ctx, nc_p2p* p, int* out);
void nc_p2p_destroy(nc_p2p* p);

#define NC_CUDA(ctx, call)                                                                     \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess)                                                                \
            return nc_set_error((ctx), NC_ECUDA, "%s failed: %s (%s:%d)", #call,               \
                                cudaGetErrorString(e__), __FILE__, __LINE__);                  \
    } while (0)

#define NC_LAUNCH_CHECK(ctx)                                                                   \
    do {                                                                                       \
        (ctx)->launches++;                                                                     \
        cudaError_t e__ = cudaGetLastError();                                                  \
        if (e__ != cudaSuccess)                                                                \
            return nc_set_error((ctx), NC_ECUDA, "kernel launch failed: %s (%s:%d)",           \
                                cudaGetErrorString(e__), __FILE__, __LINE__);                  \
    } while (0)

// ---- host-side coefficient tables (pp_tables.cpp; compiled with -ffp-contract=off) ----------
struct CvCubicAxis {      // OpenCV INTER_CUBIC, one axis, dst == 28
    int ofs[28][4];       // clamped source index of the 4 taps
    short coef[28][4];    // round(c * 2048)
};
void nc_cv_cubic_axis(int src, int dst, CvCubicAxis* out);
struct IppCubicAxis {     // the same resize in the IPP arithmetic of the wheel's default: float32 coefficients
    int ofs[28][4];
    float coef[28][4];
};
void nc_ipp_cubic_axis(int src, int dst, IppCubicAxis* out);

struct PilAxisHost {      // Pillow BICUBIC antialias tables for one axis
    int out_size = 0, ksize = 0;
    std::vector<int> xmin, cnt;   // [out]
    std::vector<int> kk;          // [out][ksize], 22-bit fixed point
};
void nc_pil_bicubic_axis(int in_size, int out_size, PilAxisHost* out);
std::vector<int> nc_gaussian_kernel_q8(int k);

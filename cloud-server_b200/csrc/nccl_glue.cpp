// NCCL bound at run time (dlopen), so the library itself links only the CUDA runtime and a host
// without NCCL can still load it.  Only what data-parallel training needs: a communicator per net
// and an in-place fp32 sum all-reduce on the caller's stream.
//
// The reference's "distribution" is a 1 ps + 1 worker gRPC parameter server
// (construct_distribute.py:344-357); the synchronous all-reduce below is its B200 replacement
// (multi-GPU section).
#include "nc_internal.h"
#include <dlfcn.h>
#include <mutex>
#include <string>

namespace {
struct UniqueId { char internal[128]; };                       // ncclUniqueId (NCCL_UNIQUE_ID_BYTES = 128)
using fn_get_unique_id = int (*)(UniqueId*);
using fn_comm_init_rank = int (*)(void**, int, UniqueId, int);
using fn_comm_destroy = int (*)(void*);
using fn_all_reduce = int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t);
using fn_error_string = const char* (*)(int);
using fn_get_version = int (*)(int*);
using fn_comm_split = int (*)(void*, int, int, void**, void*);

struct Api {
    void* handle = nullptr;
    fn_get_unique_id get_unique_id = nullptr;
    fn_comm_init_rank comm_init_rank = nullptr;
    fn_comm_destroy comm_destroy = nullptr;
    fn_all_reduce all_reduce = nullptr;
    fn_error_string error_string = nullptr;
    fn_get_version get_version = nullptr;
    fn_comm_split comm_split = nullptr;
    std::string path;
};
Api g_api;
std::mutex g_mu;

constexpr int kNcclFloat32 = 7;   // ncclDataType_t: ncclFloat32
constexpr int kNcclSum = 0;       // ncclRedOp_t: ncclSum

const char* err_text(int rc) { return g_api.error_string ? g_api.error_string(rc) : "unknown NCCL error"; }
}  // namespace

int nc_nccl_load(nc_ctx* ctx, const char* path) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_api.handle) return NC_OK;
    const char* p = (path && *path) ? path : "libnccl.so.2";
    void* h = dlopen(p, RTLD_NOW | RTLD_GLOBAL);
    if (!h) return nc_set_error(ctx, NC_EUNSUPPORTED, "cannot load NCCL from '%s': %s", p, dlerror());
    Api a;
    a.handle = h; a.path = p;
    a.get_unique_id = (fn_get_unique_id)dlsym(h, "ncclGetUniqueId");
    a.comm_init_rank = (fn_comm_init_rank)dlsym(h, "ncclCommInitRank");
    a.comm_destroy = (fn_comm_destroy)dlsym(h, "ncclCommDestroy");
    a.all_reduce = (fn_all_reduce)dlsym(h, "ncclAllReduce");
    a.error_string = (fn_error_string)dlsym(h, "ncclGetErrorString");
    a.get_version = (fn_get_version)dlsym(h, "ncclGetVersion");
    a.comm_split = (fn_comm_split)dlsym(h, "ncclCommSplit");      // NCCL >= 2.18; optional
    if (!a.get_unique_id || !a.comm_init_rank || !a.comm_destroy || !a.all_reduce)
        return nc_set_error(ctx, NC_EUNSUPPORTED, "'%s' does not export the NCCL entry points", p);
    g_api = a;
    return NC_OK;
}

int nc_nccl_version() {
    int v = 0;
    if (g_api.get_version) g_api.get_version(&v);
    return v;
}

int nc_nccl_unique_id_impl(nc_ctx* ctx, unsigned char out[128]) {
    UniqueId id;
    const int rc = g_api.get_unique_id(&id);
    if (rc != 0) return nc_set_error(ctx, NC_ECUDA, "ncclGetUniqueId failed: %s", err_text(rc));
    memcpy(out, id.internal, 128);
    return NC_OK;
}

int nc_nccl_comm_init(nc_ctx* ctx, void** comm, const unsigned char id_bytes[128], int rank, int world) {
    UniqueId id;
    memcpy(id.internal, id_bytes, 128);
    const int rc = g_api.comm_init_rank(comm, world, id, rank);
    if (rc != 0) { *comm = nullptr; return nc_set_error(ctx, NC_ECUDA, "ncclCommInitRank(rank %d of %d) failed: %s", rank, world, err_text(rc)); }
    return NC_OK;
}

// A second communicator over the same ranks (ncclCommSplit with one colour), so that collectives issued from two streams
// never share a communicator.  Collective over all ranks of `comm`.  *out stays NULL when this NCCL has no ncclCommSplit.
int nc_nccl_comm_dup(nc_ctx* ctx, void* comm, int rank, void** out) {
    *out = nullptr;
    if (!g_api.comm_split) return NC_OK;
    const int rc = g_api.comm_split(comm, 0, rank, out, nullptr);
    if (rc != 0) { *out = nullptr; return nc_set_error(ctx, NC_ECUDA, "ncclCommSplit failed: %s", err_text(rc)); }
    return NC_OK;
}

void nc_nccl_comm_destroy(void* comm) {
    if (comm && g_api.comm_destroy) g_api.comm_destroy(comm);
}

int nc_nccl_allreduce_sum_f32(nc_ctx* ctx, void* comm, float* buf, size_t count, cudaStream_t st) {
    const int rc = g_api.all_reduce(buf, buf, count, kNcclFloat32, kNcclSum, comm, st);
    if (rc != 0) return nc_set_error(ctx, NC_ECUDA, "ncclAllReduce of %zu floats failed: %s", count, err_text(rc));
    return NC_OK;
}

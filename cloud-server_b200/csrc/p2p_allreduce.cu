// Gradient all-reduce over NVLink peer memory (one process per GPU, one node, <= 8 ranks).
//
// What it replaces: the sum of the flat gradient buffer over the data-parallel ranks -- the B200 stand-in for the reference's
// ps <-> worker gRPC exchange (apps/construction/util/construct_distribute.py:344-357) -- which nc_net_train_step otherwise
// hands to ncclAllReduce.  At the sizes of this path NCCL is latency-bound: measured on 8 B200 (scripts/nccl_probe.py) 30 us
// for the 0.2 MB convolution bucket and 76-90 us for the 12.9 MB dense bucket, against a step of 280 us.
//
// One kernel per bucket, two-shot, plain loads / stores on peer pointers (every rank maps the other ranks' gradient buffers and
// flag words with CUDA IPC; NVSwitch gives every pair full bandwidth):
//   phase 0   "my gradients are final": a release-store of this call's sequence number into every peer's flag word; wait for
//             all peers' (acquire loads on local memory)
//   phase 1   rank r owns chunk r of the bucket: it reads that chunk from all W ranks, adds them in RANK ORDER (the sum is
//             bit-identical on every rank and from run to run) and stores the result into all W gradient buffers
//   phase 2   the last CTA to finish announces "chunk r is written" to every peer and waits for all peers' announcements, so
//             the kernel ends only when this rank's whole bucket is summed and no peer still reads this rank's buffer
// Sequence numbers live in device memory and are advanced by the kernel itself, so a CUDA graph that contains the launch can
// be replayed.  Waits give up after ~20 s (a peer that never launches) and raise a device-side error flag instead of hanging.
// TMA is not involved: these are ordinary ld / st to peer addresses.
#include "nc_internal.h"

namespace {

constexpr int kMaxRanks = 8;
constexpr int kThreads = 512;

struct PeerPtrs {
    float* g[kMaxRanks];        // gradient buffers (own rank: the local pointer)
    unsigned* f[kMaxRanks];     // flag words: f[p][phase * 8 + src] written by rank src
    int world, rank;
};

__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float4 ld_peer4(const float* p) {       // peer data is never served from a stale L1 line
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_peer1(const float* p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

// true = every flag reached `want`; false = gave up (sets *err)
__device__ __forceinline__ bool wait_flags(const unsigned* flags, int n, unsigned want, int* err) {
    bool ok = true;
    if ((int)threadIdx.x < n) {
        const long long t0 = clock64();
        while ((int)(ld_acquire_sys(flags + threadIdx.x) - want) < 0) {
            if (clock64() - t0 > 40000000000LL) { *err = 1; ok = false; break; }     // ~20 s: far beyond any start-up skew between ranks
            __nanosleep(64);
        }
    }
    return __syncthreads_and(ok ? 1 : 0) != 0;
}

template <int W>
__global__ void __launch_bounds__(kThreads)
p2p_allreduce_kernel(PeerPtrs a, long long off, long long count, unsigned* seq, unsigned* ticket, int* err) {
    __shared__ unsigned s_seq;
    __shared__ int s_last;
    if (threadIdx.x == 0) s_seq = *seq + 1u;          // *seq is advanced by the last CTA of the previous launch (stream order)
    __syncthreads();
    const unsigned s = s_seq;
    const int rank = a.rank;
    // ---- phase 0
    if (blockIdx.x == 0 && (int)threadIdx.x < W) st_release_sys(a.f[threadIdx.x] + rank, s);
    if (!wait_flags(a.f[rank], W, s, err)) return;
    // ---- phase 1: my chunk, summed in rank order, written everywhere
    const long long chunk = ((count + (long long)W * 4 - 1) / ((long long)W * 4)) * 4;
    const long long lo = off + (long long)rank * chunk;
    long long hi = lo + chunk;
    if (hi > off + count) hi = off + count;
    // U float4 per rank in flight per thread: with few ranks one 16-byte load per peer leaves the NVLink round trip exposed
    constexpr int U = W <= 2 ? 4 : (W <= 4 ? 2 : 1);
    const long long stride = (long long)gridDim.x * kThreads * 4;
    for (long long i0 = lo + ((long long)blockIdx.x * kThreads + threadIdx.x) * 4; i0 < hi; i0 += stride * U) {
        float4 v[U][W];
        bool full[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = i0 + u * stride;
            full[u] = i + 4 <= hi;
            if (full[u]) {
#pragma unroll
                for (int p = 0; p < W; ++p) v[u][p] = ld_peer4(a.g[p] + i);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = i0 + u * stride;
            if (full[u]) {
                float4 r = v[u][0];
#pragma unroll
                for (int p = 1; p < W; ++p) { r.x += v[u][p].x; r.y += v[u][p].y; r.z += v[u][p].z; r.w += v[u][p].w; }
#pragma unroll
                for (int p = 0; p < W; ++p) *reinterpret_cast<float4*>(a.g[p] + i) = r;
            } else if (i < hi) {
                for (long long j = i; j < hi; ++j) {
                    float r = ld_peer1(a.g[0] + j);
#pragma unroll
                    for (int p = 1; p < W; ++p) r += ld_peer1(a.g[p] + j);
#pragma unroll
                    for (int p = 0; p < W; ++p) a.g[p][j] = r;
                }
            }
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(ticket, 1u);
        s_last = (t == gridDim.x - 1);
        if (s_last) *ticket = 0u;
    }
    __syncthreads();
    if (!s_last) return;
    // ---- phase 2 (last CTA of this rank): every store of every CTA above is fenced
    __threadfence_system();
    if ((int)threadIdx.x < W) st_release_sys(a.f[threadIdx.x] + 8 + rank, s);
    wait_flags(a.f[rank] + 8, W, s, err);
    if (threadIdx.x == 0) *seq = s;
}

}  // namespace

struct nc_p2p {
    PeerPtrs ptrs;
    void* mapped[2 * kMaxRanks];      // IPC mappings to close
    int nmapped = 0;
    unsigned* flags = nullptr;        // local: [16] flag words, [16] = seq, [17] = ticket, [18] = err
    bool ready = false;
};

// local flag block: allocate once per net; returns the IPC handles of (gradient buffer, flag block): 2 x 64 bytes
int nc_p2p_export(nc_ctx* ctx, nc_p2p** pp, float* grads, unsigned char out[128]) {
    if (!*pp) *pp = new (std::nothrow) nc_p2p();
    if (!*pp) return nc_set_error(ctx, NC_ENOMEM, "out of host memory");
    nc_p2p* p = *pp;
    if (!p->flags) {
        NC_CUDA(ctx, cudaMalloc(&p->flags, 64 * sizeof(unsigned)));
        NC_CUDA(ctx, cudaMemset(p->flags, 0, 64 * sizeof(unsigned)));
    }
    cudaIpcMemHandle_t hg, hf;
    NC_CUDA(ctx, cudaIpcGetMemHandle(&hg, grads));
    NC_CUDA(ctx, cudaIpcGetMemHandle(&hf, p->flags));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
    memcpy(out, &hg, 64);
    memcpy(out + 64, &hf, 64);
    return NC_OK;
}

// all_handles: world x 128 bytes, rank order (what every rank got from nc_p2p_export, gathered by the host)
int nc_p2p_attach(nc_ctx* ctx, nc_p2p* p, float* grads, const unsigned char* all_handles, int rank, int world) {
    if (!p || !p->flags) return nc_set_error(ctx, NC_EINVAL, "nc_net_p2p_handles must be called first");
    if (world < 2 || world > kMaxRanks || rank < 0 || rank >= world) return nc_set_error(ctx, NC_EINVAL, "peer all-reduce takes 2..8 ranks");
    if (p->ready || p->nmapped) return nc_set_error(ctx, NC_EINVAL, "the peers are already mapped (detach the communicator first)");
    p->ptrs.world = world; p->ptrs.rank = rank;
    for (int r = 0; r < world; ++r) {
        if (r == rank) { p->ptrs.g[r] = grads; p->ptrs.f[r] = p->flags; continue; }
        cudaIpcMemHandle_t hg, hf;
        memcpy(&hg, all_handles + (size_t)r * 128, 64);
        memcpy(&hf, all_handles + (size_t)r * 128 + 64, 64);
        void* g = nullptr; void* f = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&g, hg, cudaIpcMemLazyEnablePeerAccess);
        if (e == cudaSuccess) e = cudaIpcOpenMemHandle(&f, hf, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            if (g) cudaIpcCloseMemHandle(g);
            for (int i = 0; i < p->nmapped; ++i) cudaIpcCloseMemHandle(p->mapped[i]);      // leave nothing half-mapped behind
            p->nmapped = 0;
            return nc_set_error(ctx, NC_EUNSUPPORTED, "cudaIpcOpenMemHandle(rank %d): %s (no peer access between these GPUs?)", r, cudaGetErrorString(e));
        }
        p->mapped[p->nmapped++] = g; p->mapped[p->nmapped++] = f;
        p->ptrs.g[r] = static_cast<float*>(g); p->ptrs.f[r] = static_cast<unsigned*>(f);
    }
    p->ready = true;
    return NC_OK;
}

bool nc_p2p_ready(const nc_p2p* p) { return p && p->ready; }

// sum grads[off, off + count) over the ranks, in place on every rank; off and the buffer 16-byte aligned
int nc_p2p_allreduce(nc_ctx* ctx, nc_p2p* p, long long off, long long count, cudaStream_t st) {
    if (!nc_p2p_ready(p)) return nc_set_error(ctx, NC_EINVAL, "peer all-reduce is not attached");
    if (count <= 0) return NC_OK;
    if (off & 3) return nc_set_error(ctx, NC_EINVAL, "peer all-reduce: offset must be a multiple of 4 floats");
    // enough CTAs to keep ~W peer loads x 16 B per thread in flight over the chunk, few enough to leave the SMs to the backward pass
    const long long chunk4 = (count / p->ptrs.world + 3) / 4;
    long long blocks = (chunk4 + kThreads - 1) / kThreads;
    static const int max_blocks = [] { const char* e = getenv("NC_P2P_CTAS"); const int v = e ? atoi(e) : 0; return v > 0 ? v : 48; }();
    if (blocks > max_blocks) blocks = max_blocks;
    if (blocks < 1) blocks = 1;
    unsigned* seq = p->flags + 16; unsigned* ticket = p->flags + 17; int* err = reinterpret_cast<int*>(p->flags + 18);
#define NC_P2P(W_) p2p_allreduce_kernel<W_><<<(int)blocks, kThreads, 0, st>>>(p->ptrs, off, count, seq, ticket, err)
    switch (p->ptrs.world) {
        case 2: NC_P2P(2); break;
        case 3: NC_P2P(3); break;
        case 4: NC_P2P(4); break;
        case 5: NC_P2P(5); break;
        case 6: NC_P2P(6); break;
        case 7: NC_P2P(7); break;
        default: NC_P2P(8); break;
    }
#undef NC_P2P
    NC_LAUNCH_CHECK(ctx);
    return NC_OK;
}

// 0 = fine, 1 = a wait gave up (a peer did not arrive within ~20 s); synchronises nothing: read after a stream synchronise
int nc_p2p_error(nc_ctx* ctx, nc_p2p* p, int* out) {
    *out = 0;
    if (!p || !p->flags) return NC_OK;
    NC_CUDA(ctx, cudaMemcpy(out, p->flags + 18, sizeof(int), cudaMemcpyDeviceToHost));
    return NC_OK;
}

void nc_p2p_destroy(nc_p2p* p) {
    if (!p) return;
    for (int i = 0; i < p->nmapped; ++i) cudaIpcCloseMemHandle(p->mapped[i]);
    cudaFree(p->flags);
    delete p;
}

// Fused "glue" of the train step: kernels that replace chains of small launches around the tensor-core GEMMs.
//
//   head_fwd_bwd_kernel   the implicit output layer of cnn() (tf.matmul(last, W) + b, construct_distribute.py:262-264),
//                         loss_function (:285-298), its gradient, the accuracy count (:381-382), the input gradient of the
//                         output layer, the relu mask of the activation in front of it and the bias gradient of the dense
//                         layer below -- one launch instead of skinny_fwd + loss + dgrad GEMM + act_bwd + colsum + reduce
//   head_wgrad_kernel     weight and bias gradient of the output layer, slices summed by the last CTA to arrive
//   maxpool_bwd_tiled_vec4_cs_kernel / act_bwd_vec4_cs_kernel
//                         the elementwise backward kernels that PRODUCE the output gradient dY of a convolution / dense
//                         layer also sum it per channel: the bias gradient costs no pass of its own
//   split_params_kernel   hi = rn_tf32(w), lo = w - hi for the tensor-core operands (see net.cu, "math modes")
// Every cross-CTA sum is taken in a fixed order by the last CTA to arrive (ticket counter): deterministic, no atomics on
// floating-point data.
#pragma once
#include "net_kernels.cuh"

namespace nck {

// round-to-nearest (ties away) onto the TF32 grid: 10 mantissa bits, low 13 bits zero
__device__ __forceinline__ float rn_tf32(float v) {
    unsigned u;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
    return __uint_as_float(u);
}

// True in every thread of the CTA that takes the last of `expected` tickets.  Callers write their partial results to
// global memory BEFORE the call; the last CTA may then read everybody's (with __ldcg: the data sits in L2, not in this
// SM's L1).  The counter is left at zero for the next launch.
__device__ __forceinline__ bool last_block_arrives(unsigned* counter, unsigned expected) {
    __shared__ int s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(counter, 1u);
        s_last = (t == expected - 1u);
        if (s_last) *counter = 0u;
    }
    __syncthreads();
    const bool last = s_last != 0;
    if (last) __threadfence();
    return last;
}

// two reduce_partials_warp_kernel launches in one: outputs [0, c1) sum ws1[split][c1], outputs [c1, c1 + c2) sum ws2[split][c2]
// (the weight and the bias gradient partials of the first-layer kernels)
__global__ void reduce_partials_warp2_kernel(float* __restrict__ dst1, const float* __restrict__ ws1, long long c1,
                                             float* __restrict__ dst2, const float* __restrict__ ws2, long long c2, int splits) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long i = warp; i < c1 + c2; i += nwarps) {
        const bool first = i < c1;
        const float* ws = first ? ws1 : ws2;
        const long long cnt = first ? c1 : c2, j = first ? i : i - c1;
        float s = 0.0f;
        for (int k = lane; k < splits; k += 32) s += ws[(long long)k * cnt + j];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) (first ? dst1 : dst2)[j] = s;
    }
}

__global__ void split_params_kernel(const float* __restrict__ w, float* __restrict__ hi, float* __restrict__ lo, long long count) {
    const long long stride = (long long)gridDim.x * blockDim.x, tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n4 = count >> 2;
    const float4* w4 = reinterpret_cast<const float4*>(w);
    for (long long i = tid; i < n4; i += stride) {
        const float4 v = __ldg(w4 + i);
        const float4 h = make_float4(rn_tf32(v.x), rn_tf32(v.y), rn_tf32(v.z), rn_tf32(v.w));
        reinterpret_cast<float4*>(hi)[i] = h;
        if (lo) reinterpret_cast<float4*>(lo)[i] = make_float4(v.x - h.x, v.y - h.y, v.z - h.z, v.w - h.w);
    }
    for (long long i = (n4 << 2) + tid; i < count; i += stride) {
        const float v = __ldg(w + i), h = rn_tf32(v);
        hi[i] = h;
        if (lo) lo[i] = v - h;
    }
}

// ---------------------------------------------------------------------------------------------
// The head of the backward pass.  One warp per batch row:
//   z = x . W + b (W[K][N] in shared memory, lanes stride K, shuffle tree)          -> logits[m][:]
//   loss_m, dz = d loss / d z (softmax cross-entropy or mse, batch mean, * grad_scale) -> dz[m][:]
//   dx[m][k] = sum_n dz[n] W[k][n], masked by x[m][k] > 0 when the layer in front is a relu whose output x is
//              (y > 0 <=> pre-activation > 0), rounded to TF32 when a tensor-core GEMM consumes it
//   CS: per-column sums of dx over the rows of this CTA (registers: K / 32 <= 32 per lane) -> the bias gradient of the
//       dense layer below, finished by the last CTA.
// ---------------------------------------------------------------------------------------------
template <int NMAX, bool CS>
__global__ void __launch_bounds__(256)
head_fwd_bwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                    const float* __restrict__ y, float* __restrict__ logits, float* __restrict__ dz, float* __restrict__ dx,
                    int M, int N, int K, int loss_name, float grad_scale, int relu_mask, int round_dx,
                    float* __restrict__ cs_partial, float* __restrict__ db_out, float* __restrict__ loss_partial,
                    int* __restrict__ hit_partial, unsigned* counter, float* __restrict__ loss_out, int* __restrict__ correct_out) {
    extern __shared__ __align__(16) float sw[];      // W[K][N]; afterwards the column sums of the 8 warps [8][K]
    {
        const int total = K * N, total4 = ((reinterpret_cast<uintptr_t>(w) & 15) == 0) ? total / 4 : 0;
        const float4* w4 = reinterpret_cast<const float4*>(w);
        float4* s4 = reinterpret_cast<float4*>(sw);
        int i = threadIdx.x;
        for (; i + 7 * 256 < total4; i += 8 * 256) {
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = __ldg(w4 + i + u * 256);
#pragma unroll
            for (int u = 0; u < 8; ++u) s4[i + u * 256] = v[u];
        }
        for (; i < total4; i += 256) s4[i] = __ldg(w4 + i);
        for (int j = total4 * 4 + threadIdx.x; j < total; j += 256) sw[j] = __ldg(w + j);
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float csum[CS ? 32 : 1];
#pragma unroll
    for (int j = 0; j < (CS ? 32 : 1); ++j) csum[j] = 0.0f;
    float lsum = 0.0f; int hits = 0;
    for (int m = blockIdx.x * 8 + warp; m < M; m += gridDim.x * 8) {
        float z[NMAX];
#pragma unroll
        for (int n = 0; n < NMAX; ++n) z[n] = 0.0f;
        const float* xr = x + (long long)m * K;
        int k = lane;
        for (; k + 224 < K; k += 256) {               // eight independent global loads in flight per lane
            float xv[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) xv[u] = __ldg(xr + k + 32 * u);
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const float* wr = sw + (k + 32 * u) * N;
#pragma unroll
                for (int n = 0; n < NMAX; ++n)
                    if (n < N) z[n] = fmaf(xv[u], wr[n], z[n]);
            }
        }
        for (; k < K; k += 32) {
            const float xv = __ldg(xr + k);
            const float* wr = sw + k * N;
#pragma unroll
            for (int n = 0; n < NMAX; ++n)
                if (n < N) z[n] = fmaf(xv, wr[n], z[n]);
        }
#pragma unroll
        for (int n = 0; n < NMAX; ++n) {
            if (n < N) {
                float v = z[n];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                z[n] = v + (bias ? __ldg(bias + n) : 0.0f);          // every lane holds the full logits row
            }
        }
        const float yl = lane < N ? __ldg(y + (long long)m * N + lane) : 0.0f;
        float yv[NMAX];
#pragma unroll
        for (int n = 0; n < NMAX; ++n) yv[n] = __shfl_sync(0xffffffffu, yl, n);
        float zmax = -INFINITY, ymax = -INFINITY; int za = 0, ya = 0;
#pragma unroll
        for (int n = 0; n < NMAX; ++n) {
            if (n < N) {
                if (z[n] > zmax) { zmax = z[n]; za = n; }
                if (yv[n] > ymax) { ymax = yv[n]; ya = n; }
            }
        }
        float d[NMAX];
        float l = 0.0f;
        if (loss_name == 0) {
            float se = 0.0f;
#pragma unroll
            for (int n = 0; n < NMAX; ++n) if (n < N) se += expf(z[n] - zmax);
            const float lse = zmax + logf(se);
#pragma unroll
            for (int n = 0; n < NMAX; ++n) {
                d[n] = 0.0f;
                if (n < N) {
                    const float logp = z[n] - lse;
                    l -= yv[n] * logp;
                    d[n] = (expf(logp) - yv[n]) * grad_scale / (float)M;
                }
            }
        } else {
#pragma unroll
            for (int n = 0; n < NMAX; ++n) {
                d[n] = 0.0f;
                if (n < N) {
                    const float e = z[n] - yv[n];
                    l += e * e;
                    d[n] = 2.0f * e * grad_scale / ((float)M * (float)N);
                }
            }
            l /= (float)N;
        }
        if (lane == 0) { lsum += l; hits += (za == ya); }
        if (lane < N) {                                // one coalesced row store each (lane n holds column n)
            float zs = 0.0f, ds = 0.0f;
#pragma unroll
            for (int n = 0; n < NMAX; ++n) if (n == lane) { zs = z[n]; ds = d[n]; }
            logits[(long long)m * N + lane] = zs;
            if (dz) dz[(long long)m * N + lane] = ds;
        }
        if (dx) {
            float* dr = dx + (long long)m * K;
            if (CS) {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int kk = lane + 32 * j;
                    if (kk < K) {
                        const float* wr = sw + kk * N;
                        float s = 0.0f;
#pragma unroll
                        for (int n = 0; n < NMAX; ++n) if (n < N) s = fmaf(d[n], wr[n], s);
                        if (relu_mask && !(__ldg(xr + kk) > 0.0f)) s = 0.0f;
                        if (round_dx) s = rn_tf32(s);
                        dr[kk] = s;
                        csum[j] += s;
                    }
                }
            } else {
                for (int kk = lane; kk < K; kk += 32) {
                    const float* wr = sw + kk * N;
                    float s = 0.0f;
#pragma unroll
                    for (int n = 0; n < NMAX; ++n) if (n < N) s = fmaf(d[n], wr[n], s);
                    if (relu_mask && !(__ldg(xr + kk) > 0.0f)) s = 0.0f;
                    if (round_dx) s = rn_tf32(s);
                    dr[kk] = s;
                }
            }
        }
    }
    __shared__ float s_l[8];
    __shared__ int s_h[8];
    if (lane == 0) { s_l[warp] = lsum; s_h[warp] = hits; }
    if (CS) {
        __syncthreads();                                // every warp is done reading W
#pragma unroll
        for (int j = 0; j < 32; ++j) { const int kk = lane + 32 * j; if (kk < K) sw[warp * K + kk] = csum[j]; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f; int hh = 0;
        for (int i = 0; i < 8; ++i) { t += s_l[i]; hh += s_h[i]; }
        loss_partial[blockIdx.x] = t; hit_partial[blockIdx.x] = hh;
    }
    if (CS) {
        for (int kk = threadIdx.x; kk < K; kk += 256) {
            float t = 0.0f;
#pragma unroll
            for (int i = 0; i < 8; ++i) t += sw[i * K + kk];
            cs_partial[(long long)blockIdx.x * K + kk] = t;
        }
    }
    if (!last_block_arrives(counter, gridDim.x)) return;
    const int G = gridDim.x;
    if (warp == 0) {
        float t = 0.0f; int hh = 0;
        for (int b = lane; b < G; b += 32) { t += __ldcg(loss_partial + b); hh += __ldcg(hit_partial + b); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { t += __shfl_xor_sync(0xffffffffu, t, o); hh += __shfl_xor_sync(0xffffffffu, hh, o); }
        if (lane == 0) {
            if (loss_out) *loss_out = t / (float)M;
            if (correct_out) *correct_out = hh;
        }
    }
    if (CS && db_out) {
        for (int kk = threadIdx.x; kk < K; kk += 256) {
            float t = 0.0f;
            int b = 0;
            for (; b + 7 < G; b += 8) {                 // eight independent L2 loads in flight, summed in block order
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = __ldcg(cs_partial + (long long)(b + u) * K + kk);
#pragma unroll
                for (int u = 0; u < 8; ++u) t += v[u];
            }
            for (; b < G; ++b) t += __ldcg(cs_partial + (long long)b * K + kk);
            db_out[kk] = t;
        }
    }
}

// dW[K][N] = x^T . dz over the batch, db[N] = column sums of dz.  grid = (ceil(K / 256), slices): a thread owns one k for
// the rows of its slice (dz rows broadcast from shared memory); the last slice to arrive for a k block sums the slices in
// slice order; the last arriver of k block 0 also sums dz into db.
template <int NMAX>
__global__ void __launch_bounds__(256)
head_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dz, float* __restrict__ partial, float* __restrict__ dW,
                  float* __restrict__ db, int M, int N, int K, int rows_per_slice, unsigned* counters) {
    extern __shared__ float sdy[];                   // dz[rows of this slice][N]
    const int m0 = blockIdx.y * rows_per_slice, m1 = min(M, m0 + rows_per_slice);
    for (int i = threadIdx.x; i < (m1 - m0) * N; i += blockDim.x) sdy[i] = __ldg(dz + (long long)m0 * N + i);
    __syncthreads();
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k < K) {
        float acc[NMAX];
#pragma unroll
        for (int n = 0; n < NMAX; ++n) acc[n] = 0.0f;
        int m = m0;
        for (; m + 3 < m1; m += 4) {                 // four independent loads in flight
            float xv[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) xv[u] = __ldg(x + (long long)(m + u) * K + k);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float* dr = sdy + (m + u - m0) * N;
#pragma unroll
                for (int n = 0; n < NMAX; ++n)
                    if (n < N) acc[n] = fmaf(xv[u], dr[n], acc[n]);
            }
        }
        for (; m < m1; ++m) {
            const float xv = __ldg(x + (long long)m * K + k);
            const float* dr = sdy + (m - m0) * N;
#pragma unroll
            for (int n = 0; n < NMAX; ++n)
                if (n < N) acc[n] = fmaf(xv, dr[n], acc[n]);
        }
        float* out = partial + ((long long)blockIdx.y * K + k) * N;
#pragma unroll
        for (int n = 0; n < NMAX; ++n)
            if (n < N) out[n] = acc[n];
    }
    if (!last_block_arrives(counters + blockIdx.x, gridDim.y)) return;
    const int S = gridDim.y;
    if (k < K) {
        float acc[NMAX];
#pragma unroll
        for (int n = 0; n < NMAX; ++n) acc[n] = 0.0f;
        for (int s = 0; s < S; ++s) {
            const float* p = partial + ((long long)s * K + k) * N;
#pragma unroll
            for (int n = 0; n < NMAX; ++n)
                if (n < N) acc[n] += __ldcg(p + n);
        }
#pragma unroll
        for (int n = 0; n < NMAX; ++n)
            if (n < N) dW[(long long)k * N + n] = acc[n];
    }
    if (blockIdx.x == 0 && db) {
        __shared__ float sdb[16][NMAX + 1];
        const int n = threadIdx.x % NMAX, r0 = threadIdx.x / NMAX;          // 256 / NMAX row lanes (16 or 8)
        float t = 0.0f;
        if (n < N)
            for (int m = r0; m < M; m += 256 / NMAX) t += __ldg(dz + (long long)m * N + n);
        if (r0 < 16) sdb[r0][n] = t;
        __syncthreads();
        if (threadIdx.x < N) {
            float u = 0.0f;
            for (int r = 0; r < 256 / NMAX; ++r) u += sdb[r][threadIdx.x];
            db[threadIdx.x] = u;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// The same head for K <= 1024, laid out for latency: the batch is small (256 rows per GPU under data parallel training)
// and the whole problem is a few MB, so what counts is the length of the dependent chain, not bytes.
//   * a thread owns k = t, t + 256, t + 512, t + 768 for EVERY row its CTA handles: its 4 x N weights live in registers
//     (staged once through shared memory with coalesced 16-byte loads), x loads and dx stores are coalesced, and the column
//     sums of dx (the bias gradient of the dense layer below) are four private accumulators -- no cross-warp reduction;
//   * R = 4 rows per pass: 4 x N partial logits per thread, warp shuffles, one shared-memory hop across the 8 warps, then
//     softmax / loss / dz in 16-lane groups (lane = class) and dz broadcast through shared memory.
// ---------------------------------------------------------------------------------------------
// R rows per pass (1 for small batches: more CTAs, a shorter dependent stream per warp); N is a compile-time constant
// (10 classes, construct_distribute.py:261-264) so that the class loops carry no predicates.
template <int R, int N>
__global__ void __launch_bounds__(256)
head_rows_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                 const float* __restrict__ y, float* __restrict__ logits, float* __restrict__ dz, float* __restrict__ dx,
                 int M, int K, int loss_name, float grad_scale, int relu_mask, int round_dx,
                 float* __restrict__ cs_partial, float* __restrict__ db_out, float* __restrict__ loss_partial,
                 int* __restrict__ hit_partial, unsigned* counter, float* __restrict__ loss_out, int* __restrict__ correct_out) {
    constexpr int NMAX = 16;
    static_assert(N <= NMAX && R <= 4, "16-lane class groups, at most four rows per pass");
    extern __shared__ __align__(16) float sw[];      // W[K][N] staging
    __shared__ float s_part[8][R][NMAX];
    __shared__ float s_dz[R][NMAX];
    __shared__ float s_l[R];
    __shared__ int s_h[R];
    const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
    {
        const int total = K * N, total4 = ((reinterpret_cast<uintptr_t>(w) & 15) == 0) ? total / 4 : 0;
        const float4* w4 = reinterpret_cast<const float4*>(w);
        float4* s4 = reinterpret_cast<float4*>(sw);
        // K * N / 4 <= 2560 float4 (K <= 1024, N = 10): at most 10 per thread, all in flight at once
        float4 v[10];
#pragma unroll
        for (int u = 0; u < 10; ++u) if (t + 256 * u < total4) v[u] = __ldg(w4 + t + 256 * u);
#pragma unroll
        for (int u = 0; u < 10; ++u) if (t + 256 * u < total4) s4[t + 256 * u] = v[u];
        for (int i = t + 2560; i < total4; i += 256) s4[i] = __ldg(w4 + i);
        for (int j = total4 * 4 + t; j < total; j += 256) sw[j] = __ldg(w + j);
    }
    __syncthreads();
    float wq[4][N];
    bool own[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        own[i] = t + 256 * i < K;
#pragma unroll
        for (int n = 0; n < N; ++n) wq[i][n] = own[i] ? sw[(t + 256 * i) * N + n] : 0.0f;
    }
    float cs[4] = {0.f, 0.f, 0.f, 0.f};
    float lsum = 0.0f; int hits = 0;           // meaningful in lane 0 / 16 of warps 0 and 1
    for (int m0 = blockIdx.x * R; m0 < M; m0 += gridDim.x * R) {
        float xv[R][4];
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int i = 0; i < 4; ++i)
                xv[r][i] = (own[i] && m0 + r < M) ? __ldg(x + (long long)(m0 + r) * K + t + 256 * i) : 0.0f;
#pragma unroll
        for (int r = 0; r < R; ++r) {
#pragma unroll
            for (int n = 0; n < N; ++n) {
                float v = fmaf(xv[r][3], wq[3][n], fmaf(xv[r][2], wq[2][n], fmaf(xv[r][1], wq[1][n], xv[r][0] * wq[0][n])));
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) s_part[warp][r][n] = v;
            }
        }
        __syncthreads();
        if (warp < 2) {
            // 16-lane group per row: lane n of the group is class n
            const int r = warp * 2 + (lane >> 4), n = lane & 15, m = m0 + r;
            const bool live = n < N && r < R && m < M;
            float z = -INFINITY, yv = -INFINITY;
            if (live) {
                float v = 0.0f;
#pragma unroll
                for (int q = 0; q < 8; ++q) v += s_part[q][r][n];
                z = v + (bias ? __ldg(bias + n) : 0.0f);
                yv = __ldg(y + (long long)m * N + n);
            }
            // first maximum wins (argmax semantics of tf.argmax): compare (value, -index)
            float zm = z, ym = yv; int za = n, ya = n;
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                const float oz = __shfl_xor_sync(0xffffffffu, zm, o), oy = __shfl_xor_sync(0xffffffffu, ym, o);
                const int oza = __shfl_xor_sync(0xffffffffu, za, o), oya = __shfl_xor_sync(0xffffffffu, ya, o);
                if (oz > zm || (oz == zm && oza < za)) { zm = oz; za = oza; }
                if (oy > ym || (oy == ym && oya < ya)) { ym = oy; ya = oya; }
            }
            float l = 0.0f, d = 0.0f;
            if (loss_name == 0) {
                float e = live ? expf(z - zm) : 0.0f, se = e;
#pragma unroll
                for (int o = 8; o > 0; o >>= 1) se += __shfl_xor_sync(0xffffffffu, se, o);
                const float lse = zm + logf(se);
                if (live) {
                    const float logp = z - lse;
                    l = -yv * logp;
                    d = (expf(logp) - yv) * grad_scale / (float)M;
                }
            } else if (live) {
                const float e = z - yv;
                l = e * e / (float)N;
                d = 2.0f * e * grad_scale / ((float)M * (float)N);
            }
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
            if (live) {
                s_dz[r][n] = d;
                logits[(long long)m * N + n] = z;
                if (dz) dz[(long long)m * N + n] = d;
            }
            if (n == 0 && r < R && m < M) { lsum += l; hits += (za == ya); }
        }
        __syncthreads();
        if (dx) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (m0 + r < M) {
                    float o[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int n = 0; n < N; ++n) {
                        const float d = s_dz[r][n];
#pragma unroll
                        for (int i = 0; i < 4; ++i) o[i] = fmaf(d, wq[i][n], o[i]);
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        if (own[i]) {
                            float v = o[i];
                            if (relu_mask && !(xv[r][i] > 0.0f)) v = 0.0f;
                            if (round_dx) v = rn_tf32(v);
                            dx[(long long)(m0 + r) * K + t + 256 * i] = v;
                            cs[i] += v;
                        }
                    }
                }
            }
        }
        __syncthreads();
    }
    if (warp < 2 && (lane & 15) == 0 && warp * 2 + (lane >> 4) < R) { s_l[warp * 2 + (lane >> 4)] = lsum; s_h[warp * 2 + (lane >> 4)] = hits; }
    if (cs_partial) {
#pragma unroll
        for (int i = 0; i < 4; ++i) if (own[i]) cs_partial[(long long)blockIdx.x * K + t + 256 * i] = cs[i];
    }
    __syncthreads();
    if (t == 0) {
        float a = 0.0f; int hh = 0;
        for (int i = 0; i < R; ++i) { a += s_l[i]; hh += s_h[i]; }
        loss_partial[blockIdx.x] = a; hit_partial[blockIdx.x] = hh;
    }
    // Two-level, fixed-order sum over the CTAs (one CTA summing every partial row alone is a chain of G dependent-latency
    // loads per column -- measured 30-70 us at G = 256-592): the last CTA of each group of 16 sums its group, the last
    // group to finish sums the group rows.  group counters: counter[1 + group]; cs_partial rows [G, G + ngroups) and
    // loss_partial / hit_partial entries [G, G + ngroups) hold the group sums.
    const int G = gridDim.x, ngroups = (G + 15) / 16, grp = blockIdx.x >> 4;
    const int g0 = grp * 16, gsize = min(16, G - g0);
    if (!last_block_arrives(counter + 1 + grp, gsize)) return;
    if (warp == 0) {
        float a = 0.0f; int hh = 0;
        if (lane < gsize) { a = __ldcg(loss_partial + g0 + lane); hh = __ldcg(hit_partial + g0 + lane); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); hh += __shfl_xor_sync(0xffffffffu, hh, o); }
        if (lane == 0) { loss_partial[G + grp] = a; hit_partial[G + grp] = hh; }
    }
    if (cs_partial) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (!own[i]) continue;
            const int kk = t + 256 * i;
            float v[16];
#pragma unroll
            for (int u = 0; u < 16; ++u) v[u] = u < gsize ? __ldcg(cs_partial + (long long)(g0 + u) * K + kk) : 0.0f;
            float a = 0.0f;
#pragma unroll
            for (int u = 0; u < 16; ++u) a += v[u];
            cs_partial[(long long)(G + grp) * K + kk] = a;
        }
    }
    if (!last_block_arrives(counter, ngroups)) return;
    if (warp == 0) {
        float a = 0.0f; int hh = 0;
        for (int b = lane; b < ngroups; b += 32) { a += __ldcg(loss_partial + G + b); hh += __ldcg(hit_partial + G + b); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); hh += __shfl_xor_sync(0xffffffffu, hh, o); }
        if (lane == 0) {
            if (loss_out) *loss_out = a / (float)M;
            if (correct_out) *correct_out = hh;
        }
    }
    if (cs_partial && db_out) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (!own[i]) continue;
            const int kk = t + 256 * i;
            float a = 0.0f;
            int b = 0;
            for (; b + 7 < ngroups; b += 8) {          // eight independent L2 loads in flight, summed in group order
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = __ldcg(cs_partial + (long long)(G + b + u) * K + kk);
#pragma unroll
                for (int u = 0; u < 8; ++u) a += v[u];
            }
            for (; b < ngroups; ++b) a += __ldcg(cs_partial + (long long)(G + b) * K + kk);
            db_out[kk] = a;
        }
    }
}

// dW[K][N] = x^T . dz and db[N] without slice partials for small batches: a CTA owns 32 consecutive k and (a slice of) the
// batch; lane = k, warp = row lane (32 of them: 1024 threads), dz rows are warp-uniform broadcast loads; the 32 row lanes are
// summed through shared memory in warp order.  gridDim.y > 1 (large batches): slices summed by the last CTA of the k block.
template <int NMAX>
__global__ void __launch_bounds__(1024)
head_wgrad_rows_kernel(const float* __restrict__ x, const float* __restrict__ dz, float* __restrict__ partial, float* __restrict__ dW,
                       float* __restrict__ db, int M, int N, int K, int rows_per_slice, unsigned* counters) {
    extern __shared__ float sred[];                   // [32 warps][32 k][N]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int k = blockIdx.x * 32 + lane;
    const int m0 = blockIdx.y * rows_per_slice, m1 = min(M, m0 + rows_per_slice);
    float acc[NMAX];
#pragma unroll
    for (int n = 0; n < NMAX; ++n) acc[n] = 0.0f;
    if (k < K) {
        int m = m0 + warp;
        for (; m + 96 < m1; m += 128) {               // four independent x loads in flight per thread
            float xv[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) xv[u] = __ldg(x + (long long)(m + 32 * u) * K + k);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float* dr = dz + (long long)(m + 32 * u) * N;
#pragma unroll
                for (int n = 0; n < NMAX; ++n)
                    if (n < N) acc[n] = fmaf(xv[u], __ldg(dr + n), acc[n]);
            }
        }
        for (; m < m1; m += 32) {
            const float xv = __ldg(x + (long long)m * K + k);
            const float* dr = dz + (long long)m * N;
#pragma unroll
            for (int n = 0; n < NMAX; ++n)
                if (n < N) acc[n] = fmaf(xv, __ldg(dr + n), acc[n]);
        }
    }
#pragma unroll
    for (int n = 0; n < NMAX; ++n)
        if (n < N) sred[(warp * 32 + lane) * N + n] = acc[n];
    __syncthreads();
    const int S = gridDim.y;
    float* out = S > 1 ? partial + (long long)blockIdx.y * K * N : dW;
    if ((int)threadIdx.x < 32 * N) {
        const int kl = threadIdx.x / N, n = threadIdx.x - kl * N;
        float v = 0.0f;
#pragma unroll 8
        for (int q = 0; q < 32; ++q) v += sred[(q * 32 + kl) * N + n];
        if (blockIdx.x * 32 + kl < K) out[(long long)(blockIdx.x * 32 + kl) * N + n] = v;
    }
    if (S > 1) {
        if (!last_block_arrives(counters + blockIdx.x, S)) return;
        if ((int)threadIdx.x < 32 * N) {
            const int kl = threadIdx.x / N, n = threadIdx.x - kl * N;
            if (blockIdx.x * 32 + kl < K) {
                float v = 0.0f;
                for (int s = 0; s < S; ++s) v += __ldcg(partial + ((long long)s * K + blockIdx.x * 32 + kl) * N + n);
                dW[(long long)(blockIdx.x * 32 + kl) * N + n] = v;
            }
        }
    }
    if (blockIdx.x == 0 && db) {
        // column sums of dz over the WHOLE batch (every slice CTA of k block 0 that gets here computes the same bytes)
        __syncthreads();
        const int n = threadIdx.x & 15, r0 = threadIdx.x >> 4;           // 64 row lanes x 16 columns
        float v = 0.0f;
        if (n < N)
            for (int m = r0; m < M; m += 64) v += __ldg(dz + (long long)m * N + n);
        sred[r0 * 16 + n] = v;
        __syncthreads();
        if ((int)threadIdx.x < N) {
            float u = 0.0f;
            for (int r = 0; r < 64; ++r) u += sred[r * 16 + threadIdx.x];
            db[threadIdx.x] = u;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// per-channel sums of what a vec4 elementwise kernel wrote: thread t only ever touches channel group t % c4n (the host
// checks 256 % c4n == 0, which also makes the grid stride a multiple of c4n), so one float4 per thread accumulates them.
// CTA partial -> partial[cta][C]; the last CTA sums the CTAs in order.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void channel_sums_finish(float4 cs, int c4n, float* __restrict__ partial, float* __restrict__ db,
                                                    unsigned* counter) {
    __shared__ float4 sh4[256];
    const int C = c4n * 4;
    sh4[threadIdx.x] = cs;
    __syncthreads();
    if ((int)threadIdx.x < c4n) {
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int j = threadIdx.x; j < 256; j += c4n) { const float4 v = sh4[j]; s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
        reinterpret_cast<float4*>(partial + (long long)blockIdx.x * C)[threadIdx.x] = s;
    }
    if (!last_block_arrives(counter, gridDim.x)) return;
    // 256 / c4n thread rows, each sums every (256 / c4n)-th CTA for its channel group; then the rows are added in order
    const int g = threadIdx.x % c4n, r = threadIdx.x / c4n, R = 256 / c4n, G = gridDim.x;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    int b = r;
    for (; b + 7 * R < G; b += 8 * R) {                 // eight independent L2 loads in flight
        float4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = __ldcg(reinterpret_cast<const float4*>(partial + (long long)(b + u * R) * C) + g);
#pragma unroll
        for (int u = 0; u < 8; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
    }
    for (; b < G; b += R) {
        const float4 v = __ldcg(reinterpret_cast<const float4*>(partial + (long long)b * C) + g);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    __syncthreads();
    sh4[threadIdx.x] = s;
    __syncthreads();
    if ((int)threadIdx.x < c4n) {
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int j = threadIdx.x; j < 256; j += c4n) { const float4 v = sh4[j]; t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w; }
        reinterpret_cast<float4*>(db)[threadIdx.x] = t;
    }
}

// maxpool_bwd_tiled_vec4_kernel + relu mask + optional TF32 rounding + the per-channel sums of the gradient it scatters
// (= the bias gradient of the convolution that feeds the pool).  blockDim.x == 256, 256 % (c / 4) == 0.
__global__ void __launch_bounds__(256)
maxpool_bwd_tiled_vec4_cs_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ arg, float* __restrict__ dx,
                                 PoolGeom g, long long total4, const float* __restrict__ pooled, int round_out,
                                 float* __restrict__ partial, float* __restrict__ db, unsigned* counter) {
    const int c4n = g.c >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    float4 cs = make_float4(0.f, 0.f, 0.f, 0.f);
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total4; idx += stride) {
        const int c4 = idx % c4n; long long t = idx / c4n;
        const int ox = t % g.ow; t /= g.ow;
        const int oy = t % g.oh; const int n = (int)(t / g.oh);
        float4 gv = __ldg(reinterpret_cast<const float4*>(dy) + idx);
        if (pooled) {
            const float4 p = __ldg(reinterpret_cast<const float4*>(pooled) + idx);
            if (!(p.x > 0.0f)) gv.x = 0.0f;
            if (!(p.y > 0.0f)) gv.y = 0.0f;
            if (!(p.z > 0.0f)) gv.z = 0.0f;
            if (!(p.w > 0.0f)) gv.w = 0.0f;
        }
        if (round_out) { gv.x = rn_tf32(gv.x); gv.y = rn_tf32(gv.y); gv.z = rn_tf32(gv.z); gv.w = rn_tf32(gv.w); }
        cs.x += gv.x; cs.y += gv.y; cs.z += gv.z; cs.w += gv.w;      // the window's argmax is always inside the picture
        const uchar4 a = reinterpret_cast<const uchar4*>(arg)[idx];
        for (int i = 0; i < g.kh; ++i) {
            const int yy = oy * g.sh + i;
            if (yy >= g.ih) continue;
            for (int j = 0; j < g.kw; ++j) {
                const int xx = ox * g.sw + j;
                if (xx >= g.iw) continue;
                const int k = i * g.kw + j;
                reinterpret_cast<float4*>(dx + (((long long)n * g.ih + yy) * g.iw + xx) * g.c)[c4] =
                    make_float4(a.x == k ? gv.x : 0.0f, a.y == k ? gv.y : 0.0f, a.z == k ? gv.z : 0.0f, a.w == k ? gv.w : 0.0f);
            }
        }
    }
    if (partial) channel_sums_finish(cs, c4n, partial, db, counter);
}

// act_bwd_kernel, four channels per thread, + optional TF32 rounding + per-channel sums (partial == nullptr: none)
__global__ void __launch_bounds__(256)
act_bwd_vec4_cs_kernel(const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ dy, float* __restrict__ dx,
                       long long count4, int c4n, int fn, float slope, int round_out, float* __restrict__ partial,
                       float* __restrict__ db, unsigned* counter) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    float4 cs = make_float4(0.f, 0.f, 0.f, 0.f);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count4; i += stride) {
        const float4 g = __ldg(reinterpret_cast<const float4*>(dy) + i);
        float4 r;
        if (fn == ACT_RELU) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
            r = make_float4(v.x > 0.0f ? g.x : 0.0f, v.y > 0.0f ? g.y : 0.0f, v.z > 0.0f ? g.z : 0.0f, v.w > 0.0f ? g.w : 0.0f);
        } else if (fn == ACT_SIGMOID) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(y) + i);
            r = make_float4(g.x * v.x * (1.0f - v.x), g.y * v.y * (1.0f - v.y), g.z * v.z * (1.0f - v.z), g.w * v.w * (1.0f - v.w));
        } else {
            const float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
            r = make_float4(v.x >= slope * v.x ? g.x : slope * g.x, v.y >= slope * v.y ? g.y : slope * g.y,
                            v.z >= slope * v.z ? g.z : slope * g.z, v.w >= slope * v.w ? g.w : slope * g.w);
        }
        if (round_out) { r.x = rn_tf32(r.x); r.y = rn_tf32(r.y); r.z = rn_tf32(r.z); r.w = rn_tf32(r.w); }
        reinterpret_cast<float4*>(dx)[i] = r;
        cs.x += r.x; cs.y += r.y; cs.z += r.z; cs.w += r.w;
    }
    if (partial) channel_sums_finish(cs, c4n, partial, db, counter);
}

}  // namespace nck

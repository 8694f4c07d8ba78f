// fp32 SIMT kernels of the CNN path (math mode 0, and every shape the tensor-core path does not
// take).  All contractions go through one tiled implicit-GEMM kernel parameterised by element
// functors (dense / im2col / transposed views), so conv fwd, dgrad, wgrad and the dense layers
// share one verified inner loop.  TF-1.x semantics as restated in oracle/tfref.py.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "tc_gemm.h"

namespace nck {

// ---------------------------------------------------------------------------------------------
// element functors: value of the virtual matrix at (row, col); rows/cols outside are never asked
// ---------------------------------------------------------------------------------------------
struct RowMajor {           // M x K (or K x N) dense, leading dimension ld
    const float* p; int ld;
    __device__ __forceinline__ float operator()(int r, int c) const { return __ldg(p + (long long)r * ld + c); }
};
struct ColMajor {           // element (r, c) stored at p[c*ld + r]   (a transposed view)
    const float* p; int ld;
    __device__ __forceinline__ float operator()(int r, int c) const { return __ldg(p + (long long)c * ld + r); }
};


// im2col view of the NHWC input: row m = (n, oy, ox), col k = (i, j, c)
struct Im2col {
    const float* x; ConvGeom g;
    __device__ __forceinline__ float operator()(int m, int k) const {
        const int c = k % g.ic; const int ij = k / g.ic; const int j = ij % g.fw, i = ij / g.fw;
        const int ox = m % g.ow; const int t = m / g.ow; const int oy = t % g.oh, n = t / g.oh;
        const int y = oy * g.sh - g.pt + i, xx = ox * g.sw - g.pl + j;
        if (y < 0 || y >= g.ih || xx < 0 || xx >= g.iw) return 0.0f;
        return __ldg(x + (((long long)n * g.ih + y) * g.iw + xx) * g.ic + c);
    }
};
// transposed im2col (for wgrad): row k = (i, j, c), col m = (n, oy, ox)
struct Im2colT {
    Im2col v;
    __device__ __forceinline__ float operator()(int k, int m) const { return v(m, k); }
};
// dgrad view of dY: row m = (n, y, x) of the INPUT, col k = (i, j, co)
struct DgradA {
    const float* dy; ConvGeom g;
    __device__ __forceinline__ float operator()(int m, int k) const {
        const int co = k % g.oc; const int ij = k / g.oc; const int j = ij % g.fw, i = ij / g.fw;
        const int xx = m % g.iw; const int t = m / g.iw; const int y = t % g.ih, n = t / g.ih;
        const int ny = y + g.pt - i, nx = xx + g.pl - j;
        if (ny < 0 || nx < 0 || ny % g.sh || nx % g.sw) return 0.0f;
        const int oy = ny / g.sh, ox = nx / g.sw;
        if (oy >= g.oh || ox >= g.ow) return 0.0f;
        return __ldg(dy + (((long long)n * g.oh + oy) * g.ow + ox) * g.oc + co);
    }
};
// dgrad view of the HWIO filter: row k = (i, j, co), col c
struct DgradB {
    const float* w; ConvGeom g;
    __device__ __forceinline__ float operator()(int k, int c) const {
        const int co = k % g.oc; const int ij = k / g.oc;
        return __ldg(w + ((long long)ij * g.ic + c) * g.oc + co);
    }
};

// ---------------------------------------------------------------------------------------------
// epilogues
// ---------------------------------------------------------------------------------------------
enum { ACT_NONE = 0, ACT_RELU = 1, ACT_SIGMOID = 2, ACT_LEAKY = 3 };

struct EpiStore {           // C[m][n] = acc (+ bias[n])
    float* c; int ldc; const float* bias;
    __device__ __forceinline__ void operator()(int m, int n, float acc, int) const {
        if (bias) acc += __ldg(bias + n);
        c[(long long)m * ldc + n] = acc;
    }
};
struct EpiPartial {         // split-K partial: ws[split][m][n]
    float* ws; int ldc; long long plane;
    __device__ __forceinline__ void operator()(int m, int n, float acc, int split) const {
        ws[(long long)split * plane + (long long)m * ldc + n] = acc;
    }
};

// ---------------------------------------------------------------------------------------------
// tiled SIMT GEMM: C[M,N] = A[M,K] * B[K,N], 64x64x16 tiles, 256 threads, 4x4 per thread.
// AKF: A's contiguous memory axis is k (true) or m (false); BNF: B's contiguous axis is n / k.
// gridDim.z = split-K slices of length kslice.
// ---------------------------------------------------------------------------------------------
constexpr int BM = 64, BN = 64, BK = 16;

template <bool AKF, bool BNF, class AF, class BF, class Epi>
__global__ void __launch_bounds__(256)
gemm_kernel(int M, int N, int K, int kslice, AF A, BF B, Epi epi) {
    __shared__ float As[BK][BM + 4];
    __shared__ float Bs[BK][BN + 4];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int kbeg = blockIdx.z * kslice;
    const int kend = min(K, kbeg + kslice);
    const int tx = tid % 16, ty = tid / 16;       // thread tile: rows ty*4.., cols tx*4..
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

    for (int k0 = kbeg; k0 < kend; k0 += BK) {
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            const int e = tid + l * 256;
            int mm, kk;
            if (AKF) { kk = e % BK; mm = e / BK; } else { mm = e % BM; kk = e / BM; }
            const int gm = m0 + mm, gk = k0 + kk;
            As[kk][mm] = (gm < M && gk < kend) ? A(gm, gk) : 0.0f;
            int nn, kb;
            if (BNF) { nn = e % BN; kb = e / BN; } else { kb = e % BK; nn = e / BK; }
            const int gn = n0 + nn, gkb = k0 + kb;
            Bs[kb][nn] = (gn < N && gkb < kend) ? B(gkb, gn) : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
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
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gm = m0 + ty * 4 + i;
        if (gm >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gn = n0 + tx * 4 + j;
            if (gn < N) epi(gm, gn, acc[i][j], blockIdx.z);
        }
    }
}

// dst[i] = sum_s ws[s*count + i]   (fixed order: deterministic)
__global__ void reduce_partials_kernel(float* __restrict__ dst, const float* __restrict__ ws, int splits, long long count,
                                       const float* __restrict__ bias, int ncols) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        float s = 0.0f;
        for (int k = 0; k < splits; ++k) s += ws[(long long)k * count + i];
        if (bias) s += __ldg(bias + (int)(i % ncols));
        dst[i] = s;
    }
}

// many splits, few outputs: one warp per output, lanes stride over the splits, fixed shuffle tree (deterministic)
__global__ void reduce_partials_warp_kernel(float* __restrict__ dst, const float* __restrict__ ws, int splits, long long count) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long i = warp; i < count; i += nwarps) {
        float s = 0.0f;
        for (int k = lane; k < splits; k += 32) s += ws[(long long)k * count + i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) dst[i] = s;
    }
}

// ---------------------------------------------------------------------------------------------
// the output layer: tf.matmul(h, W) + b with W[K][N], N = classes (10) -- construct_distribute.py:264.
// Far too narrow for a 64-wide GEMM tile; W (K*N floats) fits shared memory instead.
//   forward  one warp per row: lanes stride K (coalesced), NMAX accumulators per lane, shuffle-tree reduction
//   wgrad    one thread per k: dW[k][:] += x[m][k] * dy[m][:] over a slice of the batch, dy rows broadcast from shared
//            memory; slices are summed in fixed order by reduce_partials_kernel
// ---------------------------------------------------------------------------------------------
template <int NMAX>
__global__ void __launch_bounds__(256)
skinny_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias, float* __restrict__ y,
                  int M, int N, int K) {
    extern __shared__ __align__(16) float sw[];      // W[K][N]
    {
        // staged with 16-byte loads, eight in flight per thread (a scalar loop here is one L2 round trip per element)
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
    for (int m = blockIdx.x * 8 + warp; m < M; m += gridDim.x * 8) {
        float acc[NMAX];
#pragma unroll
        for (int n = 0; n < NMAX; ++n) acc[n] = 0.0f;
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
                    if (n < N) acc[n] = fmaf(xv[u], wr[n], acc[n]);
            }
        }
        for (; k < K; k += 32) {
            const float xv = __ldg(xr + k);
            const float* wr = sw + k * N;
#pragma unroll
            for (int n = 0; n < NMAX; ++n)
                if (n < N) acc[n] = fmaf(xv, wr[n], acc[n]);
        }
#pragma unroll
        for (int n = 0; n < NMAX; ++n) {
            if (n < N) {
                float v = acc[n];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) y[(long long)m * N + n] = v + (bias ? __ldg(bias + n) : 0.0f);
            }
        }
    }
}

// partial[slice][K][N]; grid = (ceil(K / 256), slices); rows_per_slice rows of the batch per slice
template <int NMAX>
__global__ void __launch_bounds__(256)
skinny_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ partial, int M, int N, int K,
                    int rows_per_slice) {
    extern __shared__ float sdy[];                   // dy[rows of this slice][N]
    const int m0 = blockIdx.y * rows_per_slice, m1 = min(M, m0 + rows_per_slice);
    for (int i = threadIdx.x; i < (m1 - m0) * N; i += blockDim.x) sdy[i] = __ldg(dy + (long long)m0 * N + i);
    __syncthreads();
    const int k = blockIdx.x * 256 + threadIdx.x;
    if (k >= K) return;
    float acc[NMAX];
#pragma unroll
    for (int n = 0; n < NMAX; ++n) acc[n] = 0.0f;
    for (int m = m0; m < m1; ++m) {
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

// column sums of Y[rows][cols] (bias gradients): partial[split][col]
__global__ void colsum_partial_kernel(const float* __restrict__ y, int rows, int cols, int rows_per_split, float* __restrict__ ws) {
    const int col = blockIdx.x * 32 + (threadIdx.x & 31);
    const int lane_row = threadIdx.x >> 5;            // 8 row lanes
    const int r0 = blockIdx.y * rows_per_split, r1 = min(rows, r0 + rows_per_split);
    float s = 0.0f;
    if (col < cols) {
        int r = r0 + lane_row;
        float p[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) p[u] = 0.0f;
        for (; r + 56 < r1; r += 64) {               // eight independent loads in flight per thread
#pragma unroll
            for (int u = 0; u < 8; ++u) p[u] += __ldg(y + (long long)(r + 8 * u) * cols + col);
        }
        for (; r < r1; r += 8) s += __ldg(y + (long long)r * cols + col);
        s += ((p[0] + p[1]) + (p[2] + p[3])) + ((p[4] + p[5]) + (p[6] + p[7]));
    }
    __shared__ float sh[8][33];
    sh[lane_row][threadIdx.x & 31] = s;
    __syncthreads();
    if (lane_row == 0 && col < cols) {
        float t = 0.0f;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += sh[i][threadIdx.x & 31];
        ws[(long long)blockIdx.y * cols + col] = t;
    }
}

// ---------------------------------------------------------------------------------------------
// direct convolution for few input channels (the first layer: Cin = 1 or 3).  An im2col GEMM with
// K = fh*fw*Cin = 25 wastes a 64x64x16 tile; here one thread owns one output position and 32 output
// channels (weights broadcast from shared memory), and the weight gradient keeps one output channel per
// lane with fh*fw*Cin accumulators in registers, reduced per CTA and then across CTAs in fixed order.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
conv_smallc_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                       float* __restrict__ y, ConvGeom g, long long M) {
    extern __shared__ float s_w[];                       // [fh*fw*ic][ocp], rows zero-padded to a multiple of 32
    const int T = g.fh * g.fw * g.ic;
    const int ocb = (g.oc + 31) >> 5, ocp = ocb * 32;
    for (int i = threadIdx.x; i < T * ocp; i += blockDim.x) {
        const int r = i / ocp, o = i - r * ocp;
        s_w[i] = o < g.oc ? __ldg(w + r * g.oc + o) : 0.0f;
    }
    __syncthreads();
    const long long total = M * ocb;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int ob = (int)(idx % ocb);
        const long long m = idx / ocb;
        const int ox = (int)(m % g.ow); const long long q = m / g.ow; const int oy = (int)(q % g.oh); const int n = (int)(q / g.oh);
        float acc[32];
#pragma unroll
        for (int o = 0; o < 32; ++o) acc[o] = 0.0f;
        for (int i = 0; i < g.fh; ++i) {
            const int yy = oy * g.sh - g.pt + i;
            if (yy < 0 || yy >= g.ih) continue;
            for (int j = 0; j < g.fw; ++j) {
                const int xx = ox * g.sw - g.pl + j;
                if (xx < 0 || xx >= g.iw) continue;
                const float* px = x + (((long long)n * g.ih + yy) * g.iw + xx) * g.ic;
                const float* pw = s_w + (size_t)((i * g.fw + j) * g.ic) * ocp + ob * 32;
                for (int c = 0; c < g.ic; ++c) {
                    const float v = __ldg(px + c);
#pragma unroll
                    for (int o = 0; o < 32; ++o) acc[o] = fmaf(v, pw[c * ocp + o], acc[o]);
                }
            }
        }
        float* py = y + m * g.oc + ob * 32;
#pragma unroll
        for (int o = 0; o < 32; ++o)
            if (ob * 32 + o < g.oc) py[o] = acc[o] + (bias ? __ldg(bias + ob * 32 + o) : 0.0f);
    }
}

// ---- tile versions for T = fh*fw*Cin <= 32 (the usual first layer: 5x5x1, 3x3x3): a CTA builds the im2col of 64
// output positions in shared memory once (bounds handled there), then lane = output channel:
//   forward: the 32 weights of the lane's channel sit in registers, each warp turns 8 positions into 8 coalesced
//            128-byte output rows with two broadcast LDS.128 + 8 FMA per tap;
//   wgrad:   the lane keeps dW[t][channel] for all taps in registers, 8 broadcast LDS.128 + 32 FMA per position,
//            bias gradient for free; warps are summed in shared memory, CTAs in a fixed-order reduction.
struct SmallcTile {
    int pn[64], py[64], px[64];     // decoded output positions of the tile
    int tdy[32], tdx[32], tc[32];   // decoded taps
};

__device__ __forceinline__ void smallc_decode(SmallcTile& S, const ConvGeom& g, long long m0, long long M, int T) {
    const int tid = threadIdx.x;
    if (tid < 64) {
        const long long m = m0 + tid;
        if (m < M) {
            S.px[tid] = (int)(m % g.ow); const long long q = m / g.ow;
            S.py[tid] = (int)(q % g.oh); S.pn[tid] = (int)(q / g.oh);
        } else { S.pn[tid] = -1; S.py[tid] = 0; S.px[tid] = 0; }
    } else if (tid < 96) {
        const int t = tid - 64;
        const int tt = t < T ? t : 0;
        const int c = tt % g.ic, ij = tt / g.ic;
        S.tdy[t] = ij / g.fw; S.tdx[t] = ij % g.fw; S.tc[t] = c;
    }
}
__device__ __forceinline__ float smallc_fetch(const SmallcTile& S, const ConvGeom& g, const float* __restrict__ x, int p, int t, int T) {
    if (t >= T || S.pn[p] < 0) return 0.0f;
    const int yy = S.py[p] * g.sh - g.pt + S.tdy[t], xx = S.px[p] * g.sw - g.pl + S.tdx[t];
    if (yy < 0 || yy >= g.ih || xx < 0 || xx >= g.iw) return 0.0f;
    return __ldg(x + (((long long)S.pn[p] * g.ih + yy) * g.iw + xx) * g.ic + S.tc[t]);
}

__global__ void __launch_bounds__(256)
conv_smallc_fwd_tile_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                            float* __restrict__ y, ConvGeom g, long long M, int ob, int relu) {
    __shared__ SmallcTile S;
    __shared__ __align__(16) float colT[32][64];        // [tap][position]
    const int T = g.fh * g.fw * g.ic;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int co = ob * 32 + lane;
    float wreg[32];
#pragma unroll
    for (int t = 0; t < 32; ++t) wreg[t] = (t < T && co < g.oc) ? __ldg(w + (long long)t * g.oc + co) : 0.0f;
    const float bv = (bias && co < g.oc) ? __ldg(bias + co) : 0.0f;
    const long long ntiles = (M + 63) / 64;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long m0 = tile * 64;
        __syncthreads();                                  // previous tile fully consumed
        smallc_decode(S, g, m0, M, T);
        __syncthreads();
        for (int e = threadIdx.x; e < 32 * 64; e += 256) colT[e >> 6][e & 63] = smallc_fetch(S, g, x, e & 63, e >> 6, T);
        __syncthreads();
        float acc[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = 0.0f;
#pragma unroll
        for (int t = 0; t < 32; ++t) {
            if (t < T) {
                const float4 a = *reinterpret_cast<const float4*>(&colT[t][warp * 8]);
                const float4 b = *reinterpret_cast<const float4*>(&colT[t][warp * 8 + 4]);
                acc[0] = fmaf(a.x, wreg[t], acc[0]); acc[1] = fmaf(a.y, wreg[t], acc[1]);
                acc[2] = fmaf(a.z, wreg[t], acc[2]); acc[3] = fmaf(a.w, wreg[t], acc[3]);
                acc[4] = fmaf(b.x, wreg[t], acc[4]); acc[5] = fmaf(b.y, wreg[t], acc[5]);
                acc[6] = fmaf(b.z, wreg[t], acc[6]); acc[7] = fmaf(b.w, wreg[t], acc[7]);
            }
        }
        if (co < g.oc) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const long long m = m0 + warp * 8 + k;
                if (m < M) { const float v = acc[k] + bv; y[m * g.oc + co] = relu ? fmaxf(v, 0.0f) : v; }
            }
        }
    }
}

// partial_w[block][T*oc] and partial_b[block][oc] hold this CTA's sums for output-channel block `ob`
__global__ void __launch_bounds__(256)
conv_smallc_wgrad_tile_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ partial_w,
                              float* __restrict__ partial_b, ConvGeom g, long long M, int ob, const float* __restrict__ relu_y) {
    __shared__ SmallcTile S;
    __shared__ __align__(16) float col[64][32];         // [position][tap]
    __shared__ float red[8][33];
    const int T = g.fh * g.fw * g.ic;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int co = ob * 32 + lane;
    float acc[32];
#pragma unroll
    for (int t = 0; t < 32; ++t) acc[t] = 0.0f;
    float accb = 0.0f;
    const long long ntiles = (M + 63) / 64;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long m0 = tile * 64;
        __syncthreads();
        smallc_decode(S, g, m0, M, T);
        __syncthreads();
        for (int e = threadIdx.x; e < 64 * 32; e += 256) col[e >> 5][e & 31] = smallc_fetch(S, g, x, e >> 5, e & 31, T);
        __syncthreads();
#pragma unroll 2
        for (int k = 0; k < 8; ++k) {
            const int p = warp * 8 + k;
            const long long m = m0 + p;
            float gv = (m < M && co < g.oc) ? __ldg(dy + m * g.oc + co) : 0.0f;
            // relu_y: the relu behind this convolution was applied in its epilogue and has no backward launch of its own;
            // dy is then the gradient of the relu's output and the mask comes from that output
            if (relu_y && m < M && co < g.oc && !(__ldg(relu_y + m * g.oc + co) > 0.0f)) gv = 0.0f;
            accb += gv;
#pragma unroll
            for (int t4 = 0; t4 < 8; ++t4) {
                const float4 v = *reinterpret_cast<const float4*>(&col[p][t4 * 4]);
                acc[t4 * 4 + 0] = fmaf(v.x, gv, acc[t4 * 4 + 0]); acc[t4 * 4 + 1] = fmaf(v.y, gv, acc[t4 * 4 + 1]);
                acc[t4 * 4 + 2] = fmaf(v.z, gv, acc[t4 * 4 + 2]); acc[t4 * 4 + 3] = fmaf(v.w, gv, acc[t4 * 4 + 3]);
            }
        }
    }
    // warps -> CTA partial, fixed order
#pragma unroll
    for (int t = 0; t <= 32; ++t) {
        if (t < T || t == 32) {
            __syncthreads();
            red[warp][lane] = t < 32 ? acc[t < 32 ? t : 0] : accb;
            __syncthreads();
            if (warp == 0 && co < g.oc) {
                float sum = 0.0f;
#pragma unroll
                for (int k = 0; k < 8; ++k) sum += red[k][lane];
                if (t < 32) partial_w[(size_t)blockIdx.x * T * g.oc + (size_t)t * g.oc + co] = sum;
                else partial_b[(size_t)blockIdx.x * g.oc + co] = sum;
            }
        }
    }
}

// im2col for few input channels: col[m][kp] (kp = K rounded up to 32, zero padded) so that the first
// layer can ride the tensor-core GEMM as well (forward: col . W, weight gradient: col^T . dY)
__global__ void im2col_smallc_kernel(const float* __restrict__ x, float* __restrict__ col, ConvGeom g, long long M, int K, int Kp) {
    const long long total = M * (Kp / 4);
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int k4 = (int)(idx % (Kp / 4)) * 4;
        const long long m = idx / (Kp / 4);
        const int ox = (int)(m % g.ow); const long long q = m / g.ow; const int oy = (int)(q % g.oh); const int n = (int)(q / g.oh);
        float v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int k = k4 + u;
            float r = 0.0f;
            if (k < K) {
                const int c = k % g.ic, ij = k / g.ic, j = ij % g.fw, i = ij / g.fw;
                const int yy = oy * g.sh - g.pt + i, xx = ox * g.sw - g.pl + j;
                if (yy >= 0 && yy < g.ih && xx >= 0 && xx < g.iw) r = __ldg(x + (((long long)n * g.ih + yy) * g.iw + xx) * g.ic + c);
            }
            v[u] = r;
        }
        *reinterpret_cast<float4*>(col + m * Kp + k4) = make_float4(v[0], v[1], v[2], v[3]);
    }
}

// partial_w[block][T*oc], partial_b[block][oc]; T = fh*fw*ic <= 32
__global__ void __launch_bounds__(256)
conv_smallc_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ partial_w,
                         float* __restrict__ partial_b, ConvGeom g, long long M) {
    __shared__ float s_red[8][33];
    __shared__ int s_tdy[32], s_tdx[32], s_tc[32];
    const int T = g.fh * g.fw * g.ic;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
    if (threadIdx.x < 32) {
        const int t = threadIdx.x < T ? threadIdx.x : 0;
        const int c = t % g.ic, ij = t / g.ic;
        s_tdy[threadIdx.x] = ij / g.fw; s_tdx[threadIdx.x] = ij % g.fw; s_tc[threadIdx.x] = c;
    }
    __syncthreads();
    const int ocb = (g.oc + 31) >> 5;
    for (int ob = 0; ob < ocb; ++ob) {
        const int co = ob * 32 + lane;
        float acc[32];
#pragma unroll
        for (int t = 0; t < 32; ++t) acc[t] = 0.0f;
        float accb = 0.0f;
        for (long long m = (long long)blockIdx.x * nwarp + warp; m < M; m += (long long)gridDim.x * nwarp) {
            const float gv = co < g.oc ? __ldg(dy + m * g.oc + co) : 0.0f;
            accb += gv;
            const int ox = (int)(m % g.ow); const long long q = m / g.ow; const int oy = (int)(q % g.oh); const int n = (int)(q / g.oh);
            const int y0 = oy * g.sh - g.pt, x0 = ox * g.sw - g.pl;
            const float* xb = x + (long long)n * g.ih * g.iw * g.ic;
#pragma unroll
            for (int t = 0; t < 32; ++t) {            // t is a compile-time register index; T is uniform
                if (t < T) {
                    const int yy = y0 + s_tdy[t], xx = x0 + s_tdx[t];
                    const bool ok = yy >= 0 && yy < g.ih && xx >= 0 && xx < g.iw;
                    const float v = ok ? __ldg(xb + ((long long)yy * g.iw + xx) * g.ic + s_tc[t]) : 0.0f;
                    acc[t] = fmaf(v, gv, acc[t]);
                }
            }
        }
        // reduce the warps of this CTA in fixed order, one tap at a time
        for (int t = 0; t <= T; ++t) {
            float v = accb;
            if (t < T) {
#pragma unroll
                for (int u = 0; u < 32; ++u) if (u == t) v = acc[u];
            }
            __syncthreads();
            s_red[warp][lane] = v;
            __syncthreads();
            if (warp == 0 && co < g.oc) {
                float sum = 0.0f;
                for (int k = 0; k < nwarp; ++k) sum += s_red[k][lane];
                if (t < T) partial_w[(size_t)blockIdx.x * T * g.oc + (size_t)t * g.oc + co] = sum;
                else partial_b[(size_t)blockIdx.x * g.oc + co] = sum;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// pooling  (tf.nn.max_pool, construct_distribute.py:121-130)
// ---------------------------------------------------------------------------------------------
struct PoolGeom { int ih, iw, c, oh, ow, kh, kw, sh, sw, pt, pl; };

// relu != 0: the input is a pre-activation and relu is applied on the fly (active(relu) + max_pool in one pass)
__global__ void maxpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, uint8_t* __restrict__ arg,
                                   PoolGeom g, long long total, int relu) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
        const int c = idx % g.c; long long t = idx / g.c;
        const int ox = t % g.ow; t /= g.ow;
        const int oy = t % g.oh; const int n = (int)(t / g.oh);
        float best = -INFINITY; int bi = 0;
        for (int i = 0; i < g.kh; ++i) {
            const int yy = oy * g.sh - g.pt + i;
            if (yy < 0 || yy >= g.ih) continue;
            for (int j = 0; j < g.kw; ++j) {
                const int xx = ox * g.sw - g.pl + j;
                if (xx < 0 || xx >= g.iw) continue;
                float v = __ldg(x + (((long long)n * g.ih + yy) * g.iw + xx) * g.c + c);
                if (relu) v = fmaxf(v, 0.0f);
                if (v > best) { best = v; bi = i * g.kw + j; }   // strict '>': first maximum wins
            }
        }
        y[idx] = best;
        arg[idx] = (uint8_t)bi;
    }
}

// four channels per thread (c % 4 == 0, 16-byte aligned tensors): one index decode and 128-bit accesses per window tap
__global__ void maxpool_fwd_vec4_kernel(const float* __restrict__ x, float* __restrict__ y, uint8_t* __restrict__ arg,
                                        PoolGeom g, long long total4, int relu) {
    const int c4n = g.c >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total4; idx += stride) {
        const int c4 = idx % c4n; long long t = idx / c4n;
        const int ox = t % g.ow; t /= g.ow;
        const int oy = t % g.oh; const int n = (int)(t / g.oh);
        float b0 = -INFINITY, b1 = -INFINITY, b2 = -INFINITY, b3 = -INFINITY;
        int a0 = 0, a1 = 0, a2 = 0, a3 = 0;
        for (int i = 0; i < g.kh; ++i) {
            const int yy = oy * g.sh - g.pt + i;
            if (yy < 0 || yy >= g.ih) continue;
            for (int j = 0; j < g.kw; ++j) {
                const int xx = ox * g.sw - g.pl + j;
                if (xx < 0 || xx >= g.iw) continue;
                float4 v = __ldg(reinterpret_cast<const float4*>(x + (((long long)n * g.ih + yy) * g.iw + xx) * g.c) + c4);
                if (relu) { v.x = fmaxf(v.x, 0.0f); v.y = fmaxf(v.y, 0.0f); v.z = fmaxf(v.z, 0.0f); v.w = fmaxf(v.w, 0.0f); }
                const int k = i * g.kw + j;
                if (v.x > b0) { b0 = v.x; a0 = k; }
                if (v.y > b1) { b1 = v.y; a1 = k; }
                if (v.z > b2) { b2 = v.z; a2 = k; }
                if (v.w > b3) { b3 = v.w; a3 = k; }
            }
        }
        reinterpret_cast<float4*>(y)[idx] = make_float4(b0, b1, b2, b3);
        reinterpret_cast<uchar4*>(arg)[idx] = make_uchar4((unsigned char)a0, (unsigned char)a1, (unsigned char)a2, (unsigned char)a3);
    }
}

__global__ void maxpool_bwd_tiled_vec4_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ arg, float* __restrict__ dx,
                                              PoolGeom g, long long total4, const float* __restrict__ pooled) {
    const int c4n = g.c >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
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
}

// gather form (deterministic, handles overlapping windows): each input element sums the dy of the
// windows whose argmax it is
__global__ void maxpool_bwd_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ arg, float* __restrict__ dx,
                                   PoolGeom g, long long total_in) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total_in; idx += stride) {
        const int c = idx % g.c; long long t = idx / g.c;
        const int xx = t % g.iw; t /= g.iw;
        const int yy = t % g.ih; const int n = (int)(t / g.ih);
        float s = 0.0f;
        for (int i = 0; i < g.kh; ++i) {
            const int ny = yy + g.pt - i;
            if (ny < 0 || ny % g.sh) continue;
            const int oy = ny / g.sh;
            if (oy >= g.oh) continue;
            for (int j = 0; j < g.kw; ++j) {
                const int nx = xx + g.pl - j;
                if (nx < 0 || nx % g.sw) continue;
                const int ox = nx / g.sw;
                if (ox >= g.ow) continue;
                const long long o = (((long long)n * g.oh + oy) * g.ow + ox) * g.c + c;
                if (arg[o] == i * g.kw + j) s += __ldg(dy + o);
            }
        }
        dx[idx] = s;
    }
}

// scatter form for windows that tile the input exactly (kernel == stride, no padding): one thread per
// window writes its kh*kw input positions (dy at the argmax, 0 elsewhere) -- no searching, coalesced in c
// pooled != nullptr: the pool input was relu(pre-activation); the relu mask at the argmax is pooled > 0, so the
// result is the gradient w.r.t. the pre-activation (max_pool backward + relu backward in one pass)
__global__ void maxpool_bwd_tiled_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ arg, float* __restrict__ dx,
                                         PoolGeom g, long long total_out, const float* __restrict__ pooled) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total_out; idx += stride) {
        const int c = idx % g.c; long long t = idx / g.c;
        const int ox = t % g.ow; t /= g.ow;
        const int oy = t % g.oh; const int n = (int)(t / g.oh);
        float gv = __ldg(dy + idx);
        if (pooled && !(__ldg(pooled + idx) > 0.0f)) gv = 0.0f;
        const int a = arg[idx];
        for (int i = 0; i < g.kh; ++i) {
            const int yy = oy * g.sh + i;
            if (yy >= g.ih) continue;
            for (int j = 0; j < g.kw; ++j) {
                const int xx = ox * g.sw + j;
                if (xx >= g.iw) continue;
                dx[(((long long)n * g.ih + yy) * g.iw + xx) * g.c + c] = (a == i * g.kw + j) ? gv : 0.0f;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// first block fused: conv2d (few input channels, stride 1) -> relu -> max_pool 2x2 / 2
// (cnn() on the usual LeNet-style net_config, construct_distribute.py:208-243: the first conv + active + pool).
// The un-pooled activation (N x 28 x 28 x 32 floats for LeNet) is never written or read:
//   forward   one CTA per picture: the zero-padded input tile sits in shared memory, lane = output channel keeps its
//             <= 32 weights in registers, a warp computes the 2x2 conv outputs of one pooled pixel from broadcast
//             shared loads and stores the pooled row (128 B) + argmax bytes.  Same fma order as conv_smallc_fwd_tile_kernel
//             and the same first-maximum rule as maxpool_fwd_kernel: identical bytes.
//   backward  weight / bias gradient straight from the POOLED gradient: only the argmax position of each window carries
//             gradient (and only if the pooled value is > 0), so a pooled pixel costs one 25-tap pass instead of four.
// ---------------------------------------------------------------------------------------------
struct FirstGeom {
    int ih, iw, ic;          // input
    int fh, fw, pt, pl;      // conv filter and padding before
    int coh, cow, oc;        // conv output
    int poh, pw_;            // pooled output (pw_ = pooled width)
    int th, tw;              // shared tile: rows / columns of input positions incl. padding (2*poh + fh - 1, 2*pw_ + fw - 1)
};

__device__ __forceinline__ void first_load_tile(float* tile, const float* __restrict__ xi, const FirstGeom& g) {
    const int total = g.th * g.tw * g.ic;
    for (int e = threadIdx.x; e < total; e += blockDim.x) {
        const int ch = e % g.ic; const int q = e / g.ic;
        const int c = q % g.tw, r = q / g.tw;
        const int iy = r - g.pt, ix = c - g.pl;
        tile[e] = (iy >= 0 && iy < g.ih && ix >= 0 && ix < g.iw) ? __ldg(xi + ((long long)iy * g.iw + ix) * g.ic + ch) : 0.0f;
    }
}

__global__ void __launch_bounds__(256)
first_conv_relu_pool_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                                float* __restrict__ y, uint8_t* __restrict__ arg, FirstGeom g, int n, int ob) {
    extern __shared__ float tile[];
    const int T = g.fh * g.fw * g.ic;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int co = ob * 32 + lane;
    float wreg[32];
    int off[32];                                  // tile offset of tap t relative to the window origin (uniform)
#pragma unroll
    for (int t = 0; t < 32; ++t) {
        wreg[t] = (t < T && co < g.oc) ? __ldg(w + (long long)t * g.oc + co) : 0.0f;
        const int tt = t < T ? t : 0;
        const int c = tt % g.ic, ij = tt / g.ic;
        off[t] = ((ij / g.fw) * g.tw + (ij % g.fw)) * g.ic + c;
    }
    const float bv = (bias && co < g.oc) ? __ldg(bias + co) : 0.0f;
    const int rowp = g.tw * g.ic;                 // one tile row
    const int npool = g.poh * g.pw_;
    for (int img = blockIdx.x; img < n; img += gridDim.x) {
        __syncthreads();
        first_load_tile(tile, x + (long long)img * g.ih * g.iw * g.ic, g);
        __syncthreads();
        for (int pp = warp; pp < npool; pp += 8) {
            const int py = pp / g.pw_, px = pp - py * g.pw_;
            const float* b00 = tile + (2 * py) * rowp + (2 * px) * g.ic;
            float a00 = 0.0f, a01 = 0.0f, a10 = 0.0f, a11 = 0.0f;
#pragma unroll
            for (int t = 0; t < 32; ++t) {
                if (t < T) {
                    const float* q = b00 + off[t];
                    a00 = fmaf(q[0], wreg[t], a00);
                    a01 = fmaf(q[g.ic], wreg[t], a01);
                    a10 = fmaf(q[rowp], wreg[t], a10);
                    a11 = fmaf(q[rowp + g.ic], wreg[t], a11);
                }
            }
            // relu, then the first maximum of the window in scan order; positions past the conv output do not exist
            const bool vy = 2 * py + 1 < g.coh, vx = 2 * px + 1 < g.cow;
            float best = fmaxf(a00 + bv, 0.0f); int bi = 0;
            if (vx) { const float v = fmaxf(a01 + bv, 0.0f); if (v > best) { best = v; bi = 1; } }
            if (vy) { const float v = fmaxf(a10 + bv, 0.0f); if (v > best) { best = v; bi = 2; } }
            if (vy && vx) { const float v = fmaxf(a11 + bv, 0.0f); if (v > best) { best = v; bi = 3; } }
            if (co < g.oc) {
                const long long o = ((long long)img * npool + pp) * g.oc + co;
                y[o] = best;
                arg[o] = (uint8_t)bi;
            }
        }
    }
}

// Single input channel, compile-time filter (5x5 / 3x3: the LeNet-style heads): the (FH+1) x (FW+1) input patch of a
// pooled pixel is read with 8-byte shared loads into registers (18 loads instead of 100 for 5x5) and every weight is used
// four times from its register.  Same fma order per output as the generic kernel: identical bytes.
template <int FH, int FW>
__global__ void __launch_bounds__(256, 3)
first_conv_relu_pool_fwd_c1_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                                   float* __restrict__ y, uint8_t* __restrict__ arg, FirstGeom g, int n, int ob) {
    extern __shared__ __align__(16) float tile[];
    constexpr int T = FH * FW, PWD = (FW + 2) & ~1;          // patch columns loaded (even)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int co = ob * 32 + lane;
    float wreg[T];
#pragma unroll
    for (int t = 0; t < T; ++t) wreg[t] = co < g.oc ? __ldg(w + (long long)t * g.oc + co) : 0.0f;
    const float bv = (bias && co < g.oc) ? __ldg(bias + co) : 0.0f;
    const int npool = g.poh * g.pw_;
    for (int img = blockIdx.x; img < n; img += gridDim.x) {
        __syncthreads();
        first_load_tile(tile, x + (long long)img * g.ih * g.iw, g);
        __syncthreads();
        for (int pp = warp; pp < npool; pp += 8) {
            const int py = pp / g.pw_, px = pp - py * g.pw_;
            const float* b = tile + (2 * py) * g.tw + 2 * px;        // tw even, 2 px even: 8-byte aligned
            float pt[FH + 1][PWD];
#pragma unroll
            for (int r = 0; r <= FH; ++r)
#pragma unroll
                for (int c = 0; c < PWD; c += 2) {
                    const float2 v = *reinterpret_cast<const float2*>(b + r * g.tw + c);
                    pt[r][c] = v.x; pt[r][c + 1] = v.y;
                }
            float a00 = 0.0f, a01 = 0.0f, a10 = 0.0f, a11 = 0.0f;
#pragma unroll
            for (int r = 0; r < FH; ++r)
#pragma unroll
                for (int c = 0; c < FW; ++c) {
                    const float wv = wreg[r * FW + c];
                    a00 = fmaf(pt[r][c], wv, a00);
                    a01 = fmaf(pt[r][c + 1], wv, a01);
                    a10 = fmaf(pt[r + 1][c], wv, a10);
                    a11 = fmaf(pt[r + 1][c + 1], wv, a11);
                }
            const bool vy = 2 * py + 1 < g.coh, vx = 2 * px + 1 < g.cow;
            float best = fmaxf(a00 + bv, 0.0f); int bi = 0;
            if (vx) { const float v = fmaxf(a01 + bv, 0.0f); if (v > best) { best = v; bi = 1; } }
            if (vy) { const float v = fmaxf(a10 + bv, 0.0f); if (v > best) { best = v; bi = 2; } }
            if (vy && vx) { const float v = fmaxf(a11 + bv, 0.0f); if (v > best) { best = v; bi = 3; } }
            if (co < g.oc) {
                const long long o = ((long long)img * npool + pp) * g.oc + co;
                y[o] = best;
                arg[o] = (uint8_t)bi;
            }
        }
    }
}

// Two horizontally adjacent pooled pixels per warp pass (5x5, one input channel, pooled width even, tile rows a multiple of
// 16 bytes): their common 6 x 8 input patch is 12 16-byte shared loads for 200 fmas, and the loop overhead (addresses, the
// argmax selects, the stores) is paid once per two pixels -- the one-pixel kernel above spends 60 % of its issue slots on it.
// Same fma order per output: identical bytes.
__global__ void __launch_bounds__(256, 3)
first_conv_relu_pool_fwd_c1x2_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                                     float* __restrict__ y, uint8_t* __restrict__ arg, FirstGeom g, int n, int ob) {
    extern __shared__ __align__(16) float tile[];
    constexpr int FH = 5, FW = 5, T = FH * FW;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int co = ob * 32 + lane;
    float wreg[T];
#pragma unroll
    for (int t = 0; t < T; ++t) wreg[t] = co < g.oc ? __ldg(w + (long long)t * g.oc + co) : 0.0f;
    const float bv = (bias && co < g.oc) ? __ldg(bias + co) : 0.0f;
    const int npool = g.poh * g.pw_;
    const int ppr = g.pw_ >> 1, npairs = g.poh * ppr;
    const int tw = g.tw;
    for (int img = blockIdx.x; img < n; img += gridDim.x) {
        __syncthreads();
        first_load_tile(tile, x + (long long)img * g.ih * g.iw, g);
        __syncthreads();
        int py = warp / ppr, pq = warp - py * ppr;
        for (int q = warp; q < npairs; q += 8) {
            const float* b = tile + (2 * py) * tw + 4 * pq;          // 16-byte aligned: tw % 4 == 0
            float pt[FH + 1][8];
#pragma unroll
            for (int r = 0; r <= FH; ++r) {
                const float4 v0 = *reinterpret_cast<const float4*>(b + r * tw);
                const float4 v1 = *reinterpret_cast<const float4*>(b + r * tw + 4);
                pt[r][0] = v0.x; pt[r][1] = v0.y; pt[r][2] = v0.z; pt[r][3] = v0.w;
                pt[r][4] = v1.x; pt[r][5] = v1.y; pt[r][6] = v1.z; pt[r][7] = v1.w;
            }
            float a[2][4];
#pragma unroll
            for (int h = 0; h < 2; ++h) { a[h][0] = 0.0f; a[h][1] = 0.0f; a[h][2] = 0.0f; a[h][3] = 0.0f; }
#pragma unroll
            for (int r = 0; r < FH; ++r)
#pragma unroll
                for (int c = 0; c < FW; ++c) {
                    const float wv = wreg[r * FW + c];
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        a[h][0] = fmaf(pt[r][2 * h + c], wv, a[h][0]);
                        a[h][1] = fmaf(pt[r][2 * h + c + 1], wv, a[h][1]);
                        a[h][2] = fmaf(pt[r + 1][2 * h + c], wv, a[h][2]);
                        a[h][3] = fmaf(pt[r + 1][2 * h + c + 1], wv, a[h][3]);
                    }
                }
            const bool vy = 2 * py + 1 < g.coh;
            const long long o0 = ((long long)img * npool + py * g.pw_ + 2 * pq) * g.oc + co;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const bool vx = 2 * (2 * pq + h) + 1 < g.cow;
                float best = fmaxf(a[h][0] + bv, 0.0f); int bi = 0;
                if (vx) { const float v = fmaxf(a[h][1] + bv, 0.0f); if (v > best) { best = v; bi = 1; } }
                if (vy) { const float v = fmaxf(a[h][2] + bv, 0.0f); if (v > best) { best = v; bi = 2; } }
                if (vy && vx) { const float v = fmaxf(a[h][3] + bv, 0.0f); if (v > best) { best = v; bi = 3; } }
                if (co < g.oc) {
                    y[o0 + (long long)h * g.oc] = best;
                    arg[o0 + (long long)h * g.oc] = (uint8_t)bi;
                }
            }
            pq += 8;
            while (pq >= ppr) { pq -= ppr; ++py; }
        }
    }
}

// A 2 x 2 block of pooled pixels (4 x 4 convolution outputs) per warp pass, the 8 x 8 input patch streamed row by row: two
// 16-byte shared loads per input row feed up to 80 fmas, only one patch row is live in registers, and the per-pixel overhead
// is paid once per four pooled pixels.  Launched with 224 threads (7 warps: 49 blocks per 28 x 28 picture).  Input row i meets
// filter row r = i - dy of output row dy, so every output still adds its taps in (r, c) order: identical bytes.
__global__ void __launch_bounds__(224, 4)
first_conv_relu_pool_fwd_c1x4_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                                     float* __restrict__ y, uint8_t* __restrict__ arg, FirstGeom g, int n, int ob) {
    extern __shared__ __align__(16) float tile[];
    constexpr int FH = 5, FW = 5, T = FH * FW;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
    const int co = ob * 32 + lane;
    float wreg[T];
#pragma unroll
    for (int t = 0; t < T; ++t) wreg[t] = co < g.oc ? __ldg(w + (long long)t * g.oc + co) : 0.0f;
    const float bv = (bias && co < g.oc) ? __ldg(bias + co) : 0.0f;
    const int npool = g.poh * g.pw_;
    const int qpr = g.pw_ >> 1, nquad = (g.poh >> 1) * qpr;
    const int tw = g.tw;
    for (int img = blockIdx.x; img < n; img += gridDim.x) {
        __syncthreads();
        first_load_tile(tile, x + (long long)img * g.ih * g.iw, g);
        __syncthreads();
        int qy = warp / qpr, qx = warp - qy * qpr;
        for (int q = warp; q < nquad; q += nwarp) {
            const float* b = tile + (4 * qy) * tw + 4 * qx;          // 16-byte aligned: tw % 4 == 0
            float a[4][4];
#pragma unroll
            for (int dy = 0; dy < 4; ++dy)
#pragma unroll
                for (int dx = 0; dx < 4; ++dx) a[dy][dx] = 0.0f;
#pragma unroll
            for (int i = 0; i < FH + 3; ++i) {
                const float4 v0 = *reinterpret_cast<const float4*>(b + i * tw);
                const float4 v1 = *reinterpret_cast<const float4*>(b + i * tw + 4);
                const float p[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
                for (int dy = 0; dy < 4; ++dy) {
                    const int r = i - dy;
                    if (r >= 0 && r < FH) {
#pragma unroll
                        for (int c = 0; c < FW; ++c) {
                            const float wv = wreg[r * FW + c];
#pragma unroll
                            for (int dx = 0; dx < 4; ++dx) a[dy][dx] = fmaf(p[c + dx], wv, a[dy][dx]);
                        }
                    }
                }
            }
#pragma unroll
            for (int hy = 0; hy < 2; ++hy) {
                const int py = 2 * qy + hy;
                const bool vy = 2 * py + 1 < g.coh;
#pragma unroll
                for (int hx = 0; hx < 2; ++hx) {
                    const int px = 2 * qx + hx;
                    const bool vx = 2 * px + 1 < g.cow;
                    float best = fmaxf(a[2 * hy][2 * hx] + bv, 0.0f); int bi = 0;
                    if (vx) { const float v = fmaxf(a[2 * hy][2 * hx + 1] + bv, 0.0f); if (v > best) { best = v; bi = 1; } }
                    if (vy) { const float v = fmaxf(a[2 * hy + 1][2 * hx] + bv, 0.0f); if (v > best) { best = v; bi = 2; } }
                    if (vy && vx) { const float v = fmaxf(a[2 * hy + 1][2 * hx + 1] + bv, 0.0f); if (v > best) { best = v; bi = 3; } }
                    if (co < g.oc) {
                        const long long o = ((long long)img * npool + py * g.pw_ + px) * g.oc + co;
                        y[o] = best;
                        arg[o] = (uint8_t)bi;
                    }
                }
            }
            qx += nwarp;
            while (qx >= qpr) { qx -= qpr; ++qy; }
        }
    }
}

// The matching weight gradient: the picture's masked pooled gradient and argmax bytes are staged in shared memory by the
// whole CTA first (one exposed global latency per picture instead of one per pooled pixel), taps at compile-time offsets.
// dynamic smem: tile | gm[npool][32] floats | am[npool][32] bytes
template <int FH, int FW>
__global__ void __launch_bounds__(256)
first_conv_wgrad_from_pool_c1_kernel(const float* __restrict__ x, const float* __restrict__ dpool, const uint8_t* __restrict__ arg,
                                     const float* __restrict__ pooled, float* __restrict__ partial_w, float* __restrict__ partial_b,
                                     FirstGeom g, int n, int ob) {
    extern __shared__ __align__(16) float tile[];
    __shared__ float red[8][33];
    constexpr int T = FH * FW;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int co = ob * 32 + lane;
    const int npool = g.poh * g.pw_;
    const int tile_floats = (g.th * g.tw + 3) & ~3;
    float* gm = tile + tile_floats;
    uint8_t* am = reinterpret_cast<uint8_t*>(gm + (size_t)npool * 32);
    float acc[T];
#pragma unroll
    for (int t = 0; t < T; ++t) acc[t] = 0.0f;
    float accb = 0.0f;
    for (int img = blockIdx.x; img < n; img += gridDim.x) {
        __syncthreads();
        first_load_tile(tile, x + (long long)img * g.ih * g.iw, g);
        if ((g.oc & 3) == 0 && ob * 32 + 32 <= g.oc &&
            ((reinterpret_cast<uintptr_t>(dpool) | reinterpret_cast<uintptr_t>(pooled)) & 15) == 0 && (reinterpret_cast<uintptr_t>(arg) & 3) == 0) {
            // four channels per thread and four such loads in flight: the staging is one exposed memory latency per batch of
            // 16 values instead of one per value
            const int total4 = npool * 8;
            const long long base = (long long)img * npool * g.oc + ob * 32;
            for (int e0 = threadIdx.x; e0 < total4; e0 += 4 * 256) {
                float4 gv[4], pv[4]; uchar4 av[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int e = e0 + u * 256;
                    if (e < total4) {
                        const long long o = base + (long long)(e >> 3) * g.oc + (e & 7) * 4;
                        gv[u] = __ldg(reinterpret_cast<const float4*>(dpool + o));
                        pv[u] = __ldg(reinterpret_cast<const float4*>(pooled + o));
                        av[u] = *reinterpret_cast<const uchar4*>(arg + o);
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int e = e0 + u * 256;
                    if (e < total4) {
                        float4 m = gv[u];                                  // relu under the pool
                        if (!(pv[u].x > 0.0f)) m.x = 0.0f;
                        if (!(pv[u].y > 0.0f)) m.y = 0.0f;
                        if (!(pv[u].z > 0.0f)) m.z = 0.0f;
                        if (!(pv[u].w > 0.0f)) m.w = 0.0f;
                        reinterpret_cast<float4*>(gm)[e] = m;
                        reinterpret_cast<uchar4*>(am)[e] = av[u];
                    }
                }
            }
        } else {
            for (int e = threadIdx.x; e < npool * 32; e += 256) {
                const int pp = e >> 5, c = ob * 32 + (e & 31);
                float gv = 0.0f; uint8_t a = 0;
                if (c < g.oc) {
                    const long long o = ((long long)img * npool + pp) * g.oc + c;
                    gv = __ldg(dpool + o);
                    if (!(__ldg(pooled + o) > 0.0f)) gv = 0.0f;          // relu under the pool
                    a = arg[o];
                }
                gm[e] = gv; am[e] = a;
            }
        }
        __syncthreads();
        for (int pp = warp; pp < npool; pp += 8) {
            const int py = pp / g.pw_, px = pp - py * g.pw_;
            const float gv = gm[pp * 32 + lane];
            const int a = am[pp * 32 + lane];
            accb += gv;
            const float* q0 = tile + (2 * py + (a >> 1)) * g.tw + 2 * px + (a & 1);
#pragma unroll
            for (int r = 0; r < FH; ++r)
#pragma unroll
                for (int c = 0; c < FW; ++c) acc[r * FW + c] = fmaf(q0[r * g.tw + c], gv, acc[r * FW + c]);
        }
    }
#pragma unroll
    for (int t = 0; t <= T; ++t) {
        __syncthreads();
        red[warp][lane] = t < T ? acc[t < T ? t : 0] : accb;
        __syncthreads();
        if (warp == 0 && co < g.oc) {
            float sum = 0.0f;
#pragma unroll
            for (int k = 0; k < 8; ++k) sum += red[k][lane];
            if (t < T) partial_w[(size_t)blockIdx.x * T * g.oc + (size_t)t * g.oc + co] = sum;
            else partial_b[(size_t)blockIdx.x * g.oc + co] = sum;
        }
    }
}

// Pipelined variant (32 channels per launch, 16-byte aligned rows): while the warps accumulate picture i from shared memory,
// cp.async brings picture i + 1 (input tile, pooled gradient, pooled output, argmax bytes) into the other buffers, so the
// global-memory latency that made up half of the kernel above is paid once per CTA, not once per picture.  The relu mask is
// applied in place between two pictures.  Same fma order per tap and channel: identical bytes.
// dynamic smem: tile[2][tf] | rg[2][npool * 32] floats | rp[npool * 32] floats | ra[2][npool * 32] bytes
__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_4_zfill(void* smem_dst, const void* gsrc, bool valid) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc), "r"(valid ? 4 : 0) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

template <int FH, int FW>
__global__ void __launch_bounds__(256, 2)
first_conv_wgrad_from_pool_c1p_kernel(const float* __restrict__ x, const float* __restrict__ dpool, const uint8_t* __restrict__ arg,
                                      const float* __restrict__ pooled, float* __restrict__ partial_w, float* __restrict__ partial_b,
                                      FirstGeom g, int n, int ob) {
    extern __shared__ __align__(16) float tile[];
    __shared__ float red[8][33];
    constexpr int T = FH * FW;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int co = ob * 32 + lane;
    const int npool = g.poh * g.pw_;
    const int tf = (g.th * g.tw + 3) & ~3;
    float* rg = tile + 2 * tf;
    float* rp = rg + 2 * npool * 32;
    uint8_t* ra = reinterpret_cast<uint8_t*>(rp + npool * 32);
    const int tw = g.tw;

    auto fetch = [&](int img, int buf) {
        const float* xi = x + (long long)img * g.ih * g.iw;
        float* tb = tile + buf * tf;
        const int total = g.th * tw;
        for (int e = threadIdx.x; e < total; e += 256) {
            const int r = e / tw, c = e - r * tw;
            const int iy = r - g.pt, ix = c - g.pl;
            const bool ok = iy >= 0 && iy < g.ih && ix >= 0 && ix < g.iw;
            cp_async_4_zfill(tb + e, ok ? xi + (long long)iy * g.iw + ix : xi, ok);
        }
        const long long base = (long long)img * npool * g.oc + ob * 32;
        float* gb = rg + buf * npool * 32;
        for (int e = threadIdx.x; e < npool * 8; e += 256) {
            const long long o = base + (long long)(e >> 3) * g.oc + (e & 7) * 4;
            cp_async_16(gb + e * 4, dpool + o);
            cp_async_16(rp + e * 4, pooled + o);
        }
        uint8_t* ab = ra + buf * npool * 32;
        for (int e = threadIdx.x; e < npool * 2; e += 256) cp_async_16(ab + e * 16, arg + base + (long long)(e >> 1) * g.oc + (e & 1) * 16);
    };
    auto mask = [&](int buf) {                              // relu under the pool, in place
        float4* gb = reinterpret_cast<float4*>(rg + buf * npool * 32);
        const float4* pb = reinterpret_cast<const float4*>(rp);
        for (int e = threadIdx.x; e < npool * 8; e += 256) {
            float4 m = gb[e]; const float4 pv = pb[e];
            if (!(pv.x > 0.0f)) m.x = 0.0f;
            if (!(pv.y > 0.0f)) m.y = 0.0f;
            if (!(pv.z > 0.0f)) m.z = 0.0f;
            if (!(pv.w > 0.0f)) m.w = 0.0f;
            gb[e] = m;
        }
    };

    float acc[T];
#pragma unroll
    for (int t = 0; t < T; ++t) acc[t] = 0.0f;
    float accb = 0.0f;
    int buf = 0;
    if ((int)blockIdx.x < n) {
        fetch(blockIdx.x, 0);
        cp_async_wait_all();
        __syncthreads();
        mask(0);
        __syncthreads();
    }
    for (int img = blockIdx.x; img < n; img += gridDim.x, buf ^= 1) {
        const int next = img + gridDim.x;
        if (next < n) fetch(next, buf ^ 1);
        const float* tb = tile + buf * tf;
        const float* gm = rg + buf * npool * 32;
        const uint8_t* am = ra + buf * npool * 32;
        int py = warp / g.pw_, px = warp - py * g.pw_;
        for (int pp = warp; pp < npool; pp += 8) {
            const float gv = gm[pp * 32 + lane];
            const int a = am[pp * 32 + lane];
            accb += gv;
            const float* q0 = tb + (2 * py + (a >> 1)) * tw + 2 * px + (a & 1);
#pragma unroll
            for (int r = 0; r < FH; ++r)
#pragma unroll
                for (int c = 0; c < FW; ++c) acc[r * FW + c] = fmaf(q0[r * tw + c], gv, acc[r * FW + c]);
            px += 8;
            while (px >= g.pw_) { px -= g.pw_; ++py; }
        }
        if (next < n) {
            cp_async_wait_all();
            __syncthreads();                 // everyone is done with rg[buf] / tile[buf], and picture `next` has landed
            mask(buf ^ 1);
            __syncthreads();
        }
    }
#pragma unroll
    for (int t = 0; t <= T; ++t) {
        __syncthreads();
        red[warp][lane] = t < T ? acc[t < T ? t : 0] : accb;
        __syncthreads();
        if (warp == 0 && co < g.oc) {
            float sum = 0.0f;
#pragma unroll
            for (int k = 0; k < 8; ++k) sum += red[k][lane];
            if (t < T) partial_w[(size_t)blockIdx.x * T * g.oc + (size_t)t * g.oc + co] = sum;
            else partial_b[(size_t)blockIdx.x * g.oc + co] = sum;
        }
    }
}

// ---- first layer NOT followed by a pool (conv -> relu -> conv ..., the C5 style): same picture-tile scheme, compile-time
// filter.  A warp computes 4 horizontally adjacent outputs from FH row segments of (FW + 3) * IC floats read with 16-byte
// shared loads; lane = output channel.  The 64-position im2col tiles of conv_smallc_*_tile_kernel spend two thirds of
// their instructions building the im2col; here that is one coalesced tile load per picture.  (Forward only: the same
// scheme for the weight gradient measured slower than conv_smallc_wgrad_tile_kernel -- 188 vs 120 us on C5 -- and was dropped.)
struct PatchGeom {
    int ih, iw, pt, pl;      // input, padding before
    int oh, ow, oc;          // output (stride 1)
    int th, tws;             // tile rows, tile row stride in floats (multiple of 4, >= (ow + FW + 3) * IC)
};

template <int IC>
__device__ __forceinline__ void patch_load_tile(float* tile, const float* __restrict__ xi, const PatchGeom& g) {
    const int total = g.th * g.tws;
    for (int e = threadIdx.x; e < total; e += blockDim.x) {
        const int r = e / g.tws, q = e - r * g.tws;
        const int c = q / IC, ch = q - c * IC;
        const int iy = r - g.pt, ix = c - g.pl;
        tile[e] = (iy >= 0 && iy < g.ih && ix >= 0 && ix < g.iw) ? __ldg(xi + ((long long)iy * g.iw + ix) * IC + ch) : 0.0f;
    }
}

template <int FH, int FW, int IC>
__global__ void __launch_bounds__(256)
first_conv_patch_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                            float* __restrict__ y, PatchGeom g, int n, int ob, int relu) {
    extern __shared__ __align__(16) float tile[];
    constexpr int T = FH * FW * IC, SEG = ((FW + 3) * IC + 3) / 4 * 4;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int co = ob * 32 + lane;
    float wreg[T];
#pragma unroll
    for (int t = 0; t < T; ++t) wreg[t] = co < g.oc ? __ldg(w + (long long)t * g.oc + co) : 0.0f;
    const float bv = (bias && co < g.oc) ? __ldg(bias + co) : 0.0f;
    const int gpr = (g.ow + 3) >> 2, ngroups = g.oh * gpr;
    for (int img = blockIdx.x; img < n; img += gridDim.x) {
        __syncthreads();
        patch_load_tile<IC>(tile, x + (long long)img * g.ih * g.iw * IC, g);
        __syncthreads();
        for (int gi = warp; gi < ngroups; gi += 8) {
            const int oy = gi / gpr, ox0 = (gi - oy * gpr) * 4;
            float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
            for (int r = 0; r < FH; ++r) {
                float seg[SEG];
                const float4* sp = reinterpret_cast<const float4*>(tile + (oy + r) * g.tws + ox0 * IC);
#pragma unroll
                for (int v = 0; v < SEG / 4; ++v) { const float4 q = sp[v]; seg[4 * v] = q.x; seg[4 * v + 1] = q.y; seg[4 * v + 2] = q.z; seg[4 * v + 3] = q.w; }
#pragma unroll
                for (int c = 0; c < FW; ++c)
#pragma unroll
                    for (int ch = 0; ch < IC; ++ch) {
                        const float wv = wreg[(r * FW + c) * IC + ch];
#pragma unroll
                        for (int p = 0; p < 4; ++p) acc[p] = fmaf(seg[(p + c) * IC + ch], wv, acc[p]);
                    }
            }
            if (co < g.oc) {
#pragma unroll
                for (int p = 0; p < 4; ++p)
                    if (ox0 + p < g.ow) {
                        const float v = acc[p] + bv;
                        y[(((long long)img * g.oh + oy) * g.ow + ox0 + p) * g.oc + co] = relu ? fmaxf(v, 0.0f) : v;
                    }
            }
        }
    }
}

// partial_w[block][T*oc] / partial_b[block][oc] for output-channel block `ob` (summed by reduce_partials_warp_kernel)
__global__ void __launch_bounds__(256)
first_conv_wgrad_from_pool_kernel(const float* __restrict__ x, const float* __restrict__ dpool, const uint8_t* __restrict__ arg,
                                  const float* __restrict__ pooled, float* __restrict__ partial_w, float* __restrict__ partial_b,
                                  FirstGeom g, int n, int ob) {
    extern __shared__ float tile[];
    __shared__ float red[8][33];
    const int T = g.fh * g.fw * g.ic;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int co = ob * 32 + lane;
    float acc[32];
    int off[32];
#pragma unroll
    for (int t = 0; t < 32; ++t) {
        acc[t] = 0.0f;
        const int tt = t < T ? t : 0;
        const int c = tt % g.ic, ij = tt / g.ic;
        off[t] = ((ij / g.fw) * g.tw + (ij % g.fw)) * g.ic + c;
    }
    float accb = 0.0f;
    const int rowp = g.tw * g.ic;
    const int npool = g.poh * g.pw_;
    for (int img = blockIdx.x; img < n; img += gridDim.x) {
        __syncthreads();
        first_load_tile(tile, x + (long long)img * g.ih * g.iw * g.ic, g);
        __syncthreads();
        for (int pp = warp; pp < npool; pp += 8) {
            const int py = pp / g.pw_, px = pp - py * g.pw_;
            float gv = 0.0f; int a = 0;
            if (co < g.oc) {
                const long long o = ((long long)img * npool + pp) * g.oc + co;
                gv = __ldg(dpool + o);
                if (!(__ldg(pooled + o) > 0.0f)) gv = 0.0f;        // relu under the pool
                a = arg[o];
            }
            accb += gv;
            const float* q0 = tile + (2 * py + (a >> 1)) * rowp + (2 * px + (a & 1)) * g.ic;
#pragma unroll
            for (int t = 0; t < 32; ++t)
                if (t < T) acc[t] = fmaf(q0[off[t]], gv, acc[t]);
        }
    }
#pragma unroll
    for (int t = 0; t <= 32; ++t) {
        if (t < T || t == 32) {
            __syncthreads();
            red[warp][lane] = t < 32 ? acc[t < 32 ? t : 0] : accb;
            __syncthreads();
            if (warp == 0 && co < g.oc) {
                float sum = 0.0f;
#pragma unroll
                for (int k = 0; k < 8; ++k) sum += red[k][lane];
                if (t < 32) partial_w[(size_t)blockIdx.x * T * g.oc + (size_t)t * g.oc + co] = sum;
                else partial_b[(size_t)blockIdx.x * g.oc + co] = sum;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// activations (active(), construct_distribute.py:133-152)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float act_apply(int fn, float x, float slope) {
    if (fn == ACT_RELU) return fmaxf(x, 0.0f);
    if (fn == ACT_SIGMOID) return 1.0f / (1.0f + expf(-x));
    if (fn == ACT_LEAKY) return fmaxf(x, slope * x);
    return x;
}

__global__ void act_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long long count, int fn, float slope) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        y[i] = act_apply(fn, __ldg(x + i), slope);
}

// dx = dy * f'(x): relu from y > 0 (tf ReluGrad), sigmoid y(1-y), leaky: x >= slope*x ? 1 : slope (tf.maximum grad)
__global__ void act_bwd_kernel(const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ dy,
                               float* __restrict__ dx, long long count, int fn, float slope) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        const float g = __ldg(dy + i);
        float r;
        if (fn == ACT_RELU) r = __ldg(x + i) > 0.0f ? g : 0.0f;      // x > 0 <=> relu(x) > 0; x is always kept
        else if (fn == ACT_SIGMOID) { const float v = __ldg(y + i); r = g * v * (1.0f - v); }
        else { const float v = __ldg(x + i); r = (v >= slope * v) ? g : slope * g; }
        dx[i] = r;
    }
}

// ---------------------------------------------------------------------------------------------
// batch normalisation with batch statistics (normalize(), construct_distribute.py:155-165)
// rows x C matrix (rows = N*H*W), one CTA per channel
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float block_sum(float v, float* sh) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    float t = 0.0f;
    const int nw = blockDim.x >> 5;
    for (int i = 0; i < nw; ++i) t += sh[i];
    return t;
}

__global__ void __launch_bounds__(256)
norm_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, float* __restrict__ xhat, float* __restrict__ inv_out,
                const float* __restrict__ scale, const float* __restrict__ offset, long long rows, int C, float eps) {
    __shared__ float sh[8];
    const int c = blockIdx.x;
    float s = 0.0f;
    for (long long r = threadIdx.x; r < rows; r += blockDim.x) s += __ldg(x + r * C + c);
    const float mean = block_sum(s, sh) / (float)rows;
    float q = 0.0f;
    for (long long r = threadIdx.x; r < rows; r += blockDim.x) { const float d = __ldg(x + r * C + c) - mean; q += d * d; }
    const float var = block_sum(q, sh) / (float)rows;
    const float inv = 1.0f / sqrtf(var + eps);
    if (threadIdx.x == 0) inv_out[c] = inv;
    const float sc = __ldg(scale + c), of = __ldg(offset + c);
    for (long long r = threadIdx.x; r < rows; r += blockDim.x) {
        const float h = (__ldg(x + r * C + c) - mean) * inv;
        xhat[r * C + c] = h;
        y[r * C + c] = h * sc + of;
    }
}

__global__ void __launch_bounds__(256)
norm_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ xhat, const float* __restrict__ inv_in,
                const float* __restrict__ scale, float* __restrict__ dx, float* __restrict__ dscale, float* __restrict__ doffset,
                long long rows, int C) {
    __shared__ float sh[8];
    const int c = blockIdx.x;
    float s1 = 0.0f, s2 = 0.0f;
    for (long long r = threadIdx.x; r < rows; r += blockDim.x) {
        const float g = __ldg(dy + r * C + c);
        s1 += g; s2 += g * __ldg(xhat + r * C + c);
    }
    const float sum_dy = block_sum(s1, sh);
    const float sum_dyx = block_sum(s2, sh);
    if (threadIdx.x == 0) { doffset[c] = sum_dy; dscale[c] = sum_dyx; }
    const float sc = __ldg(scale + c), inv = __ldg(inv_in + c), m = (float)rows;
    for (long long r = threadIdx.x; r < rows; r += blockDim.x) {
        const float g = __ldg(dy + r * C + c) * sc;
        dx[r * C + c] = inv / m * (m * g - sum_dy * sc - __ldg(xhat + r * C + c) * sum_dyx * sc);
    }
}

// ---- data-parallel variants: the batch statistics span all ranks (construct_distribute.py:155-165 at the GLOBAL
// batch), so each reduction is split into a per-rank partial sum, an all-reduce, and the apply pass
__global__ void __launch_bounds__(256)
norm_sum_kernel(const float* __restrict__ x, long long rows, int C, float* __restrict__ out) {
    __shared__ float sh[8];
    const int c = blockIdx.x;
    float s = 0.0f;
    for (long long r = threadIdx.x; r < rows; r += blockDim.x) s += __ldg(x + r * C + c);
    const float t = block_sum(s, sh);
    if (threadIdx.x == 0) out[c] = t;
}
__global__ void __launch_bounds__(256)
norm_sqdev_kernel(const float* __restrict__ x, long long rows, int C, const float* __restrict__ sum, float inv_total,
                  float* __restrict__ out) {
    __shared__ float sh[8];
    const int c = blockIdx.x;
    const float mean = __ldg(sum + c) * inv_total;
    float q = 0.0f;
    for (long long r = threadIdx.x; r < rows; r += blockDim.x) { const float d = __ldg(x + r * C + c) - mean; q += d * d; }
    const float t = block_sum(q, sh);
    if (threadIdx.x == 0) out[c] = t;
}
__global__ void __launch_bounds__(256)
norm_apply_kernel(const float* __restrict__ x, float* __restrict__ y, float* __restrict__ xhat, float* __restrict__ inv_out,
                  const float* __restrict__ scale, const float* __restrict__ offset, long long rows, int C, float eps,
                  const float* __restrict__ sum, const float* __restrict__ sqdev, float inv_total) {
    const int c = blockIdx.x;
    const float mean = __ldg(sum + c) * inv_total, var = __ldg(sqdev + c) * inv_total;
    const float inv = 1.0f / sqrtf(var + eps);
    if (threadIdx.x == 0) inv_out[c] = inv;
    const float sc = __ldg(scale + c), of = __ldg(offset + c);
    for (long long r = threadIdx.x; r < rows; r += blockDim.x) {
        const float h = (__ldg(x + r * C + c) - mean) * inv;
        xhat[r * C + c] = h;
        y[r * C + c] = h * sc + of;
    }
}
// sums[0..C) = sum dy, sums[C..2C) = sum dy*xhat over this rank's rows; also the LOCAL dscale / doffset
__global__ void __launch_bounds__(256)
norm_bwd_sums_kernel(const float* __restrict__ dy, const float* __restrict__ xhat, long long rows, int C, float* __restrict__ sums,
                     float* __restrict__ dscale, float* __restrict__ doffset) {
    __shared__ float sh[8];
    const int c = blockIdx.x;
    float s1 = 0.0f, s2 = 0.0f;
    for (long long r = threadIdx.x; r < rows; r += blockDim.x) {
        const float g = __ldg(dy + r * C + c);
        s1 += g; s2 += g * __ldg(xhat + r * C + c);
    }
    const float a = block_sum(s1, sh);
    const float b = block_sum(s2, sh);
    if (threadIdx.x == 0) { sums[c] = a; sums[C + c] = b; doffset[c] = a; dscale[c] = b; }
}
__global__ void __launch_bounds__(256)
norm_bwd_apply_kernel(const float* __restrict__ dy, const float* __restrict__ xhat, const float* __restrict__ inv_in,
                      const float* __restrict__ scale, float* __restrict__ dx, long long rows, int C, const float* __restrict__ sums,
                      float total_rows) {
    const int c = blockIdx.x;
    const float sc = __ldg(scale + c), inv = __ldg(inv_in + c), m = total_rows;
    const float sum_dy = __ldg(sums + c), sum_dyx = __ldg(sums + C + c);
    for (long long r = threadIdx.x; r < rows; r += blockDim.x) {
        const float g = __ldg(dy + r * C + c) * sc;
        dx[r * C + c] = inv / m * (m * g - sum_dy * sc - __ldg(xhat + r * C + c) * sum_dyx * sc);
    }
}

// ---------------------------------------------------------------------------------------------
// loss (loss_function, :285-298) fused with its gradient and the accuracy count (:381-382).
// One CTA; rows strided over threads; deterministic in-block reduction.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
loss_kernel(const float* __restrict__ z, const float* __restrict__ y, float* __restrict__ dz, int n, int classes,
            int loss_name, float grad_scale, float* __restrict__ loss_out, int* __restrict__ correct_out) {
    __shared__ float sh[32];
    __shared__ int shi[32];
    float lsum = 0.0f; int hits = 0;
    for (int r = threadIdx.x; r < n; r += blockDim.x) {
        const float* zr = z + (long long)r * classes;
        const float* yr = y + (long long)r * classes;
        float zmax = -INFINITY, ymax = -INFINITY; int za = 0, ya = 0;
        for (int j = 0; j < classes; ++j) {
            const float a = zr[j], b = yr[j];
            if (a > zmax) { zmax = a; za = j; }
            if (b > ymax) { ymax = b; ya = j; }
        }
        hits += (za == ya);
        if (loss_name == 0) {
            float se = 0.0f;
            for (int j = 0; j < classes; ++j) se += expf(zr[j] - zmax);
            const float lse = zmax + logf(se);
            float l = 0.0f;
            for (int j = 0; j < classes; ++j) {
                const float logp = zr[j] - lse;
                l -= yr[j] * logp;
                if (dz) dz[(long long)r * classes + j] = (expf(logp) - yr[j]) * grad_scale / (float)n;
            }
            lsum += l;
        } else {
            float l = 0.0f;
            for (int j = 0; j < classes; ++j) {
                const float d = zr[j] - yr[j];
                l += d * d;
                if (dz) dz[(long long)r * classes + j] = 2.0f * d * grad_scale / ((float)n * (float)classes);
            }
            lsum += l / (float)classes;
        }
    }
    for (int o = 16; o > 0; o >>= 1) { lsum += __shfl_xor_sync(0xffffffffu, lsum, o); hits += __shfl_xor_sync(0xffffffffu, hits, o); }
    if ((threadIdx.x & 31) == 0) { sh[threadIdx.x >> 5] = lsum; shi[threadIdx.x >> 5] = hits; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f; int hh = 0;
        for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { t += sh[i]; hh += shi[i]; }
        if (loss_out) *loss_out = t / (float)n;
        if (correct_out) *correct_out = hh;
    }
}

__global__ void argmax_kernel(const float* __restrict__ z, int n, int classes, int32_t* __restrict__ out) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const float* zr = z + (long long)r * classes;
    float best = zr[0]; int bi = 0;
    for (int j = 1; j < classes; ++j) if (zr[j] > best) { best = zr[j]; bi = j; }
    out[r] = bi;
}

// ---------------------------------------------------------------------------------------------
// optimizers on the flat buffers (ApplyAdagrad / ApplyGradientDescent / ApplyAdam, TF 1.x)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void adagrad_one(float& w, float g, float& acc, float lr) {
    const float a = acc + g * g;
    acc = a;
    w -= lr * g * rsqrtf(a);
}
// the flat buffers come from cudaMalloc (256-byte aligned): 16-byte accesses for the body, scalars for the last count % 4
__global__ void adagrad_kernel(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ acc, long long count, float lr) {
    const long long stride = (long long)gridDim.x * blockDim.x, tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n4 = count >> 2;
    float4* w4 = reinterpret_cast<float4*>(w);
    float4* a4 = reinterpret_cast<float4*>(acc);
    const float4* g4 = reinterpret_cast<const float4*>(g);
    for (long long i = tid; i < n4; i += stride) {
        float4 wv = w4[i], av = a4[i];
        const float4 gv = g4[i];
        adagrad_one(wv.x, gv.x, av.x, lr); adagrad_one(wv.y, gv.y, av.y, lr);
        adagrad_one(wv.z, gv.z, av.z, lr); adagrad_one(wv.w, gv.w, av.w, lr);
        a4[i] = av; w4[i] = wv;
    }
    for (long long i = (n4 << 2) + tid; i < count; i += stride) adagrad_one(w[i], g[i], acc[i], lr);
}
__global__ void sgd_kernel(float* __restrict__ w, const float* __restrict__ g, long long count, float lr) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) w[i] -= lr * g[i];
}
__global__ void adam_kernel(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                            long long count, float lr_t, float b1, float b2, float eps) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        const float gi = g[i];
        const float mi = m[i] + (gi - m[i]) * (1.0f - b1);
        const float vi = v[i] + (gi * gi - v[i]) * (1.0f - b2);
        m[i] = mi; v[i] = vi;
        w[i] -= lr_t * mi / (sqrtf(vi) + eps);
    }
}
__global__ void fill_kernel(float* __restrict__ p, long long count, float v) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) p[i] = v;
}

}  // namespace nck

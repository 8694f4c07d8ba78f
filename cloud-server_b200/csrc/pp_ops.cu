// The 14 per-picture ops of apps/preprocess/preprocess.py:33-227 on batches of 28x28 8-bit gray
// pictures resident in HBM.  One CTA per picture; the picture lives in shared memory; every op is
// integer / float32 arithmetic restated from the OpenCV / numpy routine the reference calls
// (oracle/pp_restate.py holds the CPU statement of the same arithmetic, pinned against cv2 4.13).
// These are tiny (784 B in, 784 B out per picture): launch- and latency-bound, batched so that one
// launch covers the whole dataset instead of the reference's per-file loop (views.py:118-126).
#include "nc_internal.h"
#include <cmath>
#include <limits.h>

namespace {

constexpr int S = 28;
constexpr int NPIX = S * S;
constexpr int kThreads = 256;

__device__ __forceinline__ int reflect101(int i, int n) {
    if (i < 0) i = -i;
    if (i >= n) i = 2 * (n - 1) - i;
    return i;
}
__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

struct OpParams {
    int op;
    int k;                 // kernel size for window ops
    int is_int;            // bit0: alpha is int, bit1: beta is int
    double alpha, beta;    // brightness/contrast
    int div_scale, div_delta, mean_mode;   // mean filter: 0 = ushort reciprocal, 1 = double scale
    double mean_scale;
    int gk[32];            // gaussian Q8 kernel
    const int* nl_lut;     // nl-means weight LUT (device)
    int nl_lut_size;
    const unsigned short* sp_idx;   // salt/pepper pixel indices (device), per image 2*m entries
    int sp_m;
};

__global__ void __launch_bounds__(kThreads)
ops_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, const __grid_constant__ OpParams P) {
    __shared__ unsigned char img[NPIX];
    __shared__ unsigned char res[NPIX];
    __shared__ int s_i[1024];
    extern __shared__ __align__(16) unsigned char dyn[];
    const int tid = threadIdx.x;
    const long long off = (long long)blockIdx.x * NPIX;
    for (int i = tid; i < NPIX; i += kThreads) img[i] = __ldg(in + off + i);
    __syncthreads();

    switch (P.op) {
    case NC_OP_FLIP_UP_DOWN:
        for (int i = tid; i < NPIX; i += kThreads) res[i] = img[(S - 1 - i / S) * S + i % S];
        break;
    case NC_OP_FLIP_LEFT_RIGHT:
        for (int i = tid; i < NPIX; i += kThreads) res[i] = img[(i / S) * S + (S - 1 - i % S)];
        break;
    case NC_OP_TRANSPOSE_IMAGE:
        for (int i = tid; i < NPIX; i += kThreads) res[i] = img[NPIX - 1 - i];
        break;
    case NC_OP_BRIGHTNESS_CONTRAST: {
        // preprocess.py:89-91: uint8 scalar * alpha + beta stored back into the uint8 array
        for (int i = tid; i < NPIX; i += kThreads) {
            const int p = img[i];
            int r;
            if ((P.is_int & 1) && (P.is_int & 2)) {
                r = ((p * (int)P.alpha) + (int)P.beta) & 255;          // uint8 arithmetic wraps
            } else {
                double t = (P.is_int & 1) ? (double)((p * (int)P.alpha) & 255) : __dmul_rn((double)p, P.alpha);
                t = __dadd_rn(t, P.beta);
                long long q = 0;                                        // C cast: truncate, then low byte
                if (fabs(t) < 9.2e18) q = (long long)t;
                r = (int)(q & 255);
            }
            res[i] = (unsigned char)r;
        }
        break;
    }
    case NC_OP_MEAN_FILTER: {
        const int k = P.k, a = k / 2;
        for (int i = tid; i < NPIX; i += kThreads) {
            const int y = i / S, x = i % S;
            int s = 0;
            for (int dy = 0; dy < k; ++dy) {
                const int yy = reflect101(y + dy - a, S);
                for (int dx = 0; dx < k; ++dx) s += img[yy * S + reflect101(x + dx - a, S)];
            }
            int r;
            if (P.mean_mode == 0) r = (int)(((long long)(s + P.div_delta) * P.div_scale) >> 23);
            else r = clampi(__double2int_rn(__dmul_rn((double)s, P.mean_scale)), 0, 255);
            res[i] = (unsigned char)r;
        }
        break;
    }
    case NC_OP_GAUSSIAN_BLUR: {
        // fixed-point separable filter, 8 fraction bits per pass, final (v + 2^15) >> 16
        const int k = P.k, a = k / 2;
        unsigned short* hp = reinterpret_cast<unsigned short*>(s_i);   // [28][28] Q8 (<= 65280)
        for (int i = tid; i < NPIX; i += kThreads) {
            const int y = i / S, x = i % S;
            int s = 0;
            for (int d = 0; d < k; ++d) s += img[y * S + reflect101(x + d - a, S)] * P.gk[d];
            hp[i] = (unsigned short)s;
        }
        __syncthreads();
        for (int i = tid; i < NPIX; i += kThreads) {
            const int y = i / S, x = i % S;
            unsigned s = 0;
            for (int d = 0; d < k; ++d) s += (unsigned)hp[reflect101(y + d - a, S) * S + x] * (unsigned)P.gk[d];
            res[i] = (unsigned char)clampi((int)((s + (1u << 15)) >> 16), 0, 255);
        }
        break;
    }
    case NC_OP_MEDIAN_FILTER: {
        const int k = P.k, a = k / 2, need = (k * k) / 2 + 1;
        for (int i = tid; i < NPIX; i += kThreads) {
            const int y = i / S, x = i % S;
            int lo = 0, hi = 255;          // smallest v with count(window <= v) >= need
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                int cnt = 0;
                for (int dy = 0; dy < k; ++dy) {
                    const int yy = clampi(y + dy - a, 0, S - 1);
                    for (int dx = 0; dx < k; ++dx) cnt += img[yy * S + clampi(x + dx - a, 0, S - 1)] <= mid;
                }
                if (cnt >= need) hi = mid; else lo = mid + 1;
            }
            res[i] = (unsigned char)lo;
        }
        break;
    }
    case NC_OP_ERODE:
    case NC_OP_DILATE: {
        const int k = P.k, a = k / 2;
        const bool er = P.op == NC_OP_ERODE;
        for (int i = tid; i < NPIX; i += kThreads) {
            const int y = i / S, x = i % S;
            int v = er ? 255 : 0;
            for (int dy = 0; dy < k; ++dy) {
                const int yy = y + dy - a;
                if (yy < 0 || yy >= S) continue;
                for (int dx = 0; dx < k; ++dx) {
                    const int xx = x + dx - a;
                    if (xx < 0 || xx >= S) continue;
                    const int p = img[yy * S + xx];
                    v = er ? min(v, p) : max(v, p);
                }
            }
            res[i] = (unsigned char)v;
        }
        break;
    }
    case NC_OP_EQUALIZE_HIST: {
        int* hist = s_i;            // 256 bins
        int* lut = s_i + 256;
        for (int i = tid; i < 256; i += kThreads) hist[i] = 0;
        __syncthreads();
        for (int i = tid; i < NPIX; i += kThreads) atomicAdd(&hist[img[i]], 1);
        __syncthreads();
        if (tid == 0) {
            int i0 = 0;
            while (hist[i0] == 0) ++i0;
            if (hist[i0] == NPIX) {
                for (int i = 0; i < 256; ++i) lut[i] = i0;   // cv2 fills the picture with i0
            } else {
                const float scale = __fdiv_rn(255.0f, (float)(NPIX - hist[i0]));
                int s = 0;
                for (int i = 0; i <= i0; ++i) lut[i] = 0;
                for (int i = i0 + 1; i < 256; ++i) {
                    s += hist[i];
                    lut[i] = clampi(__float2int_rn(__fmul_rn((float)s, scale)), 0, 255);
                }
            }
        }
        __syncthreads();
        for (int i = tid; i < NPIX; i += kThreads) res[i] = (unsigned char)lut[img[i]];
        break;
    }
    case NC_OP_CLAHE: {
        // createCLAHE() defaults: clipLimit 40, tiles 8x8; 28 -> padded to 32 (REFLECT_101), 4x4 tiles
        constexpr int T = 8, TS = 4, AREA = TS * TS;
        unsigned char* luts = dyn;                 // [64][256]
        unsigned char* hists = dyn + 64 * 256;     // [64][256]
        for (int i = tid; i < 64 * 256; i += kThreads) hists[i] = 0;
        __syncthreads();
        if (tid < 64) {
            const int tj = tid / T, ti = tid % T;
            unsigned char* h = hists + tid * 256;
            for (int yy = 0; yy < TS; ++yy)
                for (int xx = 0; xx < TS; ++xx) {
                    const int sy = reflect101(tj * TS + yy, S), sx = reflect101(ti * TS + xx, S);
                    h[img[sy * S + sx]]++;
                }
            int clip = (int)(40.0 * AREA / 256);
            if (clip < 1) clip = 1;
            int clipped = 0;
            for (int b = 0; b < 256; ++b)
                if (h[b] > clip) { clipped += h[b] - clip; h[b] = (unsigned char)clip; }
            const int redist = clipped / 256;
            int residual = clipped - redist * 256;
            for (int b = 0; b < 256; ++b) h[b] = (unsigned char)(h[b] + redist);
            if (residual != 0) {
                int step = 256 / residual;
                if (step < 1) step = 1;
                for (int b = 0; b < 256 && residual > 0; b += step, --residual) h[b]++;
            }
            const float lut_scale = __fdiv_rn(255.0f, (float)AREA);
            int s = 0;
            unsigned char* l = luts + tid * 256;
            for (int b = 0; b < 256; ++b) {
                s += h[b];
                l[b] = (unsigned char)clampi(__float2int_rn(__fmul_rn((float)s, lut_scale)), 0, 255);
            }
        }
        __syncthreads();
        const float inv_t = __fdiv_rn(1.0f, (float)TS);
        for (int i = tid; i < NPIX; i += kThreads) {
            const int y = i / S, x = i % S;
            const float tyf = __fsub_rn(__fmul_rn((float)y, inv_t), 0.5f);
            int ty1 = (int)floorf(tyf), ty2 = ty1 + 1;
            const float ya = __fsub_rn(tyf, (float)ty1), ya1 = __fsub_rn(1.0f, ya);
            ty1 = max(ty1, 0); ty2 = min(ty2, T - 1);
            const float txf = __fsub_rn(__fmul_rn((float)x, inv_t), 0.5f);
            int tx1 = (int)floorf(txf), tx2 = tx1 + 1;
            const float xa = __fsub_rn(txf, (float)tx1), xa1 = __fsub_rn(1.0f, xa);
            tx1 = max(tx1, 0); tx2 = min(tx2, T - 1);
            const int v = img[i];
            const float a11 = (float)luts[(ty1 * T + tx1) * 256 + v], a12 = (float)luts[(ty1 * T + tx2) * 256 + v];
            const float a21 = (float)luts[(ty2 * T + tx1) * 256 + v], a22 = (float)luts[(ty2 * T + tx2) * 256 + v];
            const float top = __fadd_rn(__fmul_rn(a11, xa1), __fmul_rn(a12, xa));
            const float bot = __fadd_rn(__fmul_rn(a21, xa1), __fmul_rn(a22, xa));
            const float r = __fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya));
            res[i] = (unsigned char)clampi(__float2int_rn(r), 0, 255);
        }
        break;
    }
    case NC_OP_NL_DENOISE: {
        // fastNlMeansDenoising(h, 7, 21): DistSquared, int fixed-point weights
        constexpr int TW = 7, SW = 21, THR = TW / 2, SHR = SW / 2, B = THR + SHR, E = S + 2 * B;  // 54
        unsigned char* ext = dyn;   // [54][54]
        for (int i = tid; i < E * E; i += kThreads) {
            const int y = i / E, x = i % E;
            ext[i] = img[reflect101(y - B, S) * S + reflect101(x - B, S)];
        }
        __syncthreads();
        for (int i = tid; i < NPIX; i += kThreads) {
            const int y = i / S + B, x = i % S + B;
            int est = 0, wsum = 0;
            for (int dy = -SHR; dy <= SHR; ++dy)
                for (int dx = -SHR; dx <= SHR; ++dx) {
                    int dist = 0;
                    for (int ty = -THR; ty <= THR; ++ty) {
                        const unsigned char* pa = ext + (y + ty) * E + x - THR;
                        const unsigned char* pb = ext + (y + dy + ty) * E + x + dx - THR;
#pragma unroll
                        for (int tx = 0; tx < TW; ++tx) {
                            const int d = (int)pa[tx] - (int)pb[tx];
                            dist += d * d;
                        }
                    }
                    const int wgt = __ldg(P.nl_lut + (dist >> 6));
                    est += wgt * (int)ext[(y + dy) * E + x + dx];
                    wsum += wgt;
                }
            const unsigned r = ((unsigned)est + (unsigned)(wsum / 2)) / (unsigned)wsum;
            res[i] = (unsigned char)(r > 255u ? 255u : r);
        }
        break;
    }
    case NC_OP_SALT_PEPPER: {
        for (int i = tid; i < NPIX; i += kThreads) res[i] = img[i];
        __syncthreads();
        if (tid == 0) {
            const unsigned short* idx = P.sp_idx + (long long)blockIdx.x * 2 * P.sp_m;
            for (int a = 0; a < P.sp_m; ++a) res[idx[a]] = 255;            // preprocess.py:163-171
            for (int a = 0; a < P.sp_m; ++a) res[idx[P.sp_m + a]] = 0;     // preprocess.py:173-183
        }
        break;
    }
    default:
        for (int i = tid; i < NPIX; i += kThreads) res[i] = img[i];
    }
    __syncthreads();
    for (int i = tid; i < NPIX; i += kThreads) out[off + i] = res[i];
}

}  // namespace

// nl-means weight LUT per h (photo/fast_nlmeans_denoising_invoker.hpp constructor)
static int get_nl_lut(nc_ctx* ctx, float h, const int** d_lut, int* size) {
    int hbits;
    memcpy(&hbits, &h, 4);
    auto key = std::make_pair(-7, hbits);
    auto it = ctx->pil_tables.find(key);
    if (it == ctx->pil_tables.end()) {
        const int tw = 7, sw = 21, tws = tw * tw;
        int shift = 0;
        while ((1 << shift) < tws) ++shift;
        const double a2a = (double)(1 << shift) / tws;
        const int max_dist = 255 * 255;
        const int almost_max = (int)(max_dist / a2a + 1);
        const int fixed_point_mult = INT_MAX / (sw * sw * 255);
        std::vector<int> lut(almost_max);
        const float hh = h * h * 1;   // float, as h[i]*h[i]*channels
        for (int a = 0; a < almost_max; ++a) {
            const double dist = a * a2a;
            double w = std::exp(-dist / hh);
            if (std::isnan(w)) w = 1.0;
            int wi = (int)std::lrint(fixed_point_mult * w);
            if (wi < 0.001 * fixed_point_mult) wi = 0;
            lut[a] = wi;
        }
        nc_dev_table tab;
        tab.bytes = lut.size() * sizeof(int);
        NC_CUDA(ctx, cudaMalloc(&tab.dptr, tab.bytes));
        NC_CUDA(ctx, cudaMemcpy(tab.dptr, lut.data(), tab.bytes, cudaMemcpyHostToDevice));
        it = ctx->pil_tables.emplace(key, tab).first;
    }
    *d_lut = (const int*)it->second.dptr;
    *size = (int)(it->second.bytes / sizeof(int));
    return NC_OK;
}

extern "C" int nc_pp_op(nc_ctx* ctx, int op_id, const uint8_t* in, uint8_t* out, int n, double v1, double v2,
                        int v_is_int, const double* host_draws, int64_t draws_per_image, void* stream) {
    if (!ctx) return nc_set_error(nullptr, NC_EINVAL, "ctx is NULL");
    if (n < 0 || (n > 0 && (!in || !out))) return nc_set_error(ctx, NC_EINVAL, "bad buffers");
    if (op_id < 0 || op_id > NC_OP_DILATE) return nc_set_error(ctx, NC_EINVAL, "unknown op %d", op_id);
    if (n == 0) return NC_OK;
    NC_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    OpParams P;
    memset(&P, 0, sizeof(P));
    P.op = op_id;
    size_t dyn = 0;
    const int k = (int)v1;
    switch (op_id) {
    case NC_OP_BRIGHTNESS_CONTRAST:
        P.alpha = v1; P.beta = v2; P.is_int = v_is_int;
        // numpy 2 (NEP 50): uint8 scalar (op) Python int stays uint8 and refuses out-of-range ints
        if ((v_is_int & 1) && (v1 < 0 || v1 > 255))
            return nc_set_error(ctx, NC_EINVAL, "OverflowError: Python integer %lld out of bounds for uint8", (long long)v1);
        if ((v_is_int & 1) && (v_is_int & 2) && (v2 < 0 || v2 > 255))
            return nc_set_error(ctx, NC_EINVAL, "OverflowError: Python integer %lld out of bounds for uint8", (long long)v2);
        break;
    case NC_OP_MEAN_FILTER: {
        if (k <= 0) return nc_set_error(ctx, NC_ECV, "cv2.error: blur ksize %d", k);
        if (k > 27) return nc_set_error(ctx, NC_EUNSUPPORTED, "blur ksize %d > 27", k);
        P.k = k;
        const int d = k * k;
        if (d <= 256) {
            // ColumnSum<ushort,uchar>: integer reciprocal (box_filter.simd.hpp)
            P.mean_mode = 0;
            if (d == 1) { P.div_scale = 1 << 23; P.div_delta = 0; }
            else {
                const double f = (double)(1 << 23) / d;
                P.div_scale = (int)std::floor(f);
                const double frac = f - P.div_scale;
                P.div_delta = d / 2;
                if (frac < 0.5) P.div_delta++; else P.div_scale++;
            }
        } else {
            P.mean_mode = 1;
            P.mean_scale = 1.0 / d;
        }
        break;
    }
    case NC_OP_GAUSSIAN_BLUR: {
        if (k <= 0 || (k & 1) == 0) return nc_set_error(ctx, NC_ECV, "cv2.error: GaussianBlur ksize %d must be odd and positive", k);
        if (k > 27) return nc_set_error(ctx, NC_EUNSUPPORTED, "GaussianBlur ksize %d > 27", k);
        P.k = k;
        std::vector<int> q = nc_gaussian_kernel_q8(k);
        for (int i = 0; i < k; ++i) P.gk[i] = q[i];
        break;
    }
    case NC_OP_MEDIAN_FILTER:
        if (k <= 0 || (k & 1) == 0) return nc_set_error(ctx, NC_ECV, "cv2.error: medianBlur ksize %d must be odd and positive", k);
        if (k > 27) return nc_set_error(ctx, NC_EUNSUPPORTED, "medianBlur ksize %d > 27", k);
        P.k = k;
        break;
    case NC_OP_ERODE:
    case NC_OP_DILATE:
        if (k < 0) return nc_set_error(ctx, NC_EINVAL, "ValueError: negative dimensions are not allowed");
        P.k = (k == 0) ? 3 : k;   // an empty kernel makes cv2 fall back to the 3x3 default
        if (P.k > 55) return nc_set_error(ctx, NC_EUNSUPPORTED, "morphology ksize %d > 55", k);
        break;
    case NC_OP_NL_DENOISE: {
        int rc = get_nl_lut(ctx, (float)v1, &P.nl_lut, &P.nl_lut_size);
        if (rc != NC_OK) return rc;
        dyn = 54 * 54;
        break;
    }
    case NC_OP_CLAHE:
        dyn = 2 * 64 * 256;
        break;
    case NC_OP_SALT_PEPPER: {
        long long m = (long long)(28 * 28 * v1);   // int(28*28*percent), preprocess.py:162
        if (m < 0) m = 0;
        if (m > 0) {
            if (!host_draws || draws_per_image < 4 * m)
                return nc_set_error(ctx, NC_EINVAL, "salt/pepper needs %lld host draws per picture", 4 * m);
            const size_t count = (size_t)n * 2 * m;
            int rc = nc_pinned_reserve(ctx, count * sizeof(unsigned short));
            if (rc != NC_OK) return rc;
            rc = nc_ws_reserve(ctx, count * sizeof(unsigned short));
            if (rc != NC_OK) return rc;
            // the pinned buffer is reused by the next call: wait for earlier copies out of it
            NC_CUDA(ctx, cudaStreamSynchronize(st));
            unsigned short* hp = (unsigned short*)ctx->pinned;
            for (int im = 0; im < n; ++im) {
                const double* d = host_draws + (size_t)im * draws_per_image;
                for (long long a = 0; a < 2 * m; ++a) {
                    const int i = (int)(d[2 * a] * 28), j = (int)(d[2 * a + 1] * 28);   // i = column, j = row
                    if (i < 0 || i >= 28 || j < 0 || j >= 28)
                        return nc_set_error(ctx, NC_EINVAL, "IndexError: draw outside [0,1)");
                    hp[(size_t)im * 2 * m + a] = (unsigned short)(j * 28 + i);
                }
            }
            NC_CUDA(ctx, cudaMemcpyAsync(ctx->ws, hp, count * sizeof(unsigned short), cudaMemcpyHostToDevice, st));
            P.sp_idx = (const unsigned short*)ctx->ws;
        }
        P.sp_m = (int)m;
        break;
    }
    default:
        break;
    }
    if (dyn > 48 * 1024)
        NC_CUDA(ctx, cudaFuncSetAttribute(ops_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
    ops_kernel<<<n, kThreads, dyn, st>>>(in, out, P);
    NC_LAUNCH_CHECK(ctx);
    return NC_OK;
}

// Host-side coefficient tables for the preprocess kernels.
//
// These are the few hundred coefficients that cv2 / Pillow derive from the image shape before
// touching pixels; they are computed once per shape on the host with the same expression order
// and precision as the libraries (this file must be compiled with -ffp-contract=off so that no
// multiply-add is fused) and handed to the kernels.
//   cv2.resize(INTER_CUBIC)      apps/preprocess/preprocess.py:29
//   PIL Image.resize((20,20))    apps/construction/util/construct_inference.py:315
//   cv2.GaussianBlur(k,k,0)      apps/preprocess/preprocess.py:134
#include <cmath>
#include <algorithm>
#include "nc_internal.h"

void nc_cv_cubic_axis(int src, int dst, CvCubicAxis* out) {
    // OpenCV: inv_scale = dsize/ssize (double); scale = 1./inv_scale;
    //         fx = (float)((dx+0.5)*scale - 0.5); sx = cvFloor(fx); fx -= sx;
    //         interpolateCubic(fx, cbuf) with A = -0.75f in float;
    //         alpha = saturate_cast<short>(cbuf[k] * INTER_RESIZE_COEF_SCALE)  (cvRound: half even)
    const double inv_scale = (double)dst / (double)src;
    const double scale = 1.0 / inv_scale;
    const float A = -0.75f;
    for (int d = 0; d < dst && d < 28; ++d) {
        volatile float fx = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor((double)fx);
        volatile float x = fx - (float)s;
        volatile float xp = x + 1.0f;
        volatile float t0 = A * xp;
        volatile float t1 = 5.0f * A;
        volatile float t2 = t0 - t1;
        volatile float t3 = t2 * xp;
        volatile float t4 = 8.0f * A;
        volatile float t5 = t3 + t4;
        volatile float t6 = t5 * xp;
        volatile float t7 = 4.0f * A;
        volatile float c0 = t6 - t7;
        volatile float a2 = A + 2.0f, a3 = A + 3.0f;
        volatile float u0 = a2 * x;
        volatile float u1 = u0 - a3;
        volatile float u2 = u1 * x;
        volatile float u3 = u2 * x;
        volatile float c1 = u3 + 1.0f;
        volatile float xm = 1.0f - x;
        volatile float v0 = a2 * xm;
        volatile float v1 = v0 - a3;
        volatile float v2 = v1 * xm;
        volatile float v3 = v2 * xm;
        volatile float c2 = v3 + 1.0f;
        volatile float w0 = 1.0f - c0;
        volatile float w1 = w0 - c1;
        volatile float c3 = w1 - c2;
        const float cs[4] = {c0, c1, c2, c3};
        for (int k = 0; k < 4; ++k) {
            volatile float scaled = cs[k] * 2048.0f;
            long r = std::lrint((double)scaled);  // nearbyint, round half to even
            r = std::max(-32768L, std::min(32767L, r));
            out->coef[d][k] = (short)r;
            int idx = s - 1 + k;
            out->ofs[d][k] = std::max(0, std::min(src - 1, idx));
        }
    }
}

static double pil_bicubic_filter(double x) {
    const double a = -0.5;
    if (x < 0.0) x = -x;
    if (x < 1.0) return ((a + 2.0) * x - (a + 3.0)) * x * x + 1;
    if (x < 2.0) return (((x - 5) * x + 8) * x - 4) * a;
    return 0.0;
}

void nc_pil_bicubic_axis(int in_size, int out_size, PilAxisHost* out) {
    // Pillow src/libImaging/Resample.c precompute_coeffs + normalize_coeffs_8bpc (PRECISION_BITS 22)
    const double scale = (double)in_size / (double)out_size;
    double filterscale = scale;
    if (filterscale < 1.0) filterscale = 1.0;
    const double support = 2.0 * filterscale;
    const int ksize = (int)std::ceil(support) * 2 + 1;
    out->out_size = out_size;
    out->ksize = ksize;
    out->xmin.assign(out_size, 0);
    out->cnt.assign(out_size, 0);
    out->kk.assign((size_t)out_size * ksize, 0);
    std::vector<double> w(ksize);
    const double ss = 1.0 / filterscale;
    for (int xx = 0; xx < out_size; ++xx) {
        const double center = (xx + 0.5) * scale;
        double ww = 0.0;
        int xmin = (int)(center - support + 0.5);
        if (xmin < 0) xmin = 0;
        int xmax = (int)(center + support + 0.5);
        if (xmax > in_size) xmax = in_size;
        xmax -= xmin;
        for (int x = 0; x < xmax; ++x) {
            double v = pil_bicubic_filter((x + xmin - center + 0.5) * ss);
            w[x] = v;
            ww += v;
        }
        for (int x = 0; x < xmax; ++x) {
            if (ww != 0.0) w[x] /= ww;
            double q = w[x] < 0 ? (-0.5 + w[x] * (double)(1 << 22)) : (0.5 + w[x] * (double)(1 << 22));
            out->kk[(size_t)xx * ksize + x] = (int)q;  // C truncation toward zero
        }
        out->xmin[xx] = xmin;
        out->cnt[xx] = xmax;
    }
}

std::vector<int> nc_gaussian_kernel_q8(int k) {
    // cv::getGaussianKernelBitExact(k, 0) + getGaussianKernelFixedPoint_ED(.., 8 fraction bits)
    std::vector<double> w(k);
    static const double t1[] = {1.0};
    static const double t3[] = {0.25, 0.5, 0.25};
    static const double t5[] = {0.0625, 0.25, 0.375, 0.25, 0.0625};
    static const double t7[] = {0.03125, 0.109375, 0.21875, 0.28125, 0.21875, 0.109375, 0.03125};
    static const double t9[] = {4 / 256., 13 / 256., 30 / 256., 51 / 256., 60 / 256., 51 / 256., 30 / 256., 13 / 256., 4 / 256.};
    const double* tab = k == 1 ? t1 : k == 3 ? t3 : k == 5 ? t5 : k == 7 ? t7 : k == 9 ? t9 : nullptr;
    if (tab) {
        for (int i = 0; i < k; ++i) w[i] = tab[i];
    } else {
        const double sigma = k * 0.15 + 0.35;
        const double scale2x = -0.125 / (sigma * sigma);
        const int n2 = (k - 1) / 2;
        double sum = 0.0;
        std::vector<double> vals(n2);
        for (int i = 0, x = 1 - k; i < n2; ++i, x += 2) {
            vals[i] = std::exp((double)(x * x) * scale2x);
            sum += vals[i];
        }
        sum = sum * 2.0 + 1.0;
        const double mul1 = 1.0 / sum;
        for (int i = 0; i < n2; ++i) w[i] = w[k - 1 - i] = vals[i] * mul1;
        w[n2] = mul1;
    }
    std::vector<int> q(k);
    const int n2 = k / 2;
    double err = 0.0;
    long total = 0;
    for (int i = 0; i < n2; ++i) {
        double adj = w[i] * 256.0 + err;
        long v0 = std::lrint(adj);
        err = adj - (double)v0;
        q[i] = q[k - 1 - i] = (int)v0;
        total += v0;
    }
    q[n2] = (int)(256 - 2 * total);
    return q;
}

extern "C" int nc_pp_cubic_table(int src, int32_t* ofs, int16_t* coef) {
    if (src <= 0 || !ofs || !coef) return nc_set_error(nullptr, NC_EINVAL, "bad arguments");
    CvCubicAxis ax;
    nc_cv_cubic_axis(src, 28, &ax);
    for (int d = 0; d < 28; ++d)
        for (int k = 0; k < 4; ++k) { ofs[d * 4 + k] = ax.ofs[d][k]; coef[d * 4 + k] = ax.coef[d][k]; }
    return NC_OK;
}

// IPP arithmetic of cv2.resize INTER_CUBIC (the wheel's default): position and Keys cubic in double, coefficients rounded to
// float32, never quantised (model fitted to the live call, see oracle/pp_restate.py:cv_resize_cubic_ipp)
void nc_ipp_cubic_axis(int src, int dst, IppCubicAxis* out) {
    const double scale = (double)src / (double)dst;
    const double A = -0.75;
    for (int d = 0; d < dst && d < 28; ++d) {
        const double f = (d + 0.5) * scale - 0.5;
        const int s = (int)std::floor(f);
        const double x = f - s;
        const double c0 = ((A * (x + 1) - 5 * A) * (x + 1) + 8 * A) * (x + 1) - 4 * A;
        const double c1 = ((A + 2) * x - (A + 3)) * x * x + 1;
        const double c2 = ((A + 2) * (1 - x) - (A + 3)) * (1 - x) * (1 - x) + 1;
        const double c3 = 1.0 - c0 - c1 - c2;
        const double cs[4] = {c0, c1, c2, c3};
        for (int k = 0; k < 4; ++k) {
            int o = s - 1 + k;
            o = o < 0 ? 0 : (o > src - 1 ? src - 1 : o);
            out->ofs[d][k] = o;
            out->coef[d][k] = (float)cs[k];
        }
    }
}

extern "C" int nc_pp_cubic_table_f32(int src, int32_t* ofs, float* coef) {
    if (src <= 0 || !ofs || !coef) return nc_set_error(nullptr, NC_EINVAL, "bad arguments");
    IppCubicAxis ax;
    nc_ipp_cubic_axis(src, 28, &ax);
    for (int d = 0; d < 28; ++d)
        for (int k = 0; k < 4; ++k) { ofs[d * 4 + k] = ax.ofs[d][k]; coef[d * 4 + k] = ax.coef[d][k]; }
    return NC_OK;
}

extern "C" int nc_pp_pil_table(int in_size, int out_size, int32_t* xmin, int32_t* cnt, int32_t* kk, int kk_capacity) {
    if (in_size <= 0 || out_size <= 0) return nc_set_error(nullptr, NC_EINVAL, "bad arguments");
    PilAxisHost ax;
    nc_pil_bicubic_axis(in_size, out_size, &ax);
    if (kk) {
        if (kk_capacity < out_size * ax.ksize || !xmin || !cnt) return nc_set_error(nullptr, NC_EINVAL, "kk buffer too small");
        for (int i = 0; i < out_size; ++i) { xmin[i] = ax.xmin[i]; cnt[i] = ax.cnt[i]; }
        for (size_t i = 0; i < ax.kk.size(); ++i) kk[i] = ax.kk[i];
    }
    return ax.ksize;
}

extern "C" int nc_pp_gaussian_q8(int k, int32_t* q) {
    if (k <= 0 || (k & 1) == 0 || !q) return nc_set_error(nullptr, NC_EINVAL, "k must be odd and positive");
    std::vector<int> v = nc_gaussian_kernel_q8(k);
    for (int i = 0; i < k; ++i) q[i] = v[i];
    return NC_OK;
}

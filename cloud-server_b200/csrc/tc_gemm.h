// Internal interface of the tcgen05 TF32 GEMM (gemm_tcgen05.cu).
#pragma once
#include "nc_internal.h"

// x3: 3xTF32 split (fp32-class accuracy) instead of one rounded TF32 product.
// a_mn / b_mn: operand is "MN-major" (stored [K][M] resp. [K][N], the non-reduction dimension
// contiguous) instead of "K-major" (stored [M][K] resp. [N][K]).
bool nc_tc_gemm_supported(const float* A, const float* B, int M, int N, int K, bool a_mn, bool b_mn);
size_t nc_tc_gemm_workspace(int sm_count, int M, int N, int K, int* splits_out, int* kbps_out);
int nc_tc_gemm(nc_ctx* ctx, const float* A, const float* B, float* C, const float* bias, int M, int N, int K, bool a_mn,
               bool b_mn, int act, float slope, float* ws, size_t ws_bytes, cudaStream_t st, bool x3);
int nc_tc_gemm_ex(nc_ctx* ctx, const float* A, const float* B, float* C, const float* bias, int M, int N, int K, bool a_mn,
                  bool b_mn, int lda, int ldb, int ka, int kb, int act, float slope, float* ws, size_t ws_bytes, cudaStream_t st,
                  bool x3);

namespace nck {
struct ConvGeom {
    int ih, iw, ic, oh, ow, oc, fh, fw, sh, sw, pt, pl;
};
}  // namespace nck

// implicit-GEMM convolution on the tensor cores (conv_tcgen05.cu).  mode: 0 forward (src = x, wmat = W,
// out = y), 1 dgrad (src = dy, wmat = W, out = dx), 2 wgrad (src = x, wmat = dy, out = dW)
bool nc_tc_conv_supported(int mode, const nck::ConvGeom& g, const float* src, const float* wmat);
void nc_tc_conv_dims(int mode, const nck::ConvGeom& g, int batch, int* M, int* N, int* K);
int nc_tc_conv(nc_ctx* ctx, int mode, const nck::ConvGeom& g, int batch, const float* src, const float* wmat, float* out,
               const float* bias, float* ws, size_t ws_bytes, cudaStream_t st, bool x3);

// the same forward / dgrad products with the A operand loaded by TMA in im2col mode (conv_im2col.cu):
// stride 1, output size == input size, reduced channel count a multiple of 32
bool nc_tc_conv_i2c_supported(int mode, const nck::ConvGeom& g, const float* src, const float* wmat);
// act: 0 none, 1 relu applied to the result (after the bias)
int nc_tc_conv_i2c(nc_ctx* ctx, int mode, const nck::ConvGeom& g, int batch, const float* src, const float* wmat, float* out,
                   const float* bias, float* ws, size_t ws_bytes, cudaStream_t st, bool x3, int act = 0);


// input gradient with the zero-padded picture resident in shared memory (conv_strip.cu): stride 1, SAME, Cin == 32,
// Cout 32 or 64, small pictures (H * (W + fw - 1) <= 384); single rounded TF32 products only
bool nc_tc_conv_strip_supported(int mode, const nck::ConvGeom& g, const float* src, const float* wmat, float* out);
int nc_tc_conv_strip_dgrad(nc_ctx* ctx, const nck::ConvGeom& g, int batch, const float* dy, const float* wmat, float* dx, cudaStream_t st);

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

// Internal engine entry points shared between the translation units of libse_b200.so.
#pragma once
#include "common.cuh"

namespace se {

struct ConvGeom {
  int Cin2, H, W, Cout2, kh, kw, sh, sw, ph, pw, Ho, Wo;
};

// stft_simt.cu
int stft_fwd_simt(const float* x, int64_t B, int64_t L, const float* W, int C2, int win, int hop,
                  int pad, float* out, int64_t T, cudaStream_t s);
int istft_fwd_simt(const float* spec, const float* phase, const float* A_packed, const float* window,
                   int64_t B, int C2, int64_t T, int win, int hop, int64_t crop, int64_t out_len,
                   float* out, cudaStream_t s);

// cconv_simt.cu
int cconv_fwd_simt(bool transposed, const float* x, const float* packed, const float* bias,
                   int64_t B, const ConvGeom& g, float* y, cudaStream_t s);

// cconv_umma.cu (tcgen05 bf16x3 implicit GEMM)
bool cconv_umma_supported(bool transposed, int64_t B, const ConvGeom& g);
int64_t cconv_umma_workspace_bytes(bool transposed, int64_t B, int Cin2, int H, int W, int Cout2,
                                   int kh, int kw, int sh, int sw);
int cconv_fwd_umma(bool transposed, const float* x, const float* packed, const float* bias,
                   int64_t B, const ConvGeom& g, float* y, void* ws, int64_t ws_bytes, cudaStream_t s);

// cconv_direct.cu (HBM-bound degenerate layers, exact fp32)
int cconv_fwd_small_cout(bool transposed, const float* x, const float* packed, const float* bias, int64_t B,
                         const ConvGeom& g, float* y, cudaStream_t s);

}  // namespace se

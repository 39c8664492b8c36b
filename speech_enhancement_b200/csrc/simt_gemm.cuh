// CUDA-core fp32 implicit GEMM used (a) for the degenerate layers that are HBM-bound by
// construction (K <= 20 or M = 2), (b) as the exact-fp32 engine for shapes the
// tcgen05 path does not take, and (c) as the on-device cross-check of the tensor path.
//
//   D[m, n] = sum_k A[m, k] * Bop(k, n)          A row-major [M, K] fp32 in global memory
//
// `Prob` supplies the implicit B operand and the epilogue:
//   struct Prob { int M, K; int64_t N;
//     struct Col;                                            // per-column context (decoded n)
//     __device__ Col  col(int64_t n) const;
//     __device__ float b(int k, const Col&) const;           // gathered operand element
//     __device__ void store(int m, int64_t n, float acc) const; };
#pragma once
#include "common.cuh"

namespace se {

constexpr int SG_BM = 64, SG_BN = 64, SG_BK = 16, SG_THREADS = 256;

template <class Prob>
__global__ void __launch_bounds__(SG_THREADS)
simt_gemm_kernel(const float* __restrict__ A, Prob p) {
  __shared__ float As[2][SG_BK][SG_BM + 4];
  __shared__ float Bs[2][SG_BK][SG_BN + 4];

  const int tid = threadIdx.x;
  const int64_t n0 = (int64_t)blockIdx.x * SG_BN;
  const int m0 = blockIdx.y * SG_BM;

  // B loader: this thread always gathers column (n0 + tid % 64), rows tid/64 + 4*i
  const int bn = tid & 63, bk0 = tid >> 6;
  const int64_t my_n = n0 + bn;
  const bool n_ok = my_n < p.N;
  typename Prob::Col colctx = p.col(n_ok ? my_n : 0);
  // A loader: k = tid % 16, rows tid/16 + 16*i
  const int ak = tid & 15, am0 = tid >> 4;

  const int ty = tid >> 4, tx = tid & 15;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  float ra[4], rb[4];
  auto gload = [&](int k0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int m = m0 + am0 + 16 * i, k = k0 + ak;
      ra[i] = (m < p.M && k < p.K) ? __ldg(A + (int64_t)m * p.K + k) : 0.f;
      int kb = k0 + bk0 + 4 * i;
      rb[i] = (n_ok && kb < p.K) ? p.b(kb, colctx) : 0.f;
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      As[buf][ak][am0 + 16 * i] = ra[i];
      Bs[buf][bk0 + 4 * i][bn] = rb[i];
    }
  };

  const int nk = (p.K + SG_BK - 1) / SG_BK;
  gload(0);
  sstore(0);
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) gload((kt + 1) * SG_BK);
#pragma unroll
    for (int k = 0; k < SG_BK; ++k) {
      float4 a = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
      float4 b = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (kt + 1 < nk) sstore(buf ^ 1);
    __syncthreads();
  }

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int m = m0 + ty * 4 + i;
    if (m >= p.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int64_t n = n0 + tx * 4 + j;
      if (n < p.N) p.store(m, n, acc[i][j]);
    }
  }
}

template <class Prob>
inline int launch_simt_gemm(const float* A, const Prob& p, cudaStream_t s, const char* what) {
  if (p.M <= 0 || p.N <= 0) return SE_OK;
  int64_t gx = (p.N + SG_BN - 1) / SG_BN;
  int gy = (p.M + SG_BM - 1) / SG_BM;
  if (gx > 2147483647LL) return fail(SE_ERR_UNSUPPORTED, "%s: N too large", what);
  simt_gemm_kernel<Prob><<<dim3((unsigned)gx, (unsigned)gy), SG_THREADS, 0, s>>>(A, p);
  return check_launch(what);
}

}  // namespace se

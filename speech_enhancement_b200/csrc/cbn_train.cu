// Train-mode complex batch norm (the half that cannot be folded into the conv weights):
// batch statistics and the 2x2 whitening of /root/reference/models/modules/complex_nn.py:244-322 as four HBM-bound
// passes over an NCHW fp32 tensor x [B, 2C, H, W] (channels [0, C) real, [C, 2C) imaginary) instead of the ~40
// full-tensor ATen launches the reference's forward + autograd graph issue per layer.
//
//   forward   cbn_stats_kernel   per complex channel: sum xr, xi, xr^2, xi^2, xr*xi           (1 read)
//             [host, C-sized]    mean M, covariance V, Z = W V^-1/2, c = B - Z M                (complex_nn.py:266-322)
//             cbn_apply_kernel   y = Z x + c                                                     (1 read, 1 write)
//   backward  cbn_bwd_stats      per channel: sum gr, gi, gr*xr, gr*xi, gi*xr, gi*xi            (2 reads)
//             [host, C-sized]    dL/dZ, dL/dc -> dL/d(sums), dL/dW, dL/dB through the closed form (autograd on [C] tensors)
//             cbn_bwd_apply      dx = Z^T g + a0 + A x  (the statistics' dependence on x)        (2 reads, 1 write)
//
// Partial sums are accumulated in fp64 (raw second moments of un-centred data would cancel in fp32) and written as
// [C][splits][K] partials that the host adds in a fixed order: deterministic, no atomics.
#include "common.cuh"

namespace se {
namespace {

constexpr int CBN_THREADS = 256;

template <int K>
__device__ __forceinline__ void block_reduce_store(double (&v)[K], double* __restrict__ dst) {
  __shared__ double red[K][CBN_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double s = v[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) red[k][warp] = s;
  }
  __syncthreads();
  if (threadIdx.x < K) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < CBN_THREADS / 32; ++w) s += red[threadIdx.x][w];
    dst[threadIdx.x] = s;
  }
}

// grid (C, splits): block (c, s) walks the (b, pixel) pairs of channel c with stride `splits`
__global__ void __launch_bounds__(CBN_THREADS)
cbn_stats_kernel(const float* __restrict__ x, int64_t B, int C, int64_t P, double* __restrict__ partial) {
  const int c = blockIdx.x, splits = gridDim.y;
  double acc[5] = {0, 0, 0, 0, 0};
  const int64_t chunks = (P + CBN_THREADS - 1) / CBN_THREADS;          // per (b): pixel chunks of one block width
  for (int64_t w = blockIdx.y; w < B * chunks; w += splits) {
    const int64_t b = w / chunks, p = (w - b * chunks) * CBN_THREADS + threadIdx.x;
    if (p < P) {
      const double xr = (double)__ldg(x + ((b * 2 * C + c) * P + p)), xi = (double)__ldg(x + ((b * 2 * C + C + c) * P + p));
      acc[0] += xr; acc[1] += xi; acc[2] += xr * xr; acc[3] += xi * xi; acc[4] += xr * xi;
    }
  }
  block_reduce_store<5>(acc, partial + ((int64_t)c * splits + blockIdx.y) * 5);
}

// coef [6][C] = Zrr, Zri, Zir, Zii, cr, ci.   y = Z x + c
__global__ void __launch_bounds__(CBN_THREADS)
cbn_apply_kernel(const float* __restrict__ x, const float* __restrict__ coef, int64_t B, int C, int64_t P,
                 float* __restrict__ y) {
  const int c = blockIdx.x;
  const float zrr = coef[c], zri = coef[C + c], zir = coef[2 * C + c], zii = coef[3 * C + c], cr = coef[4 * C + c],
              ci = coef[5 * C + c];
  const int64_t chunks = (P + CBN_THREADS - 1) / CBN_THREADS;
  for (int64_t w = blockIdx.y; w < B * chunks; w += gridDim.y) {
    const int64_t b = w / chunks, p = (w - b * chunks) * CBN_THREADS + threadIdx.x;
    if (p < P) {
      const int64_t ir = (b * 2 * C + c) * P + p, ii = (b * 2 * C + C + c) * P + p;
      const float xr = __ldg(x + ir), xi = __ldg(x + ii);
      y[ir] = fmaf(zrr, xr, fmaf(zri, xi, cr));
      y[ii] = fmaf(zir, xr, fmaf(zii, xi, ci));
    }
  }
}

__global__ void __launch_bounds__(CBN_THREADS)
cbn_bwd_stats_kernel(const float* __restrict__ x, const float* __restrict__ g, int64_t B, int C, int64_t P,
                     double* __restrict__ partial) {
  const int c = blockIdx.x, splits = gridDim.y;
  double acc[6] = {0, 0, 0, 0, 0, 0};
  const int64_t chunks = (P + CBN_THREADS - 1) / CBN_THREADS;
  for (int64_t w = blockIdx.y; w < B * chunks; w += splits) {
    const int64_t b = w / chunks, p = (w - b * chunks) * CBN_THREADS + threadIdx.x;
    if (p < P) {
      const int64_t ir = (b * 2 * C + c) * P + p, ii = (b * 2 * C + C + c) * P + p;
      const double xr = (double)__ldg(x + ir), xi = (double)__ldg(x + ii);
      const double gr = (double)__ldg(g + ir), gi = (double)__ldg(g + ii);
      acc[0] += gr; acc[1] += gi; acc[2] += gr * xr; acc[3] += gr * xi; acc[4] += gi * xr; acc[5] += gi * xi;
    }
  }
  block_reduce_store<6>(acc, partial + ((int64_t)c * splits + blockIdx.y) * 6);
}

// coef [10][C] = Zrr, Zri, Zir, Zii, ar0, arr, ari, ai0, air, aii:
//   dxr = Zrr gr + Zir gi + ar0 + arr xr + ari xi;   dxi = Zri gr + Zii gi + ai0 + air xr + aii xi
__global__ void __launch_bounds__(CBN_THREADS)
cbn_bwd_apply_kernel(const float* __restrict__ x, const float* __restrict__ g, const float* __restrict__ coef, int64_t B,
                     int C, int64_t P, float* __restrict__ dx) {
  const int c = blockIdx.x;
  float k[10];
#pragma unroll
  for (int i = 0; i < 10; ++i) k[i] = coef[i * C + c];
  const int64_t chunks = (P + CBN_THREADS - 1) / CBN_THREADS;
  for (int64_t w = blockIdx.y; w < B * chunks; w += gridDim.y) {
    const int64_t b = w / chunks, p = (w - b * chunks) * CBN_THREADS + threadIdx.x;
    if (p < P) {
      const int64_t ir = (b * 2 * C + c) * P + p, ii = (b * 2 * C + C + c) * P + p;
      const float xr = __ldg(x + ir), xi = __ldg(x + ii), gr = __ldg(g + ir), gi = __ldg(g + ii);
      dx[ir] = fmaf(k[0], gr, fmaf(k[2], gi, fmaf(k[5], xr, fmaf(k[6], xi, k[4]))));
      dx[ii] = fmaf(k[1], gr, fmaf(k[3], gi, fmaf(k[8], xr, fmaf(k[9], xi, k[7]))));
    }
  }
}

inline int cbn_splits_for(int64_t B, int C, int64_t P) {
  const int64_t chunks = B * ((P + CBN_THREADS - 1) / CBN_THREADS);
  int64_t want = ((int64_t)num_sms() * 8 + C - 1) / C;            // ~8 resident blocks per SM over all channels
  if (want > chunks) want = chunks;
  if (want < 1) want = 1;
  if (want > 1024) want = 1024;
  return (int)want;
}

}  // namespace
}  // namespace se

extern "C" int se_cbn_splits(int64_t B, int C, int H, int W) {
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0) return 0;
  return se::cbn_splits_for(B, C, (int64_t)H * W);
}

#define SE_CBN_CHECK(nm)                                                                                       \
  SE_REQUIRE(B >= 0 && C > 0 && H > 0 && W > 0 && C <= 65535, nm ": bad dims");                                 \
  if (B == 0) return SE_OK;                                                                                    \
  const int64_t P = (int64_t)H * W;                                                                            \
  const int splits = se::cbn_splits_for(B, C, P);                                                              \
  const dim3 grid((unsigned)C, (unsigned)splits)

extern "C" int se_cbn_stats(const float* x, int64_t B, int C, int H, int W, double* partial, void* stream) {
  SE_REQUIRE(x && partial, "se_cbn_stats: null pointer");
  SE_CBN_CHECK("se_cbn_stats");
  se::cbn_stats_kernel<<<grid, se::CBN_THREADS, 0, (cudaStream_t)stream>>>(x, B, C, P, partial);
  return se::check_launch("se_cbn_stats");
}

extern "C" int se_cbn_apply(const float* x, const float* coef, int64_t B, int C, int H, int W, float* y, void* stream) {
  SE_REQUIRE(x && coef && y, "se_cbn_apply: null pointer");
  SE_CBN_CHECK("se_cbn_apply");
  se::cbn_apply_kernel<<<grid, se::CBN_THREADS, 0, (cudaStream_t)stream>>>(x, coef, B, C, P, y);
  return se::check_launch("se_cbn_apply");
}

extern "C" int se_cbn_bwd_stats(const float* x, const float* g, int64_t B, int C, int H, int W, double* partial,
                                void* stream) {
  SE_REQUIRE(x && g && partial, "se_cbn_bwd_stats: null pointer");
  SE_CBN_CHECK("se_cbn_bwd_stats");
  se::cbn_bwd_stats_kernel<<<grid, se::CBN_THREADS, 0, (cudaStream_t)stream>>>(x, g, B, C, P, partial);
  return se::check_launch("se_cbn_bwd_stats");
}

extern "C" int se_cbn_bwd_apply(const float* x, const float* g, const float* coef, int64_t B, int C, int H, int W,
                                float* dx, void* stream) {
  SE_REQUIRE(x && g && coef && dx, "se_cbn_bwd_apply: null pointer");
  SE_CBN_CHECK("se_cbn_bwd_apply");
  se::cbn_bwd_apply_kernel<<<grid, se::CBN_THREADS, 0, (cudaStream_t)stream>>>(x, g, coef, B, C, P, dx);
  return se::check_launch("se_cbn_bwd_apply");
}

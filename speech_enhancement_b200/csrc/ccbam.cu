// CCBAM attention on a skip connection (/root/reference/models/modules/ccbam.py:28-106) as three
// HBM-bound passes over the CL2 activation instead of ~25 PyTorch launches over NCHW copies
// (side item of rank 4):
//   1. per (b, channel) sum and max over the pixels            (channel attention descriptors, :45-63)
//   2. per pixel mean / max over the real and the imaginary channels of x * ca   (spatial pooling, :71-80)
//   3. out = x * ca + sa (real channels + sa[0], imaginary + sa[1]), written back as CL2   (:98-106)
// The shared MLP, the 7x7 attention convolution and the sigmoids act on tiny tensors and stay on the host
// framework.  CL2 = channels-last bf16 hi|lo planes [B][2][H*W][Cp], value = hi + lo.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <algorithm>
#include <cfloat>
#include <cstdint>
#include "common.cuh"
#include "../../include/se_b200.h"

namespace se {

__device__ __forceinline__ void load8(const __nv_bfloat16* hi, const __nv_bfloat16* lo, float (&v)[8]) {
  const uint4 a = *reinterpret_cast<const uint4*>(hi), b = *reinterpret_cast<const uint4*>(lo);
  const __nv_bfloat16* ph = reinterpret_cast<const __nv_bfloat16*>(&a);
  const __nv_bfloat16* pl = reinterpret_cast<const __nv_bfloat16*>(&b);
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = __bfloat162float(ph[j]) + __bfloat162float(pl[j]);
}

// grid (nblk, B), block 256 = 16 pixel lanes x (Cp/8 <= 16) channel groups; partial[b][blk][2][Cp]
__global__ void __launch_bounds__(256)
ccbam_channel_stats_kernel(const __nv_bfloat16* __restrict__ x, int64_t P, int Cp, float* __restrict__ partial) {
  __shared__ float s_sum[16][128 + 1], s_max[16][128 + 1];
  const int groups = Cp / 8;
  const int cg = threadIdx.x % groups, pl = threadIdx.x / groups, npl = 256 / groups;
  const int64_t b = blockIdx.y, nblk = gridDim.x;
  const int64_t per = (P + nblk - 1) / nblk, p0 = blockIdx.x * per, p1 = min(p0 + per, P);
  const __nv_bfloat16* hi = x + (b * 2) * P * Cp;
  const __nv_bfloat16* lo = hi + P * Cp;
  float sum[8], mx[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { sum[j] = 0.f; mx[j] = -FLT_MAX; }
  if (pl < npl) {
    for (int64_t p = p0 + pl; p < p1; p += npl) {
      float v[8];
      load8(hi + p * Cp + cg * 8, lo + p * Cp + cg * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) { sum[j] += v[j]; mx[j] = fmaxf(mx[j], v[j]); }
    }
  }
  // reduce over the pixel lanes through shared memory (Cp <= 128)
  for (int r = 0; r < (npl + 15) / 16; ++r) {
    __syncthreads();
    const int row = pl - r * 16;
    if (row >= 0 && row < 16 && pl < npl) {
#pragma unroll
      for (int j = 0; j < 8; ++j) { s_sum[row][cg * 8 + j] = sum[j]; s_max[row][cg * 8 + j] = mx[j]; }
    }
    __syncthreads();
    const int rows = min(16, npl - r * 16);
    if (threadIdx.x < Cp) {
      float a = 0.f, m = -FLT_MAX;
      for (int i = 0; i < rows; ++i) { a += s_sum[i][threadIdx.x]; m = fmaxf(m, s_max[i][threadIdx.x]); }
      float* out = partial + ((b * nblk + blockIdx.x) * 2) * Cp + threadIdx.x;
      if (r == 0) { out[0] = a; out[Cp] = m; } else { out[0] += a; out[Cp] = fmaxf(out[Cp], m); }
    }
  }
}

// one block per (pixel chunk, b); 16 lanes per pixel, 8 channels per lane: lanes [0, C/16) real, next C/16 imaginary
__global__ void __launch_bounds__(256)
ccbam_spatial_pool_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ ca, int64_t P, int C, int Cp,
                          float* __restrict__ pooled) {
  __shared__ float s_ca[128];
  const int64_t b = blockIdx.y;
  for (int i = threadIdx.x; i < Cp; i += 256) s_ca[i] = i < C ? ca[b * C + i] : 0.f;
  __syncthreads();
  const int half_lanes = C / 16;                     // lanes per half (C/2 channels / 8)
  const int lane16 = threadIdx.x & 15;
  const __nv_bfloat16* hi = x + (b * 2) * P * Cp;
  const __nv_bfloat16* lo = hi + P * Cp;
  const float inv = 2.f / (float)C;
  for (int64_t p = (int64_t)blockIdx.x * 16 + (threadIdx.x >> 4); p < P; p += (int64_t)gridDim.x * 16) {
    float sum = 0.f, mx = -FLT_MAX;
    if (lane16 < 2 * half_lanes) {
      float v[8];
      load8(hi + p * Cp + lane16 * 8, lo + p * Cp + lane16 * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) { const float t = v[j] * s_ca[lane16 * 8 + j]; sum += t; mx = fmaxf(mx, t); }
    }
    // reduce inside each half (lanes 0..half_lanes-1 and half_lanes..2*half_lanes-1); half_lanes is 1, 2, 4 or 8
    for (int off = half_lanes >> 1; off > 0; off >>= 1) {
      sum += __shfl_xor_sync(0xffffffffu, sum, off);
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    }
    if (lane16 == 0 || lane16 == half_lanes) {
      const int part = lane16 == 0 ? 0 : 1;
      pooled[((b * 4 + part * 2 + 0) * P) + p] = sum * inv;      // mean over C/2 channels
      pooled[((b * 4 + part * 2 + 1) * P) + p] = mx;
    }
  }
}

__global__ void __launch_bounds__(256)
ccbam_apply_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ ca, const float* __restrict__ sa,
                   int64_t P, int C, int Cp, __nv_bfloat16* __restrict__ y) {
  __shared__ float s_ca[128];
  const int64_t b = blockIdx.y;
  for (int i = threadIdx.x; i < Cp; i += 256) s_ca[i] = i < C ? ca[b * C + i] : 0.f;
  __syncthreads();
  const int groups = Cp / 8;
  const int64_t total = P * groups;
  const __nv_bfloat16* hi = x + (b * 2) * P * Cp;
  const __nv_bfloat16* lo = hi + P * Cp;
  __nv_bfloat16* yh = y + (b * 2) * P * Cp;
  __nv_bfloat16* yl = yh + P * Cp;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int cg = (int)(i % groups);
    const int64_t p = i / groups;
    float v[8];
    load8(hi + p * Cp + cg * 8, lo + p * Cp + cg * 8, v);
    const float s0 = __ldg(sa + (b * 2 + 0) * P + p), s1 = __ldg(sa + (b * 2 + 1) * P + p);
    __align__(16) __nv_bfloat16 oh[8], ol[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = cg * 8 + j;
      float o = 0.f;
      if (c < C) o = v[j] * s_ca[c] + (c < C / 2 ? s0 : s1);
      oh[j] = __float2bfloat16_rn(o);
      ol[j] = __float2bfloat16_rn(o - __bfloat162float(oh[j]));
    }
    *reinterpret_cast<uint4*>(yh + p * Cp + cg * 8) = *reinterpret_cast<const uint4*>(oh);
    *reinterpret_cast<uint4*>(yl + p * Cp + cg * 8) = *reinterpret_cast<const uint4*>(ol);
  }
}

}  // namespace se

static int ccbam_check(const char* nm, int64_t B, int64_t P, int C, int Cp) {
  if (B < 0 || P <= 0 || C <= 0 || C % 16 != 0 || C > 128 || Cp < C || Cp % 64 != 0 || Cp > 128 || B > 65535)
    return se::fail(SE_ERR_UNSUPPORTED, "%s: needs C %% 16 == 0, C <= Cp <= 128, B <= 65535", nm);
  if ((C / 16) & (C / 16 - 1)) return se::fail(SE_ERR_UNSUPPORTED, "%s: C / 16 must be a power of two", nm);
  return SE_OK;
}

extern "C" int se_ccbam_num_partials(int64_t P) { return (int)std::min<int64_t>(std::max<int64_t>(P / 2048, 1), 64); }

extern "C" int se_ccbam_channel_stats(const void* x_cl2, int64_t B, int64_t P, int C, int Cp, float* partial, void* stream) {
  SE_REQUIRE(x_cl2 && partial, "se_ccbam_channel_stats: null pointer");
  int rc = ccbam_check("se_ccbam_channel_stats", B, P, C, Cp);
  if (rc != SE_OK || B == 0) return rc;
  dim3 grid((unsigned)se_ccbam_num_partials(P), (unsigned)B);
  se::ccbam_channel_stats_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const __nv_bfloat16*>(x_cl2), P, Cp,
                                                                         partial);
  return se::check_launch("se_ccbam_channel_stats");
}

extern "C" int se_ccbam_spatial_pool(const void* x_cl2, const float* ca, int64_t B, int64_t P, int C, int Cp, float* pooled,
                                     void* stream) {
  SE_REQUIRE(x_cl2 && ca && pooled, "se_ccbam_spatial_pool: null pointer");
  int rc = ccbam_check("se_ccbam_spatial_pool", B, P, C, Cp);
  if (rc != SE_OK || B == 0) return rc;
  dim3 grid((unsigned)std::min<int64_t>((P + 15) / 16, 1024), (unsigned)B);
  se::ccbam_spatial_pool_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const __nv_bfloat16*>(x_cl2), ca, P, C,
                                                                        Cp, pooled);
  return se::check_launch("se_ccbam_spatial_pool");
}

extern "C" int se_ccbam_apply(const void* x_cl2, const float* ca, const float* sa, int64_t B, int64_t P, int C, int Cp,
                              void* y_cl2, void* stream) {
  SE_REQUIRE(x_cl2 && ca && sa && y_cl2, "se_ccbam_apply: null pointer");
  int rc = ccbam_check("se_ccbam_apply", B, P, C, Cp);
  if (rc != SE_OK || B == 0) return rc;
  dim3 grid((unsigned)std::min<int64_t>((P * (Cp / 8) + 255) / 256, 2048), (unsigned)B);
  se::ccbam_apply_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const __nv_bfloat16*>(x_cl2), ca, sa, P, C, Cp,
                                                                 reinterpret_cast<__nv_bfloat16*>(y_cl2));
  return se::check_launch("se_ccbam_apply");
}

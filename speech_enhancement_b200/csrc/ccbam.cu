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

// channel attention map from the partial statistics (ccbam.py:45-63): mean / max descriptors -> shared MLP
// (w1 [R2][C], ReLU, w2 [C][R2]; dense real matrices -- the reference's ComplexLinear is block diagonal over the real
// and imaginary halves, complex_nn.py:106-113) on both -> sigmoid of the sum.  One block per batch row.
__global__ void __launch_bounds__(128)
ccbam_channel_attn_kernel(const float* __restrict__ partial, int nblk, float inv_p, const float* __restrict__ w1,
                          const float* __restrict__ w2, int C, int Cp, int R2, float* __restrict__ ca) {
  __shared__ float v[2][128], hdn[2][64];
  const int64_t b = blockIdx.x;
  const int c = threadIdx.x;
  if (c < C) {
    float a = 0.f, m = -FLT_MAX;
    for (int k = 0; k < nblk; ++k) {
      const float* pp = partial + ((b * nblk + k) * 2) * Cp + c;
      a += pp[0];
      m = fmaxf(m, pp[Cp]);
    }
    v[0][c] = a * inv_p;
    v[1][c] = m;
  }
  __syncthreads();
  if (c < 2 * R2) {
    const int which = c / R2, j = c - which * R2;
    float a = 0.f;
    for (int k = 0; k < C; ++k) a = fmaf(__ldg(w1 + j * C + k), v[which][k], a);
    hdn[which][j] = fmaxf(a, 0.f);
  }
  __syncthreads();
  if (c < C) {
    float a = 0.f, m = 0.f;
    for (int j = 0; j < R2; ++j) {
      const float w = __ldg(w2 + c * R2 + j);
      a = fmaf(w, hdn[0][j], a);
      m = fmaf(w, hdn[1][j], m);
    }
    ca[b * C + c] = 1.f / (1.f + expf(-(a + m)));
  }
}

// spatial attention map (ccbam.py:65-86): 7x7 conv over the four pooled maps (zero padding 3) with the complex
// conv + eval-mode complex BN folded into `packed` [2][4*49] / `bias` [2], ReLU, sigmoid.  32 x 8 output pixels per
// block, the 14 x 38 x 4 input tile and the weights in shared memory.
__global__ void __launch_bounds__(256)
ccbam_spatial_attn_kernel(const float* __restrict__ pooled, const float* __restrict__ packed, const float* __restrict__ bias,
                          int H, int W, float* __restrict__ sa) {
  constexpr int TW = 32, TH = 8, R = 3, SW = TW + 2 * R, SH = TH + 2 * R;
  __shared__ float tile[4][SH][SW + 1];
  __shared__ float2 wts[4 * 49];
  const int64_t b = blockIdx.z, P = (int64_t)H * W;
  const int w0 = blockIdx.x * TW, h0 = blockIdx.y * TH;
  for (int i = threadIdx.x; i < 4 * 49; i += 256) wts[i] = make_float2(__ldg(packed + i), __ldg(packed + 4 * 49 + i));
  for (int i = threadIdx.x; i < 4 * SH * SW; i += 256) {
    const int ch = i / (SH * SW), r = (i / SW) % SH, col = i % SW;
    const int h = h0 + r - R, w = w0 + col - R;
    tile[ch][r][col] = (h >= 0 && h < H && w >= 0 && w < W) ? __ldg(pooled + (b * 4 + ch) * P + (int64_t)h * W + w) : 0.f;
  }
  __syncthreads();
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int h = h0 + ty, w = w0 + tx;
  float a0 = __ldg(bias), a1 = __ldg(bias + 1);
#pragma unroll
  for (int ch = 0; ch < 4; ++ch)
#pragma unroll
    for (int dh = 0; dh < 7; ++dh)
#pragma unroll
      for (int dw = 0; dw < 7; ++dw) {
        const float x = tile[ch][ty + dh][tx + dw];
        const float2 wv = wts[(ch * 7 + dh) * 7 + dw];
        a0 = fmaf(wv.x, x, a0);
        a1 = fmaf(wv.y, x, a1);
      }
  if (h < H && w < W) {
    sa[(b * 2 + 0) * P + (int64_t)h * W + w] = 1.f / (1.f + expf(-fmaxf(a0, 0.f)));
    sa[(b * 2 + 1) * P + (int64_t)h * W + w] = 1.f / (1.f + expf(-fmaxf(a1, 0.f)));
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

extern "C" int se_ccbam_channel_attn(const float* partial, int nblk, int64_t P, const float* w1, const float* w2, int64_t B,
                                     int C, int Cp, int R2, float* ca, void* stream) {
  SE_REQUIRE(partial && w1 && w2 && ca, "se_ccbam_channel_attn: null pointer");
  int rc = ccbam_check("se_ccbam_channel_attn", B, P, C, Cp);
  if (rc != SE_OK || B == 0) return rc;
  if (nblk <= 0 || R2 <= 0 || R2 > 64) return se::fail(SE_ERR_UNSUPPORTED, "se_ccbam_channel_attn: needs 0 < R2 <= 64");
  se::ccbam_channel_attn_kernel<<<(unsigned)B, 128, 0, (cudaStream_t)stream>>>(partial, nblk, 1.0f / (float)P, w1, w2, C, Cp,
                                                                               R2, ca);
  return se::check_launch("se_ccbam_channel_attn");
}

extern "C" int se_ccbam_spatial_attn(const float* pooled, const float* packed, const float* bias, int64_t B, int H, int W,
                                     float* sa, void* stream) {
  SE_REQUIRE(pooled && packed && bias && sa, "se_ccbam_spatial_attn: null pointer");
  SE_REQUIRE(B >= 0 && H > 0 && W > 0 && B <= 65535 && (H + 7) / 8 <= 65535, "se_ccbam_spatial_attn: bad dims");
  if (B == 0) return SE_OK;
  dim3 grid((unsigned)((W + 31) / 32), (unsigned)((H + 7) / 8), (unsigned)B);
  se::ccbam_spatial_attn_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(pooled, packed, bias, H, W, sa);
  return se::check_launch("se_ccbam_spatial_attn");
}

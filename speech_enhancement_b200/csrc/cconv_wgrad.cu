// Weight gradient of the packed complex convolution on tcgen05 (autograd of ComplexConv2d /
// ComplexConvTranspose2d, /root/reference/models/modules/complex_nn.py:52-91; the reference gets it
// from ATen's conv backward).
//
// Both flavours reduce to ONE form.  With U the operand sampled on the un-shifted grid and V the one
// sampled through the taps,
//     G[tap][cu][cv] = sum_{b, hu, wu} U[b, cu, hu, wu] * V[b, cv, hu*sh + dh - ph, wu*sw + dw - pw]
//   conv :  U = dY (2Co channels, output grid), V = X  (2Ci)  ->  dWp[co][ci][tap] = G[tap][co][ci]
//   convT:  U = X  (2Ci channels, input grid),  V = dY (2Co)  ->  dWt[ci][co][tap] = G[tap][ci][co]
// This is a GEMM whose K dimension is the pixels.  Activations are staged as "NC2": NCHW bf16 hi|lo
// planes [B][2][C][H][Wp] (Wp = W padded to 8 so TMA strides are 16-byte multiples) -- pixels of one
// image row are contiguous, i.e. K-major, so a TMA box {64 px, 1 row, C channels} lands in shared
// memory directly as a SWIZZLE_128B UMMA operand tile.  TMA wants the innermost start coordinate 16-byte
// aligned, so the column shift of a tap cannot be a box coordinate: V is staged once per kernel COLUMN dw,
// already gathered along W (Vg[dw][.., wu] = V[.., wu*sw + dw - pw], zero outside), and the row shift /
// row stride is the box's row coordinate hu*sh + dh - ph; zero padding in H is out-of-bounds fill.  Same bf16x3 split as the forward
// (lo*hi + hi*hi + hi*lo), fp32 accumulation in TMEM over the CTA's share of the pixels, then one
// red.global.add per element into G (split-K over image rows).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include "common.cuh"
#include "umma.cuh"
#include "../../include/se_b200.h"

namespace se {

__device__ __forceinline__ void split_bf16_w(float v, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(v);
  lo = __float2bfloat16_rn(v - __bfloat162float(hi));
}

// NCHW fp32 -> NC2 (bf16 hi|lo planes, rows padded to Wp)
__global__ void __launch_bounds__(256)
nc2_from_nchw_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int64_t rows, int W, int Wp,
                     int64_t plane_rows) {
  // rows = B*C*H image rows; out index: ((b*2 + s)*C*H + ch) * Wp + w  with ch = row % (C*H)
  const int vec = Wp / 8;
  const int64_t total = rows * vec;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int v8 = (int)(i % vec);
    const int64_t row = i / vec;
    const int64_t b = row / plane_rows, ch = row - b * plane_rows;
    __align__(16) __nv_bfloat16 hi[8], lo[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int w = v8 * 8 + j;
      split_bf16_w(w < W ? __ldg(x + row * W + w) : 0.f, hi[j], lo[j]);
    }
    const int64_t o = ((b * 2) * plane_rows + ch) * Wp + v8 * 8;
    *reinterpret_cast<uint4*>(out + o) = *reinterpret_cast<const uint4*>(hi);
    *reinterpret_cast<uint4*>(out + o + plane_rows * Wp) = *reinterpret_cast<const uint4*>(lo);
  }
}

// NCHW fp32 -> kw column-gathered NC2 copies: out[dw][b][s][c][h][wu] = x[b][c][h][wu*sw + dw - pw]
__global__ void __launch_bounds__(256)
nc2_gather_w_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int64_t rows, int W, int Wu, int Wup,
                    int64_t plane_rows, int kw, int sw, int pw, int64_t copy_elems) {
  const int vec = Wup / 8;
  const int64_t total = rows * vec * kw;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int v8 = (int)(i % vec);
    int64_t r = i / vec;
    const int64_t row = r % rows;
    const int dw = (int)(r / rows);
    const int64_t b = row / plane_rows, ch = row - b * plane_rows;
    __align__(16) __nv_bfloat16 hi[8], lo[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int wu = v8 * 8 + j;
      const int w = wu * sw + dw - pw;
      split_bf16_w((wu < Wu && w >= 0 && w < W) ? __ldg(x + row * W + w) : 0.f, hi[j], lo[j]);
    }
    const int64_t o = (int64_t)dw * copy_elems + ((b * 2) * plane_rows + ch) * Wup + v8 * 8;
    *reinterpret_cast<uint4*>(out + o) = *reinterpret_cast<const uint4*>(hi);
    *reinterpret_cast<uint4*>(out + o + plane_rows * Wup) = *reinterpret_cast<const uint4*>(lo);
  }
}

struct WgradParams {
  int Cu, Cv, Hu, Wu, kw, sh, sw, ph, pw, taps;
  int64_t B;
  int m_tiles, n_tiles, n_tile, tg, tap_groups;      // tg = taps per group (accumulators per CTA)
  int ksplit;
  int64_t rows_total, rows_per_split;                // rows = (b, hu) pairs
  int a_bytes, b_bytes, stage_bytes, stages;
  int tmem_cols;
  float* G;                                          // [taps][Cu][Cv]
};

constexpr int kWgThreads = 192;                      // warp 0 TMA, warp 1 MMA, warps 2-5 epilogue

__global__ void __launch_bounds__(kWgThreads, 1)
cconv_wgrad_kernel(const __grid_constant__ CUtensorMap tmU, const __grid_constant__ CUtensorMap tmV,
                   const __grid_constant__ WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full[4], empty[4], done;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  // work item
  int item = blockIdx.x;
  const int tgi = item % p.tap_groups; item /= p.tap_groups;
  const int nt = item % p.n_tiles; item /= p.n_tiles;
  const int mt = item;
  const int tap0 = tgi * p.tg, ntap = min(p.tg, p.taps - tap0);
  const int64_t row_begin = (int64_t)blockIdx.y * p.rows_per_split;
  const int64_t row_end = min(row_begin + p.rows_per_split, p.rows_total);
  const int wblocks = (p.Wu + 63) / 64;
  const int64_t nkb = (row_end > row_begin ? row_end - row_begin : 0) * wblocks;

  if (threadIdx.x == 0) {
    ptx::prefetch_tmap(&tmU);
    ptx::prefetch_tmap(&tmV);
    for (int s = 0; s < p.stages; ++s) { ptx::mbar_init(&full[s], 1); ptx::mbar_init(&empty[s], 1); }
    ptx::mbar_init(&done, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(&tmem_slot, (uint32_t)p.tmem_cols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (nkb > 0) {
    if (warp == 0) {
      if (ptx::elect_one()) {
        uint32_t s = 0, ph = 1;
        for (int64_t row = row_begin; row < row_end; ++row) {
          const int b = (int)(row / p.Hu), hu = (int)(row - (int64_t)b * p.Hu);
          for (int wb = 0; wb < wblocks; ++wb) {
            ptx::mbar_wait(&empty[s], ph);
            uint8_t* st = smem + (size_t)s * p.stage_bytes;
            ptx::mbar_arrive_expect_tx(&full[s], (uint32_t)(2 * p.a_bytes + 2 * ntap * p.b_bytes));
            {
            ptx::tma_load_4d(st, &tmU, &full[s], wb * 64, hu, mt * 128, b * 2 + 0);
            ptx::tma_load_4d(st + p.a_bytes, &tmU, &full[s], wb * 64, hu, mt * 128, b * 2 + 1);
            }
            for (int t = 0; t < ntap; ++t) {
              const int tap = tap0 + t, dh = tap / p.kw, dw = tap - dh * p.kw;
              const int vy = hu * p.sh + dh - p.ph;
              uint8_t* bt = st + 2 * p.a_bytes + (size_t)t * 2 * p.b_bytes;
              ptx::tma_load_5d(bt, &tmV, &full[s], wb * 64, vy, nt * p.n_tile, b * 2 + 0, dw);
              ptx::tma_load_5d(bt + p.b_bytes, &tmV, &full[s], wb * 64, vy, nt * p.n_tile, b * 2 + 1, dw);
            }
            if (++s == (uint32_t)p.stages) { s = 0; ph ^= 1u; }
          }
        }
      }
    } else if (warp == 1) {
      if (ptx::elect_one()) {
        const uint32_t idesc = ptx::make_idesc_bf16_f32(128, p.n_tile);
        const uint32_t base32 = (ptx::smem_u32(smem) >> 4) | 0x10000u;
        const uint32_t stage16 = (uint32_t)p.stage_bytes >> 4, a16 = (uint32_t)p.a_bytes >> 4, b16 = (uint32_t)p.b_bytes >> 4;
        uint32_t s = 0, ph = 0, keep = 0;
        for (int64_t kb = 0; kb < nkb; ++kb) {
          ptx::mbar_wait(&full[s], ph);
          ptx::tc_fence_after();
          const uint32_t ah = base32 + s * stage16, al = ah + a16;
          for (int t = 0; t < ntap; ++t) {
            const uint32_t bh = ah + 2 * a16 + (uint32_t)t * 2 * b16, bl = bh + b16;
            const uint32_t d = tmem + (uint32_t)(t * p.n_tile);
            ptx::umma_bf16_lo32(d, al, bh, idesc, keep);
            ptx::umma_bf16_lo32(d, ah, bh, idesc, 1);
            ptx::umma_bf16_lo32(d, ah, bl, idesc, 1);
#pragma unroll
            for (uint32_t k = 2; k < 8; k += 2) {
              ptx::umma_bf16_lo32(d, al + k, bh + k, idesc, 1);
              ptx::umma_bf16_lo32(d, ah + k, bh + k, idesc, 1);
              ptx::umma_bf16_lo32(d, ah + k, bl + k, idesc, 1);
            }
          }
          keep = 1;
          ptx::umma_commit(&empty[s]);
          if (++s == (uint32_t)p.stages) { s = 0; ph ^= 1u; }
        }
        ptx::umma_commit(&done);
      }
    } else {
      // epilogue: wait for the whole K range, then add this CTA's partial sums into G
      const int q = warp & 3;
      ptx::mbar_wait(&done, 0);
      ptx::tc_fence_after();
      const int cu = mt * 128 + q * 32 + lane;
      for (int t = 0; t < ntap; ++t) {
        float* g = p.G + ((int64_t)(tap0 + t) * p.Cu + cu) * p.Cv + nt * p.n_tile;
        for (int c0 = 0; c0 < p.n_tile; c0 += 32) {
          uint32_t r[32];
          ptx::tmem_ld_32x32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(t * p.n_tile + c0), r);
          ptx::tmem_ld_wait();
          if (cu < p.Cu) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (nt * p.n_tile + c0 + j < p.Cv) atomicAdd(g + c0 + j, __uint_as_float(r[j]));
          }
        }
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem, (uint32_t)p.tmem_cols);
}

typedef CUresult (*EncodeTiledFnW)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFnW encode_fn_w() {
  static EncodeTiledFnW fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFnW>(ptr);
  }
  return fn;
}

// NC2 tensor [copies][B*2][C][H][Wp] with logical row length W; box {64 px, 1 row, cbox channels, 1, 1}
static int make_nc2_map(CUtensorMap* tm, const void* base, int64_t B, int Cc, int H, int W, int Wp, int cbox, int copies) {
  EncodeTiledFnW fn = encode_fn_w();
  if (fn == nullptr) return fail(SE_ERR_CUDA, "se_cconv_wgrad: cuTensorMapEncodeTiled unavailable");
  const cuuint32_t rank = copies > 0 ? 5 : 4;
  cuuint64_t dims[5] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)Cc, (cuuint64_t)(B * 2), (cuuint64_t)(copies > 0 ? copies : 1)};
  cuuint64_t str[4] = {(cuuint64_t)Wp * 2, (cuuint64_t)H * Wp * 2, (cuuint64_t)Cc * H * Wp * 2,
                       (cuuint64_t)(B * 2) * Cc * H * Wp * 2};
  cuuint32_t box[5] = {64, 1, (cuuint32_t)cbox, 1, 1};
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), dims, str, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(SE_ERR_CUDA, "se_cconv_wgrad: cuTensorMapEncodeTiled failed (%d)", (int)r);
  return SE_OK;
}

}  // namespace se

extern "C" int64_t se_nc2_elems(int64_t B, int C, int H, int W) {
  const int Wp = (W + 7) / 8 * 8;
  return B * 2 * (int64_t)C * H * Wp;
}

extern "C" int se_nc2_from_nchw(const float* x, int64_t B, int C, int H, int W, void* nc2, void* stream) {
  SE_REQUIRE(x && nc2, "se_nc2_from_nchw: null pointer");
  SE_REQUIRE(B >= 0 && C > 0 && H > 0 && W > 0, "se_nc2_from_nchw: bad dims");
  if (B == 0) return SE_OK;
  const int Wp = (W + 7) / 8 * 8;
  const int64_t rows = B * C * H, total = rows * (Wp / 8);
  const unsigned blocks = (unsigned)std::min<int64_t>((total + 255) / 256, (int64_t)se::num_sms() * 16);
  se::nc2_from_nchw_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, reinterpret_cast<__nv_bfloat16*>(nc2), rows, W, Wp,
                                                                     (int64_t)C * H);
  return se::check_launch("se_nc2_from_nchw");
}

extern "C" int64_t se_nc2_gather_elems(int64_t B, int C, int H, int Wu, int kw) {
  const int Wup = (Wu + 7) / 8 * 8;
  return (int64_t)kw * B * 2 * (int64_t)C * H * Wup;
}

// column-gathered copies of the tap-sampled operand: out[dw][b][s][c][h][wu] = x[b][c][h][wu*sw + dw - pw]
extern "C" int se_nc2_gather_w(const float* x, int64_t B, int C, int H, int W, int Wu, int kw, int sw, int pw, void* out,
                               void* stream) {
  SE_REQUIRE(x && out, "se_nc2_gather_w: null pointer");
  SE_REQUIRE(B >= 0 && C > 0 && H > 0 && W > 0 && Wu > 0 && kw > 0 && sw > 0, "se_nc2_gather_w: bad dims");
  if (B == 0) return SE_OK;
  const int Wup = (Wu + 7) / 8 * 8;
  const int64_t rows = B * C * H, total = rows * (Wup / 8) * kw;
  const unsigned blocks = (unsigned)std::min<int64_t>((total + 255) / 256, (int64_t)se::num_sms() * 16);
  se::nc2_gather_w_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, reinterpret_cast<__nv_bfloat16*>(out), rows, W, Wu, Wup,
                                                                    (int64_t)C * H, kw, sw, pw, B * 2 * (int64_t)C * H * Wup);
  return se::check_launch("se_nc2_gather_w");
}

// G[tap][Cu][Cv] += sum over pixels (G must be zeroed by the caller before the first call).
// u_nc2: se_nc2_from_nchw of U [B,Cu,Hu,Wu]; v_gather: se_nc2_gather_w of V [B,Cv,Hv,Wv] with this (Wu, kw, sw, pw).
extern "C" int se_cconv_wgrad(const void* u_nc2, int Cu, int Hu, int Wu, const void* v_gather, int Cv, int Hv,
                              int64_t B, int kh, int kw, int sh, int ph, float* G, void* stream) {
  SE_REQUIRE(u_nc2 && v_gather && G, "se_cconv_wgrad: null pointer");
  SE_REQUIRE(B >= 0 && Cu > 0 && Cv > 0 && Hu > 0 && Wu > 0 && Hv > 0 && kh > 0 && kw > 0 && sh > 0,
             "se_cconv_wgrad: bad dims");
  if (B == 0) return SE_OK;
  se::WgradParams p{};
  p.Cu = Cu; p.Cv = Cv; p.Hu = Hu; p.Wu = Wu; p.kw = kw; p.sh = sh; p.sw = 1; p.ph = ph; p.pw = 0; p.taps = kh * kw;
  p.B = B; p.G = G;
  p.n_tile = Cv >= 128 ? 128 : (Cv + 15) / 16 * 16;
  p.m_tiles = (Cu + 127) / 128;
  p.n_tiles = (Cv + p.n_tile - 1) / p.n_tile;
  p.a_bytes = 128 * 128;                       // 128 channels x 64 px bf16
  p.b_bytes = p.n_tile * 128;
  const int smem_budget = 200 * 1024;
  int tg = std::min(p.taps, 512 / p.n_tile);
  while (tg > 1 && 2 * (2 * p.a_bytes + 2 * tg * p.b_bytes) > smem_budget) --tg;
  p.tg = tg;
  p.tap_groups = (p.taps + tg - 1) / tg;
  p.stage_bytes = 2 * p.a_bytes + 2 * tg * p.b_bytes;
  p.stages = std::min(4, smem_budget / p.stage_bytes);
  if (p.stages < 2) return se::fail(SE_ERR_UNSUPPORTED, "se_cconv_wgrad: tile does not fit shared memory");
  p.tmem_cols = 32; while (p.tmem_cols < tg * p.n_tile) p.tmem_cols <<= 1;
  p.rows_total = B * Hu;
  const int64_t items = (int64_t)p.m_tiles * p.n_tiles * p.tap_groups;
  // split K so that the grid is a few waves of the machine, but keep >= 8 image rows per CTA
  int64_t ks = std::max<int64_t>(1, (4LL * se::num_sms() + items - 1) / items);
  ks = std::min<int64_t>(ks, std::max<int64_t>(1, p.rows_total / 8));
  ks = std::min<int64_t>(ks, 65535);
  p.rows_per_split = (p.rows_total + ks - 1) / ks;
  p.ksplit = (int)((p.rows_total + p.rows_per_split - 1) / p.rows_per_split);
  const int Wup = (Wu + 7) / 8 * 8;
  CUtensorMap tmU, tmV;
  int rc = se::make_nc2_map(&tmU, u_nc2, B, Cu, Hu, Wu, Wup, 128, 0);
  if (rc != SE_OK) return rc;
  rc = se::make_nc2_map(&tmV, v_gather, B, Cv, Hv, Wu, Wup, p.n_tile, kw);
  if (rc != SE_OK) return rc;
  const int smem = p.stages * p.stage_bytes + 1024;
  static se::PerDevice<int> smem_attr_pd(0);
  int& smem_attr = smem_attr_pd.get();
  if (smem > smem_attr) {
    SE_CUDA(cudaFuncSetAttribute(se::cconv_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    smem_attr = smem;
  }
  if (items > 2147483647LL) return se::fail(SE_ERR_UNSUPPORTED, "se_cconv_wgrad: too many tiles");
  dim3 grid((unsigned)items, (unsigned)p.ksplit);
  se::cconv_wgrad_kernel<<<grid, se::kWgThreads, smem, (cudaStream_t)stream>>>(tmU, tmV, p);
  return se::check_launch("se_cconv_wgrad");
}

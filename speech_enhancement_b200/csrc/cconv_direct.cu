// Direct (CUDA-core) kernels for the degenerate complex-conv layers that are HBM-bound by
// construction: the first encoder layer of every model has 2 input channels
// (K = Cin2*kh*kw <= 20), so its cost is writing the output once.  One thread owns one output
// pixel, keeps <= 32 output channels in registers per pass, reads the packed fp32 weights as
// shared-memory broadcasts and stores either NCHW fp32 (module contract) or CL2 bf16 hi|lo (chain).
// Reference semantics: complex_nn.py:52-78 through the packed operand [[Wr,-Wi],[Wi,Wr]].
#include "engines.cuh"
#include <cuda_bf16.h>

namespace se {

struct SmallCinParams {
  const float* x; int64_t xs_b, xs_c, xs_h;      // input strides in elements (w contiguous)
  int Cin2, H, W, pad_top, pad_left;
  int Cout2, kh, kw, sh, sw, ph, pw, Ho, Wo;
  const float* w;                                 // packed [Cout2][Cin2*kh*kw]
  const float* bias; float slope;
  float* y; __nv_bfloat16* y_cl2; int y_Cp;
};

constexpr int kSmallCinMaxK = 512;

// CW = output channels held in registers per pass: 32 for the encoder layers (and any CL2 store),
// 4 / 8 for the few-channel convs (the CCBAM 7x7 spatial-attention conv has Cout2 = 2), where a
// 32-wide pass would spend 16x the FMAs on zero weights.
template <int CW>
__global__ void __launch_bounds__(128)
cconv_small_cin_kernel(const SmallCinParams p) {
  extern __shared__ float w_s[];                  // [K][Cout2r]  (Cout2 rounded up to CW)
  const int K = p.Cin2 * p.kh * p.kw;
  const int Cr = (p.Cout2 + CW - 1) / CW * CW;
  for (int i = threadIdx.x; i < K * Cr; i += blockDim.x) {
    int k = i / Cr, co = i - k * Cr;
    w_s[i] = co < p.Cout2 ? __ldg(p.w + (int64_t)co * K + k) : 0.f;
  }
  __syncthreads();
  const int wo = blockIdx.x * blockDim.x + threadIdx.x;
  const int ho = blockIdx.y;
  const int64_t b = blockIdx.z;
  if (wo >= p.Wo) return;
  const int64_t plane = (int64_t)p.Ho * p.Wo;
  const int64_t pix = (int64_t)ho * p.Wo + wo;
  const int cmax = p.y_cl2 != nullptr && p.y_Cp > Cr ? p.y_Cp : Cr;
  for (int c0 = 0; c0 < cmax; c0 += CW) {
    float acc[CW];
#pragma unroll
    for (int j = 0; j < CW; ++j) acc[j] = 0.f;
    if (c0 < Cr) {
      // walk the K = Cin2*kh*kw taps; inputs come from L1 after the first pass (zero padding + the
      // caller's causal pad are the bounds checks)
      int k = 0;
      for (int c = 0; c < p.Cin2; ++c)
        for (int dh = 0; dh < p.kh; ++dh) {
          const int h = ho * p.sh - p.ph - p.pad_top + dh;
          const bool h_ok = h >= 0 && h < p.H;
          const float* xr = p.x + b * p.xs_b + c * p.xs_c + (int64_t)h * p.xs_h;
          for (int dw = 0; dw < p.kw; ++dw, ++k) {
            const int w = wo * p.sw - p.pw - p.pad_left + dw;
            const float xk = (h_ok && w >= 0 && w < p.W) ? __ldg(xr + w) : 0.f;
            const float4* wr = reinterpret_cast<const float4*>(w_s + k * Cr + c0);
#pragma unroll
            for (int j4 = 0; j4 < CW / 4; ++j4) {
              const float4 w4 = wr[j4];
              acc[4 * j4 + 0] = fmaf(w4.x, xk, acc[4 * j4 + 0]);
              acc[4 * j4 + 1] = fmaf(w4.y, xk, acc[4 * j4 + 1]);
              acc[4 * j4 + 2] = fmaf(w4.z, xk, acc[4 * j4 + 2]);
              acc[4 * j4 + 3] = fmaf(w4.w, xk, acc[4 * j4 + 3]);
            }
          }
        }
    }
#pragma unroll
    for (int j = 0; j < CW; ++j) {
      const int co = c0 + j;
      float f = acc[j];
      if (co < p.Cout2 && p.bias != nullptr) f += __ldg(p.bias + co);
      acc[j] = co < p.Cout2 ? (f > 0.f ? f : f * p.slope) : 0.f;
    }
    if (p.y != nullptr) {
#pragma unroll
      for (int j = 0; j < CW; ++j)
        if (c0 + j < p.Cout2) p.y[(b * p.Cout2 + c0 + j) * plane + pix] = acc[j];
    }
    if constexpr (CW == 32) if (p.y_cl2 != nullptr && c0 < p.y_Cp) {
      __align__(16) __nv_bfloat16 hi[32];
      __align__(16) __nv_bfloat16 lo[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        hi[j] = __float2bfloat16_rn(acc[j]);
        lo[j] = __float2bfloat16_rn(acc[j] - __bfloat162float(hi[j]));
      }
      __nv_bfloat16* cb = p.y_cl2 + (b * 2 * plane + pix) * p.y_Cp + c0;
      uint4* dh = reinterpret_cast<uint4*>(cb);
      uint4* dl = reinterpret_cast<uint4*>(cb + plane * p.y_Cp);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        dh[j] = reinterpret_cast<const uint4*>(hi)[j];
        dl[j] = reinterpret_cast<const uint4*>(lo)[j];
      }
    }
  }
}

// Compile-time-shaped version for the first encoder layer of the three complex families (2 input channels; 5x2
// stride 2x1 in DCCRN / FRCRN, 3x3 stride 1 in DCUNet): every tap loop unrolls, the K input samples of a pixel are
// loaded once into registers and reused by each 32-channel pass, bias + leaky + the bf16 hi|lo split cost ~5
// instructions per channel (packed cvt.rn.bf16x2), and CL2 pad channels are plain zero stores.  The generic kernel
// above spends ~4x the FMA count in index arithmetic (profiles/r1e_hbm_full_summary.txt: 466 M instructions for
// 110 M FFMAs).
__device__ __forceinline__ void st_global_v8(void* ptr, const uint32_t* v) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(ptr), "r"(v[0]), "r"(v[1]), "r"(v[2]),
               "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
               : "memory");
}

// PX output pixels per thread: every broadcast LDS.128 of four weights (two shared-memory wavefronts) feeds 4 * PX FMAs.
// Measured at 64 x 4 s: PX = 2 wins where the weight reads dominate (Cout2 = 64: 1.15 -> 0.99 ms, 128: 1.14 -> 0.97 ms)
// and loses where the CL2 stores do (Cout2 = 32, half of the 128-byte rows is channel padding: 0.50 -> 0.56 ms).
template <int KH, int KW, int SH, int SW, int PX>
__global__ void __launch_bounds__(128)
cconv_first_kernel(const SmallCinParams p) {
  constexpr int CIN2 = 2, K = CIN2 * KH * KW;
  extern __shared__ float w_s[];                  // [K][Cr] weights, then [Cr] bias
  const int Cr = (p.Cout2 + 31) / 32 * 32;
  float* b_s = w_s + K * Cr;
  for (int i = threadIdx.x; i < K * Cr; i += blockDim.x) {
    int k = i / Cr, co = i - k * Cr;
    w_s[i] = co < p.Cout2 ? __ldg(p.w + (int64_t)co * K + k) : 0.f;
  }
  for (int i = threadIdx.x; i < Cr; i += blockDim.x) b_s[i] = (i < p.Cout2 && p.bias != nullptr) ? __ldg(p.bias + i) : 0.f;
  __syncthreads();
  // PX * 128 consecutive output pixels of one image per block (rows are not padded to the block size); thread t owns
  // pixels base + t, base + 128 + t, ...
  const int64_t plane = (int64_t)p.Ho * p.Wo;
  const int64_t pix0 = (int64_t)blockIdx.x * (blockDim.x * PX) + threadIdx.x;
  const int64_t b = blockIdx.y;
  if (pix0 >= plane) return;

  // Packed FP32 (scripts/ffma2_rate.cu: a 3-register FFMA issues at 73 FMA lanes / clk / SM on B200, fma.rn.f32x2 at
  // 113): two neighbouring output channels share one FFMA2 -- their weights are a natural register pair of the LDS.128,
  // the input sample is kept as a (x, x) pair.  Per-lane arithmetic and order are those of the scalar loop.
  float2 xk[PX][K];
  bool ok[PX];
#pragma unroll
  for (int i = 0; i < PX; ++i) {
    const int64_t pix = pix0 + (int64_t)i * blockDim.x;
    ok[i] = pix < plane;
    const int64_t pc = ok[i] ? pix : pix0;
    const int ho = (int)(pc / p.Wo), wo = (int)(pc - (int64_t)ho * p.Wo);
    const int h0 = ho * SH - p.ph - p.pad_top, w0 = wo * SW - p.pw - p.pad_left;
    const float* xb = p.x + b * p.xs_b;
#pragma unroll
    for (int c = 0; c < CIN2; ++c)
#pragma unroll
      for (int dh = 0; dh < KH; ++dh) {
        const int h = h0 + dh;
        const bool h_ok = h >= 0 && h < p.H;
        const float* xr = xb + c * p.xs_c + (int64_t)h * p.xs_h + w0;
#pragma unroll
        for (int dw = 0; dw < KW; ++dw) {
          const int w = w0 + dw;
          const float v = (h_ok && w >= 0 && w < p.W) ? __ldg(xr + dw) : 0.f;
          xk[i][(c * KH + dh) * KW + dw] = make_float2(v, v);
        }
      }
  }
  const float slope = p.slope;
  for (int c0 = 0; c0 < Cr; c0 += 32) {
    float2 acc2[PX][16];
    {
      const float4* br = reinterpret_cast<const float4*>(b_s + c0);
#pragma unroll
      for (int j4 = 0; j4 < 8; ++j4) {
        const float4 b4 = br[j4];
#pragma unroll
        for (int i = 0; i < PX; ++i) {
          acc2[i][2 * j4] = make_float2(b4.x, b4.y);
          acc2[i][2 * j4 + 1] = make_float2(b4.z, b4.w);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const float4* wr = reinterpret_cast<const float4*>(w_s + k * Cr + c0);
#pragma unroll
      for (int j4 = 0; j4 < 8; ++j4) {
        const float4 w4 = wr[j4];
#pragma unroll
        for (int i = 0; i < PX; ++i) {
          acc2[i][2 * j4] = __ffma2_rn(make_float2(w4.x, w4.y), xk[i][k], acc2[i][2 * j4]);
          acc2[i][2 * j4 + 1] = __ffma2_rn(make_float2(w4.z, w4.w), xk[i][k], acc2[i][2 * j4 + 1]);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < PX; ++i) {
      if (!ok[i]) continue;
      const int64_t pix = pix0 + (int64_t)i * blockDim.x;
      float acc_i[32];
#pragma unroll
      for (int j = 0; j < 16; ++j) { acc_i[2 * j] = acc2[i][j].x; acc_i[2 * j + 1] = acc2[i][j].y; }
#pragma unroll
      for (int j = 0; j < 32; ++j) acc_i[j] = acc_i[j] > 0.f ? acc_i[j] : acc_i[j] * slope;
      if (p.y != nullptr) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (c0 + j < p.Cout2) p.y[(b * p.Cout2 + c0 + j) * plane + pix] = acc_i[j];
      }
      if (p.y_cl2 != nullptr) {
        // channels >= Cout2 inside this pass carry zero weights and zero bias: they come out as exact zeros
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const __nv_bfloat162 h2 = __floats2bfloat162_rn(acc_i[2 * j], acc_i[2 * j + 1]);
          const float2 hf = __bfloat1622float2(h2);
          const __nv_bfloat162 l2 = __floats2bfloat162_rn(acc_i[2 * j] - hf.x, acc_i[2 * j + 1] - hf.y);
          hi[j] = *reinterpret_cast<const uint32_t*>(&h2);
          lo[j] = *reinterpret_cast<const uint32_t*>(&l2);
        }
        // 32-byte stores: one full sector per lane and instruction (lanes are a pixel = 128+ bytes apart)
        __nv_bfloat16* dh = p.y_cl2 + (b * 2 * plane + pix) * p.y_Cp + c0;
        __nv_bfloat16* dl = dh + plane * p.y_Cp;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          st_global_v8(dh + 16 * j, hi + 8 * j);
          st_global_v8(dl + 16 * j, lo + 8 * j);
        }
      }
    }
  }
  if (p.y_cl2 != nullptr) {                        // CL2 pad channels [Cr, y_Cp)
    const uint32_t z[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
#pragma unroll
    for (int i = 0; i < PX; ++i) {
      if (!ok[i]) continue;
      __nv_bfloat16* cb = p.y_cl2 + (b * 2 * plane + pix0 + (int64_t)i * blockDim.x) * p.y_Cp;
      for (int c0 = Cr; c0 < p.y_Cp; c0 += 16) {     // Cr and y_Cp are multiples of 32
        st_global_v8(cb + c0, z);
        st_global_v8(cb + plane * p.y_Cp + c0, z);
      }
    }
  }
}

template <int KH, int KW, int SH, int SW>
static int launch_first(const SmallCinParams& p, int64_t B, cudaStream_t s) {
  const int Cr = (p.Cout2 + 31) / 32 * 32;
  const int smem = (2 * KH * KW * Cr + Cr) * 4;
  const int px = Cr > 32 ? 2 : 1;
  dim3 grid((unsigned)(((int64_t)p.Ho * p.Wo + 128 * px - 1) / (128 * px)), (unsigned)B);
  if (px == 2) cconv_first_kernel<KH, KW, SH, SW, 2><<<grid, 128, smem, s>>>(p);
  else cconv_first_kernel<KH, KW, SH, SW, 1><<<grid, 128, smem, s>>>(p);
  return check_launch("se_cconv_small_cin(first)");
}

}  // namespace se

extern "C" int se_cconv_small_cin(const float* x, int64_t xs_b, int64_t xs_c, int64_t xs_h, int64_t B, int Cin2,
                                  int H, int W, int pad_top, int pad_left, const float* packed, const float* bias,
                                  float slope, int Cout2, int kh, int kw, int sh, int sw, int ph, int pw,
                                  float* y_nchw, void* y_cl2, int y_Cp, int Ho, int Wo, void* stream) {
  SE_REQUIRE(x && packed && (y_nchw || y_cl2), "se_cconv_small_cin: null pointer");
  SE_REQUIRE(B >= 0 && Cin2 > 0 && Cout2 > 0 && H > 0 && W > 0 && kh > 0 && kw > 0 && sh > 0 && sw > 0 &&
             Ho > 0 && Wo > 0 && pad_top >= 0 && pad_left >= 0, "se_cconv_small_cin: bad geometry");
  SE_REQUIRE(Ho == (H + pad_top + 2 * ph - kh) / sh + 1 && Wo == (W + pad_left + 2 * pw - kw) / sw + 1,
             "se_cconv_small_cin: output size does not match geometry");
  SE_REQUIRE(y_cl2 == nullptr || (y_Cp >= Cout2 && y_Cp % 64 == 0), "se_cconv_small_cin: bad y_Cp");
  const int CW = (y_cl2 != nullptr || Cout2 > 8) ? 32 : (Cout2 > 4 ? 8 : 4);
  const int K = Cin2 * kh * kw, Cr = (Cout2 + CW - 1) / CW * CW;
  if (K > se::kSmallCinMaxK || (size_t)K * Cr * 4 > 160 * 1024)
    return se::fail(SE_ERR_UNSUPPORTED, "se_cconv_small_cin: needs Cin2*kh*kw <= %d", se::kSmallCinMaxK);
  if (B == 0) return SE_OK;
  if (Ho > 65535 || B > 65535) return se::fail(SE_ERR_UNSUPPORTED, "se_cconv_small_cin: tensor too large");
  se::SmallCinParams p{x, xs_b, xs_c, xs_h, Cin2, H, W, pad_top, pad_left, Cout2, kh, kw, sh, sw, ph, pw, Ho, Wo,
                       packed, bias, slope, y_nchw, reinterpret_cast<__nv_bfloat16*>(y_cl2), y_Cp};
  // the first encoder layers of DCCRN / FRCRN (5x2, stride 2x1) and DCUNet (3x3, stride 1): compile-time-shaped kernel
  if (Cin2 == 2 && Cout2 >= 16 && Cout2 <= 256) {
    if (kh == 5 && kw == 2 && sh == 2 && sw == 1) return se::launch_first<5, 2, 2, 1>(p, B, (cudaStream_t)stream);
    if (kh == 3 && kw == 3 && sh == 1 && sw == 1) return se::launch_first<3, 3, 1, 1>(p, B, (cudaStream_t)stream);
  }
  const int smem = K * Cr * 4;
  static se::PerDevice<int> smem_attr_pd(48 * 1024);
  int& smem_attr = smem_attr_pd.get();
  if (smem > smem_attr) {                         // only the 32-wide pass can exceed the default
    SE_CUDA(cudaFuncSetAttribute(se::cconv_small_cin_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    smem_attr = smem;
  }
  dim3 grid((unsigned)((Wo + 127) / 128), (unsigned)Ho, (unsigned)B);
  if (CW == 32) se::cconv_small_cin_kernel<32><<<grid, 128, smem, (cudaStream_t)stream>>>(p);
  else if (CW == 8) se::cconv_small_cin_kernel<8><<<grid, 128, smem, (cudaStream_t)stream>>>(p);
  else se::cconv_small_cin_kernel<4><<<grid, 128, smem, (cudaStream_t)stream>>>(p);
  return se::check_launch("se_cconv_small_cin");
}

// ---- last decoder layer: Cout2 <= 4 (M = 2 in SURVEY.md 8a), cost = reading the input once ----
namespace se {

struct SmallCoutParams {
  const float* x; const float* w; const float* bias; float* y;
  int transposed, Cin2, H, W, Cout2, kh, kw, sh, sw, ph, pw, Ho, Wo;
  float slope;
};

// one thread per output pixel; weights in shared memory as float4 per (tap, cin): [taps][Cin2][4]
__global__ void __launch_bounds__(128)
cconv_small_cout_kernel(const SmallCoutParams p) {
  extern __shared__ float4 w4_s[];
  const int taps = p.kh * p.kw;
  for (int i = threadIdx.x; i < taps * p.Cin2; i += blockDim.x) {
    const int tap = i / p.Cin2, c = i - tap * p.Cin2;
    float v[4];
#pragma unroll
    for (int co = 0; co < 4; ++co)
      v[co] = co < p.Cout2 ? __ldg(p.w + ((int64_t)co * p.Cin2 + c) * taps + tap) : 0.f;
    w4_s[i] = make_float4(v[0], v[1], v[2], v[3]);
  }
  __syncthreads();
  const int wo = blockIdx.x * blockDim.x + threadIdx.x;
  const int ho = blockIdx.y;
  const int64_t b = blockIdx.z;
  if (wo >= p.Wo) return;
  const int64_t in_plane = (int64_t)p.H * p.W;
  const float* xb = p.x + b * p.Cin2 * in_plane;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int dh = 0; dh < p.kh; ++dh) {
    int hi;
    if (p.transposed) {
      const int nh = ho + p.ph - dh;
      if (nh < 0 || nh % p.sh != 0) continue;
      hi = nh / p.sh;
    } else {
      hi = ho * p.sh - p.ph + dh;
    }
    if (hi < 0 || hi >= p.H) continue;
    for (int dw = 0; dw < p.kw; ++dw) {
      int wi;
      if (p.transposed) {
        const int nw = wo + p.pw - dw;
        if (nw < 0 || nw % p.sw != 0) continue;
        wi = nw / p.sw;
      } else {
        wi = wo * p.sw - p.pw + dw;
      }
      if (wi < 0 || wi >= p.W) continue;
      const float* xp = xb + (int64_t)hi * p.W + wi;
      const float4* wp = w4_s + (dh * p.kw + dw) * p.Cin2;
#pragma unroll 8
      for (int c = 0; c < p.Cin2; ++c) {
        const float xv = __ldg(xp + c * in_plane);
        const float4 w = wp[c];
        acc.x = fmaf(w.x, xv, acc.x); acc.y = fmaf(w.y, xv, acc.y);
        acc.z = fmaf(w.z, xv, acc.z); acc.w = fmaf(w.w, xv, acc.w);
      }
    }
  }
  const float a4[4] = {acc.x, acc.y, acc.z, acc.w};
  const int64_t plane = (int64_t)p.Ho * p.Wo;
#pragma unroll
  for (int co = 0; co < 4; ++co) {
    if (co >= p.Cout2) break;
    float f = a4[co] + (p.bias ? __ldg(p.bias + co) : 0.f);
    p.y[(b * p.Cout2 + co) * plane + (int64_t)ho * p.Wo + wo] = f > 0.f ? f : f * p.slope;
  }
}

int cconv_fwd_small_cout(bool transposed, const float* x, const float* packed, const float* bias, int64_t B,
                         const ConvGeom& g, float* y, cudaStream_t s) {
  const int smem = g.kh * g.kw * g.Cin2 * 16;
  if (g.Cout2 > 4 || smem > 200 * 1024)
    return fail(SE_ERR_UNSUPPORTED, "small-Cout direct kernel: needs Cout2 <= 4 and kh*kw*Cin2 <= 12800");
  if (g.Ho > 65535 || B > 65535) return fail(SE_ERR_UNSUPPORTED, "small-Cout direct kernel: tensor too large");
  SmallCoutParams p{x, packed, bias, y, transposed ? 1 : 0, g.Cin2, g.H, g.W, g.Cout2, g.kh, g.kw, g.sh, g.sw,
                    g.ph, g.pw, g.Ho, g.Wo, 1.f};
  static se::PerDevice<int> smem_attr_pd(48 * 1024);
  int& smem_attr = smem_attr_pd.get();
  if (smem > smem_attr) {
    SE_CUDA(cudaFuncSetAttribute(cconv_small_cout_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    smem_attr = smem;
  }
  dim3 grid((unsigned)((g.Wo + 127) / 128), (unsigned)g.Ho, (unsigned)B);
  cconv_small_cout_kernel<<<grid, 128, smem, s>>>(p);
  return check_launch("cconv_small_cout_kernel");
}

}  // namespace se


// ------------------------------------------------------------------------------------------------
// Scatter form of a transposed convolution with very few output channels (the last decoder layer of
// DCCRN / DCUNet: Cout2 = 2; /root/reference/models/_2008_00264_dccrn.py:104-111,
// _1903_03107_dcunet.py:79-81).  On the tensor cores such a layer costs one >= 84-cycle MMA per tap and
// k-step for a handful of useful columns; instead the taps are folded into the GEMM's N dimension
// (a 1x1 convolution with taps*Cout2 output columns, one pass over the input, no shifted re-reads):
//   P[b, hi, wi, tap*Cout2 + co] = sum_c x[b, c, hi, wi] * w[c, co, tap]
// and this kernel gathers the partial products that land on every output pixel:
//   y[b, co, ho, wo] = act(bias[co] + sum_{dh,dw : sh | ho+ph-dh, sw | wo+pw-dw} P[b, (ho+ph-dh)/sh, (wo+pw-dw)/sw, ...])
namespace se {

struct Col2imParams {
  const float* P; float* y; const float* bias;
  int64_t B; int H, W, Np, Cout2, kh, kw, sh, sw, ph, pw, Ho, Wo;
  float slope;
};

template <int CO>
__global__ void __launch_bounds__(256)
col2im_small_kernel(const Col2imParams p) {
  const int64_t total = p.B * p.Ho * (int64_t)p.Wo;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int wo = (int)(i % p.Wo);
    const int64_t r = i / p.Wo;
    const int ho = (int)(r % p.Ho);
    const int64_t b = r / p.Ho;
    float acc[CO];
#pragma unroll
    for (int c = 0; c < CO; ++c) acc[c] = p.bias != nullptr ? __ldg(p.bias + c) : 0.f;
    for (int dh = 0; dh < p.kh; ++dh) {
      const int nh = ho + p.ph - dh;
      if (nh < 0 || nh % p.sh != 0) continue;
      const int hi = nh / p.sh;
      if (hi >= p.H) continue;
      for (int dw = 0; dw < p.kw; ++dw) {
        const int nw = wo + p.pw - dw;
        if (nw < 0 || nw % p.sw != 0) continue;
        const int wi = nw / p.sw;
        if (wi >= p.W) continue;
        const float* src = p.P + ((b * p.H + hi) * p.W + wi) * p.Np + (dh * p.kw + dw) * CO;
        if (CO == 2) {
          const float2 v = __ldg(reinterpret_cast<const float2*>(src));
          acc[0] += v.x; acc[1] += v.y;
        } else {
#pragma unroll
          for (int c = 0; c < CO; ++c) acc[c] += __ldg(src + c);
        }
      }
    }
    const int64_t plane = (int64_t)p.Ho * p.Wo;
#pragma unroll
    for (int c = 0; c < CO; ++c) {
      const float v = acc[c];
      p.y[(b * CO + c) * plane + (int64_t)ho * p.Wo + wo] = v > 0.f ? v : v * p.slope;
    }
  }
}

}  // namespace se

extern "C" int se_cconvT_col2im(const float* P, int64_t B, int H, int W, int Np, int Cout2, int kh, int kw, int sh,
                                int sw, int ph, int pw, int Ho, int Wo, const float* bias, float slope, float* y,
                                void* stream) {
  SE_REQUIRE(P && y, "se_cconvT_col2im: null pointer");
  SE_REQUIRE(B >= 0 && H > 0 && W > 0 && kh > 0 && kw > 0 && sh > 0 && sw > 0 && Ho > 0 && Wo > 0 &&
             Np >= kh * kw * Cout2 && Np % 2 == 0, "se_cconvT_col2im: bad dims");
  if (Cout2 != 2 && Cout2 != 4) return se::fail(SE_ERR_UNSUPPORTED, "se_cconvT_col2im: Cout2 must be 2 or 4");
  if (B == 0) return SE_OK;
  se::Col2imParams p{P, y, bias, B, H, W, Np, Cout2, kh, kw, sh, sw, ph, pw, Ho, Wo, slope};
  const int64_t total = B * Ho * (int64_t)Wo;
  const unsigned blocks = (unsigned)std::min<int64_t>((total + 255) / 256, (int64_t)se::num_sms() * 32);
  if (Cout2 == 2) se::col2im_small_kernel<2><<<blocks, 256, 0, (cudaStream_t)stream>>>(p);
  else se::col2im_small_kernel<4><<<blocks, 256, 0, (cudaStream_t)stream>>>(p);
  return se::check_launch("se_cconvT_col2im");
}

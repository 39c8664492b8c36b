// tcgen05 / TMEM / TMA engine for ComplexConv2d and ComplexConvTranspose2d
// (reference semantics: /root/reference/models/modules/complex_nn.py:52-91).
//
// One real implicit GEMM per layer against the packed operand [[Wr,-Wi],[Wi,Wr]]:
//     D[pixel, cout2] = sum_{tap} sum_{cin2} X[pixel + tap, cin2] * Wp[tap][cout2, cin2]
// fp32 accuracy on bf16 tensor cores through a 3-term split: x = x_hi + x_lo (both bf16),
//     x*w ~= x_hi*w_hi + x_hi*w_lo + x_lo*w_hi          (fp32 accumulate in TMEM)
// which keeps ~16 mantissa bits per operand (measured 4e-5..8e-5 end-to-end waveform rel-L2).
//
// Data flow per layer
//   1. prep kernel: NCHW fp32 activations -> channels-last bf16 hi|lo planes, de-interleaved by
//      stride phase ("polyphase"), so every kernel tap of a strided conv is a DENSE box:
//         xs[(b*2 + hi/lo)*P + phase][Hq][Wq][Cp]     P = sh*sw (conv) or 1 (transposed)
//      Zero padding of the conv = TMA out-of-bounds zero fill (negative / overflowing coords).
//   2. weight prep: packed fp32 [Cout2][Cin2][taps] -> wq[tap*2 + hi/lo][Npad][Cp] bf16.
//   3. GEMM kernel, one CTA per 128-pixel output tile (TH x TW), N = Cout2 padded to 16:
//      warp 0 lane: TMA producer (A = 128 pixels x 64 ch, B = Npad x 64 ch, hi and lo, 128B swizzle)
//      warp 1 lane: tcgen05.mma issuer, 4 K-steps x 3 split terms per stage, accumulator in TMEM
//      all 4 warps: epilogue tcgen05.ld -> +bias -> coalesced NCHW fp32 stores.
//   A transposed conv is run as sh*sw stride-1 sub-convolutions (output parity classes), so no
//   multiply by inserted zeros is ever issued.
#include "engines.cuh"
#include "umma.cuh"
#include <cuda_bf16.h>

namespace se {

constexpr int kMaxTaps = 49;
constexpr int kKBlock = 64;                 // bf16 elements per 128-byte swizzle row
constexpr int kTileM = 128;                 // pixels per CTA tile = UMMA M
constexpr int kABytes = kTileM * 128;       // one A (hi or lo) stage tile

struct Tap { int16_t dh, dw, nsel, wslab; };

struct UmmaConvParams {
  int ntaps, kblocks, P;
  int TH, TW, tiles_w, tiles_h;
  int Hq_out, Wq_out;
  int oh_mul, oh_add, ow_mul, ow_add;
  int Ho, Wo, Cout2, npad, nstages, tmem_cols;
  const float* bias;
  float* y;
  Tap taps[kMaxTaps];
};

// ------------------------------------------------------------------------------ prep kernels
__device__ __forceinline__ void split_bf16(float v, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(v);
  lo = __float2bfloat16_rn(v - __bfloat162float(hi));
}

// grid: (ceil(Wq/32), Hq, B*P*(Cp/64)); block 256
__global__ void __launch_bounds__(256)
cconv_prep_x_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ xs, int C, int H, int W,
                    int Cp, int Hq, int Wq, int sh, int sw, int P) {
  __shared__ __align__(16) __nv_bfloat16 s_hi[32][kKBlock + 8];
  __shared__ __align__(16) __nv_bfloat16 s_lo[32][kKBlock + 8];
  const int cblocks = Cp / kKBlock;
  int z = blockIdx.z;
  const int cb = z % cblocks; z /= cblocks;
  const int p = z % P;
  const int b = z / P;
  const int hq = blockIdx.y, wq0 = blockIdx.x * 32;
  const int rh = p / sw, rw = p % sw;
  const int hi_src = hq * sh + rh;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int wi_src = (wq0 + tx) * sw + rw;
  const bool pix_ok = hi_src < H && wi_src < W && (wq0 + tx) < Wq;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int cl = ty + 8 * i, c = cb * kKBlock + cl;
    float v = (pix_ok && c < C) ? __ldg(x + (((int64_t)b * C + c) * H + hi_src) * W + wi_src) : 0.f;
    split_bf16(v, s_hi[tx][cl], s_lo[tx][cl]);
  }
  __syncthreads();
  const int pix = threadIdx.x >> 3, chunk = threadIdx.x & 7;
  if (wq0 + pix < Wq) {
    const int64_t plane = (int64_t)Hq * Wq * Cp;
    int64_t off = ((int64_t)hq * Wq + wq0 + pix) * Cp + cb * kKBlock + chunk * 8;
    int64_t n_hi = ((int64_t)b * 2 + 0) * P + p, n_lo = ((int64_t)b * 2 + 1) * P + p;
    *reinterpret_cast<uint4*>(xs + n_hi * plane + off) = *reinterpret_cast<const uint4*>(&s_hi[pix][chunk * 8]);
    *reinterpret_cast<uint4*>(xs + n_lo * plane + off) = *reinterpret_cast<const uint4*>(&s_lo[pix][chunk * 8]);
  }
}

// packed [Cout2][Cin2][taps] fp32 -> wq[(tap*2 + s)][Npad][Cp] bf16
__global__ void cconv_prep_w_kernel(const float* __restrict__ packed, __nv_bfloat16* __restrict__ wq,
                                    int Cout2, int Cin2, int taps, int npad, int Cp) {
  int64_t total = (int64_t)taps * npad * Cp;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  int c = (int)(i % Cp); int64_t r = i / Cp;
  int co = (int)(r % npad); int tap = (int)(r / npad);
  float v = (co < Cout2 && c < Cin2) ? __ldg(packed + ((int64_t)co * Cin2 + c) * taps + tap) : 0.f;
  __nv_bfloat16 hi, lo;
  split_bf16(v, hi, lo);
  int64_t slab = (int64_t)npad * Cp;
  wq[((int64_t)tap * 2 + 0) * slab + (int64_t)co * Cp + c] = hi;
  wq[((int64_t)tap * 2 + 1) * slab + (int64_t)co * Cp + c] = lo;
}

// ------------------------------------------------------------------------------ GEMM kernel
__global__ void __launch_bounds__(128)
cconv_umma_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                  const __grid_constant__ UmmaConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[8];
  __shared__ __align__(8) uint64_t empty_bar[8];
  __shared__ __align__(8) uint64_t accum_bar;
  __shared__ uint32_t tmem_base_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t b_bytes = (uint32_t)p.npad * 128;
  const uint32_t stage_bytes = 2 * kABytes + 2 * b_bytes;
  const int NS = p.nstages;

  const int tile = blockIdx.x, b = blockIdx.y;
  const int th = tile / p.tiles_w, tw = tile - th * p.tiles_w;
  const int h0 = th * p.TH, w0 = tw * p.TW;
  const int nit = p.ntaps * p.kblocks;

  if (threadIdx.x == 0) {
    ptx::prefetch_tmap(&tmX);
    ptx::prefetch_tmap(&tmW);
    for (int s = 0; s < NS; ++s) { ptx::mbar_init(&full_bar[s], 1); ptx::mbar_init(&empty_bar[s], 1); }
    ptx::mbar_init(&accum_bar, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(&tmem_base_slot, (uint32_t)p.tmem_cols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_acc = tmem_base_slot;

  if (warp == 0) {
    if (ptx::elect_one()) {
      // ===== TMA producer =====
      for (int it = 0; it < nit; ++it) {
        const int s = it % NS; const uint32_t ph = (uint32_t)(it / NS) & 1u;
        ptx::mbar_wait(&empty_bar[s], ph ^ 1u);
        const int t = it / p.kblocks, kb = it - t * p.kblocks;
        const Tap tp = p.taps[t];
        uint8_t* st = smem + (size_t)s * stage_bytes;
        ptx::mbar_arrive_expect_tx(&full_bar[s], stage_bytes);
        const int c0 = kb * kKBlock, wc = w0 + tp.dw, hc = h0 + tp.dh;
        ptx::tma_load_4d(st, &tmX, &full_bar[s], c0, wc, hc, (b * 2 + 0) * p.P + tp.nsel);
        ptx::tma_load_4d(st + kABytes, &tmX, &full_bar[s], c0, wc, hc, (b * 2 + 1) * p.P + tp.nsel);
        ptx::tma_load_3d(st + 2 * kABytes, &tmW, &full_bar[s], c0, 0, tp.wslab * 2 + 0);
        ptx::tma_load_3d(st + 2 * kABytes + b_bytes, &tmW, &full_bar[s], c0, 0, tp.wslab * 2 + 1);
      }
    }
  } else if (warp == 1) {
    if (ptx::elect_one()) {
      // ===== MMA issuer =====
      const uint32_t idesc = ptx::make_idesc_bf16_f32(kTileM, p.npad);
      for (int it = 0; it < nit; ++it) {
        const int s = it % NS; const uint32_t ph = (uint32_t)(it / NS) & 1u;
        ptx::mbar_wait(&full_bar[s], ph);
        ptx::tc_fence_after();
        const uint32_t sa = ptx::smem_u32(smem + (size_t)s * stage_bytes);
        const uint64_t a_hi = ptx::make_desc_k_sw128(sa), a_lo = ptx::make_desc_k_sw128(sa + kABytes);
        const uint64_t b_hi = ptx::make_desc_k_sw128(sa + 2 * kABytes);
        const uint64_t b_lo = ptx::make_desc_k_sw128(sa + 2 * kABytes + b_bytes);
#pragma unroll
        for (int k = 0; k < kKBlock / 16; ++k) {
          const uint64_t ko = (uint64_t)(k * 32 >> 4);      // 16 bf16 = 32 B along the swizzled row
          ptx::umma_bf16(tmem_acc, a_lo + ko, b_hi + ko, idesc, (it | k) != 0);
          ptx::umma_bf16(tmem_acc, a_hi + ko, b_lo + ko, idesc, 1);
          ptx::umma_bf16(tmem_acc, a_hi + ko, b_hi + ko, idesc, 1);
        }
        ptx::umma_commit(&empty_bar[s]);        // frees the smem stage once these MMAs have read it
      }
      ptx::umma_commit(&accum_bar);             // accumulator complete
    }
  }
  __syncwarp();

  // ===== epilogue: TMEM -> registers -> (+bias) -> NCHW fp32 =====
  if (nit > 0) {
    ptx::mbar_wait(&accum_bar, 0);
    ptx::tc_fence_after();
  }
  const int m = warp * 32 + lane;
  const int hl = m / p.TW, wl = m - hl * p.TW;
  const int qh = h0 + hl, qw = w0 + wl;
  const int ho = qh * p.oh_mul + p.oh_add, wo = qw * p.ow_mul + p.ow_add;
  const bool ok = qh < p.Hq_out && qw < p.Wq_out && ho < p.Ho && wo < p.Wo;
  const int64_t plane = (int64_t)p.Ho * p.Wo;
  float* yb = p.y + (int64_t)b * p.Cout2 * plane + (int64_t)ho * p.Wo + wo;
  for (int c0 = 0; c0 < p.npad; c0 += 32) {
    uint32_t r[32];
    if (nit > 0) {
      ptx::tmem_ld_32x32(tmem_acc + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
      ptx::tmem_ld_wait();
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) r[j] = 0u;
    }
    if (ok) {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const int co = c0 + j;
        if (co < p.Cout2) {
          float v = __uint_as_float(r[j]);
          if (p.bias != nullptr) v += __ldg(p.bias + co);
          yb[(int64_t)co * plane] = v;
        }
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_acc, (uint32_t)p.tmem_cols);
}

}  // namespace se

// =================================================================================== host side
namespace se {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 tensor, dims innermost-first, dense row-major strides, 128-byte swizzle, zero OOB fill
static int make_tmap_bf16(CUtensorMap* m, void* base, int rank, const uint64_t* dims, const uint32_t* box) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr) return fail(SE_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t gdim[5], gstride[4];
  cuuint32_t bdim[5], estr[5];
  uint64_t stride = 2;
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i]; bdim[i] = box[i]; estr[i] = 1;
    stride *= dims[i];
    if (i < rank - 1) gstride[i] = stride;
  }
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, base, gdim, gstride, bdim, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(SE_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return SE_OK;
}

static inline int round_up(int a, int b) { return (a + b - 1) / b * b; }
static inline int64_t round_up64(int64_t a, int64_t b) { return (a + b - 1) / b * b; }
static inline int pymod(int a, int b) { int r = a % b; return r < 0 ? r + b : r; }

struct UmmaPlan {
  int Cp, npad, P, Hq, Wq, taps;
  int64_t xs_bytes, wq_bytes;
};

static UmmaPlan make_plan(bool transposed, int64_t B, int Cin2, int H, int W, int Cout2, int kh, int kw,
                          int sh, int sw) {
  UmmaPlan pl;
  pl.Cp = round_up(Cin2, kKBlock);
  pl.npad = round_up(Cout2, 16);
  pl.taps = kh * kw;
  if (transposed) { pl.P = 1; pl.Hq = H; pl.Wq = W; }
  else { pl.P = sh * sw; pl.Hq = (H + sh - 1) / sh; pl.Wq = (W + sw - 1) / sw; }
  pl.xs_bytes = round_up64(B * 2 * pl.P * (int64_t)pl.Hq * pl.Wq * pl.Cp * 2, 1024);
  pl.wq_bytes = round_up64((int64_t)pl.taps * 2 * pl.npad * pl.Cp * 2, 1024);
  return pl;
}

// hard limits of the kernel
static bool cconv_umma_can_run(int64_t B, const ConvGeom& g) {
  if (g.Cout2 > 256 || g.kh * g.kw > kMaxTaps) return false;
  if (B > 65535 || B * 2 * g.sh * g.sw > 2000000000LL) return false;
  return true;
}

// SE_ENGINE_AUTO heuristic: layers whose 64-channel K-blocks would be mostly padding stay on SIMT
bool cconv_umma_supported(bool transposed, int64_t B, const ConvGeom& g) {
  (void)transposed;
  return cconv_umma_can_run(B, g) && g.Cin2 >= 16;
}

int64_t cconv_umma_workspace_bytes(bool transposed, int64_t B, int Cin2, int H, int W, int Cout2, int kh,
                                   int kw, int sh, int sw) {
  UmmaPlan pl = make_plan(transposed, B, Cin2, H, W, Cout2, kh, kw, sh, sw);
  return pl.xs_bytes + pl.wq_bytes + 1024;
}

static void pick_tile(int hq, int wq, int& TH, int& TW) {
  int64_t best = -1;
  for (int tw = 128; tw >= 8; tw >>= 1) {
    int th = kTileM / tw;
    int64_t cost = (int64_t)round_up(hq, th) * round_up(wq, tw);
    // prefer wide tiles (coalesced epilogue) unless a narrower one saves >= 3 % of the padded work
    if (best < 0 || cost * 100 < best * 97) { best = cost; TH = th; TW = tw; }
  }
}

int cconv_fwd_umma(bool transposed, const float* x, const float* packed, const float* bias, int64_t B,
                   const ConvGeom& g, float* y, void* ws, int64_t ws_bytes, cudaStream_t s) {
  if (!cconv_umma_can_run(B, g))
    return fail(SE_ERR_UNSUPPORTED, "se_cconv*_fwd: shape not supported by the tcgen05 engine "
                "(needs Cout2 <= 256, kh*kw <= %d)", kMaxTaps);
  UmmaPlan pl = make_plan(transposed, B, g.Cin2, g.H, g.W, g.Cout2, g.kh, g.kw, g.sh, g.sw);
  uintptr_t base = (reinterpret_cast<uintptr_t>(ws) + 1023) & ~uintptr_t(1023);
  if (ws == nullptr || (int64_t)(base - reinterpret_cast<uintptr_t>(ws)) + pl.xs_bytes + pl.wq_bytes > ws_bytes)
    return fail(SE_ERR_WORKSPACE, "se_cconv*_fwd: workspace of %lld bytes needed (se_cconv_workspace_bytes)",
                (long long)(pl.xs_bytes + pl.wq_bytes + 1024));
  __nv_bfloat16* xs = reinterpret_cast<__nv_bfloat16*>(base);
  __nv_bfloat16* wq = reinterpret_cast<__nv_bfloat16*>(base + pl.xs_bytes);

  // 1. operand preparation
  {
    int sh = transposed ? 1 : g.sh, sw = transposed ? 1 : g.sw;
    dim3 grid((unsigned)((pl.Wq + 31) / 32), (unsigned)pl.Hq, (unsigned)(B * pl.P * (pl.Cp / kKBlock)));
    if (grid.y > 65535 || grid.z > 65535)
      return fail(SE_ERR_UNSUPPORTED, "se_cconv*_fwd: activation too large for the prep grid");
    cconv_prep_x_kernel<<<grid, 256, 0, s>>>(x, xs, g.Cin2, g.H, g.W, pl.Cp, pl.Hq, pl.Wq, sh, sw, pl.P);
    int rc = check_launch("cconv_prep_x_kernel");
    if (rc != SE_OK) return rc;
    int64_t total = (int64_t)pl.taps * pl.npad * pl.Cp;
    cconv_prep_w_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(packed, wq, g.Cout2, g.Cin2, pl.taps,
                                                                       pl.npad, pl.Cp);
    rc = check_launch("cconv_prep_w_kernel");
    if (rc != SE_OK) return rc;
  }

  // 2. tensor maps
  CUtensorMap tmW;
  {
    uint64_t dims[3] = {(uint64_t)pl.Cp, (uint64_t)pl.npad, (uint64_t)pl.taps * 2};
    uint32_t box[3] = {(uint32_t)kKBlock, (uint32_t)pl.npad, 1};
    int rc = make_tmap_bf16(&tmW, wq, 3, dims, box);
    if (rc != SE_OK) return rc;
  }

  // 3. one GEMM launch per output parity class (a single class for the forward conv)
  const int nch = transposed ? g.sh : 1, ncw = transposed ? g.sw : 1;
  for (int ch = 0; ch < nch; ++ch) {
    for (int cw = 0; cw < ncw; ++cw) {
      UmmaConvParams p;
      p.ntaps = 0; p.kblocks = pl.Cp / kKBlock; p.P = pl.P;
      for (int dh = 0; dh < g.kh; ++dh) {
        for (int dw = 0; dw < g.kw; ++dw) {
          Tap t;
          if (transposed) {
            // ho = hi*sh - ph + dh with ho = sh*q + ch  =>  hi = q + (ch + ph - dh)/sh when divisible
            int nh = ch + g.ph - dh, nw = cw + g.pw - dw;
            if (pymod(nh, g.sh) != 0 || pymod(nw, g.sw) != 0) continue;
            t.dh = (int16_t)(nh / g.sh); t.dw = (int16_t)(nw / g.sw); t.nsel = 0;
          } else {
            // hi = ho*sh - ph + dh = sh*(ho + a) + r
            int rh = pymod(dh - g.ph, g.sh), rw = pymod(dw - g.pw, g.sw);
            t.dh = (int16_t)((dh - g.ph - rh) / g.sh); t.dw = (int16_t)((dw - g.pw - rw) / g.sw);
            t.nsel = (int16_t)(rh * g.sw + rw);
          }
          t.wslab = (int16_t)(dh * g.kw + dw);
          p.taps[p.ntaps++] = t;
        }
      }
      if (transposed) {
        p.Hq_out = (g.Ho - ch + g.sh - 1) / g.sh; p.Wq_out = (g.Wo - cw + g.sw - 1) / g.sw;
        p.oh_mul = g.sh; p.oh_add = ch; p.ow_mul = g.sw; p.ow_add = cw;
      } else {
        p.Hq_out = g.Ho; p.Wq_out = g.Wo; p.oh_mul = 1; p.oh_add = 0; p.ow_mul = 1; p.ow_add = 0;
      }
      if (p.Hq_out <= 0 || p.Wq_out <= 0) continue;
      pick_tile(p.Hq_out, p.Wq_out, p.TH, p.TW);
      p.tiles_h = (p.Hq_out + p.TH - 1) / p.TH; p.tiles_w = (p.Wq_out + p.TW - 1) / p.TW;
      p.Ho = g.Ho; p.Wo = g.Wo; p.Cout2 = g.Cout2; p.npad = pl.npad;
      p.tmem_cols = 32; while (p.tmem_cols < pl.npad) p.tmem_cols <<= 1;
      p.bias = bias; p.y = y;
      const int stage_bytes = 2 * kABytes + 2 * pl.npad * 128;
      const int nit = p.ntaps * p.kblocks;
      // 2 CTAs/SM when the stages fit in ~110 KB each (hides the epilogue), else up to 4 stages in 220 KB
      int ns = (110 * 1024) / stage_bytes;
      if (ns < 2) ns = (220 * 1024) / stage_bytes;
      if (ns > 4) ns = 4;
      if (ns > nit && nit > 0) ns = nit;
      if (ns < 1) ns = 1;
      p.nstages = ns;
      const int smem = ns * stage_bytes + 1024;

      CUtensorMap tmX;
      uint64_t dims[4] = {(uint64_t)pl.Cp, (uint64_t)pl.Wq, (uint64_t)pl.Hq, (uint64_t)(B * 2 * pl.P)};
      uint32_t box[4] = {(uint32_t)kKBlock, (uint32_t)p.TW, (uint32_t)p.TH, 1};
      int rc = make_tmap_bf16(&tmX, xs, 4, dims, box);
      if (rc != SE_OK) return rc;

      static int smem_attr = 0;
      if (smem > smem_attr) {
        SE_CUDA(cudaFuncSetAttribute(cconv_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        smem_attr = smem;
      }
      dim3 grid((unsigned)(p.tiles_h * p.tiles_w), (unsigned)B);
      cconv_umma_kernel<<<grid, 128, smem, s>>>(tmX, tmW, p);
      rc = check_launch("cconv_umma_kernel");
      if (rc != SE_OK) return rc;
    }
  }
  return SE_OK;
}

}  // namespace se

// tcgen05 / TMEM / TMA engine for ComplexConv2d and ComplexConvTranspose2d
// (reference semantics: /root/reference/models/modules/complex_nn.py:52-91).
//
// One real implicit GEMM per layer against the packed operand [[Wr,-Wi],[Wi,Wr]]:
//     D[pixel, cout2] = sum_{tap} sum_{cin2} X[pixel (+) tap, cin2] * Wp[tap][cout2, cin2]
// fp32 accuracy on bf16 tensor cores through a 3-term split: x = x_hi + x_lo (both bf16),
//     x*w ~= x_lo*w_hi + x_hi*w_lo + x_hi*w_hi          (fp32 accumulate in TMEM)
//
// Activation format ("CL2"): channels-last bf16 hi|lo planes  [B][2][H][W][Cp]  (Cp = C padded to
// 64).  Strided convolutions read it through TMA *element strides* (the box walks every sh-th row /
// sw-th pixel), zero padding is TMA out-of-bounds fill, a transposed convolution is run as sh*sw
// stride-1 sub-convolutions (output parity classes) so no multiply by an inserted zero is issued.
// The epilogue can emit NCHW fp32 (the nn.Module contract) and/or CL2 for the next layer, with the
// folded bias and a leaky activation applied, so a chain of layers never leaves this format.
//
// What bounds the kernel is L2->SMEM fill (measured ~38 B/cycle/SM), not the tensor pipe, so the
// CTA tile is shaped to reuse every byte it stages:
//   * A (activations): taps that differ only by a row offset share ONE halo box
//     [nacc*TH + span rows][TW][kb]; each tap's MMA reads it through a row-shifted descriptor.
//   * B (weights): one [npad x kb] slab per tap feeds `nacc` accumulators (vertically stacked
//     128-pixel sub-tiles, nacc*npad <= 512 TMEM columns).
//   * A and B travel in separate mbarrier rings (2 A slots, up to 6 B slots).
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + tcgen05.mma issuer,
// warps 2-5 = epilogue (tcgen05.ld -> bias/activation -> NCHW fp32 and/or CL2 bf16 stores).
#include "engines.cuh"
#include "umma.cuh"
#include <cuda_bf16.h>
#include <cmath>
#include <cstdlib>

namespace se {

constexpr int kMaxTaps = 49;
constexpr int kMaxGroups = 48;
constexpr int kMaxClasses = 4;
constexpr int kSmemBudget = 226 * 1024;      // dynamic smem we allow ourselves incl. alignment slack (227 KB limit)

struct GTap { uint16_t arow, wslab, acol, pad_; };   // acol: column shift inside a shared halo box (wshare)
struct GGroup { int16_t row_add, col_add; uint8_t src, ntaps, tap0, pad_; };
struct GClass {
  int Hq_out, Wq_out, oh_mul, oh_add, ow_mul, ow_add;
  int g0, ng, tiles_h, tiles_w, tile_begin;
};

struct GemmParams {
  GClass cls[kMaxClasses];
  GGroup groups[kMaxGroups];
  GTap taps[kMaxTaps];
  int ncls, row_mul, col_mul;
  int kblocks[2], kcol0[2];
  int klast[2];               // descriptor units (16 B = 8 bf16) of real channels in the LAST k-block of a source: 2, 4, 6 or 8
  int TH, TW, nacc;
  int TWb;                    // pixels per row of the staged A box: TW, or TW + column span rounded up to 8 (wshare)
  int concat, acc_cols;       // concat: a_hi x [b_hi|b_lo] in one MMA (npad <= 128), 2*npad TMEM columns per accumulator
  int nsplit, unit_rows;      // split mode: every hi / lo slab travels as nsplit column slices of unit_rows rows
  int a_half, a_slot, b_unit, SA, SB;
  int ntiles, nbuf, buf_cols;  // persistent tile loop; nbuf TMEM accumulator buffers of buf_cols columns
  int Ho, Wo, Cout2, npad, tmem_cols, y_Cp;
  // phase folding (transposed convolution as ONE stride-1 convolution whose accumulator columns are the sh*sw output
  // phases side by side): column group ph = r * fold_w + s of quotient pixel (qh, qw) is output pixel
  // (qh * fold_h + r, qw * fold_w + s).  Cg = accumulator columns per phase, Cout2g = real channels per phase.
  // Not folded: fold_h = fold_w = 1, Cg = npad, Cout2g = Cout2.
  int fold_h, fold_w, Cg, Cout2g;
  int ph_h_mul, ph_w_mul;     // output distance between neighbouring folded phases (1 unless the launch covers a sub-grid)
  const float* bias;
  float slope;
  float* y;
  __nv_bfloat16* y_cl2;
  int64_t y_bs, y_cs, y_hs, y_ws, y_off, y_lim;   // fp32 output addressing (defaults = NCHW)
  const float* y_scale;                           // optional per-position scale (iSTFT normaliser)
  int y_vec4;                                     // fp32 output rows are channel-contiguous and 16-byte (1) / 32-byte (2) aligned
  unsigned long long* prof;                       // SE_GEMM_PROF=1: per-role wait/work cycle counters (tuning aid)
};

// ------------------------------------------------------------------------------ prep kernels
__device__ __forceinline__ void split_bf16(float v, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(v);
  lo = __float2bfloat16_rn(v - __bfloat162float(hi));
}

// NCHW fp32 -> CL2.  grid: (ceil(W/32), H, B*(Cp/64)); block 256
__global__ void __launch_bounds__(256)
cl2_from_nchw_kernel(const float* __restrict__ x, int64_t xs_b, int64_t xs_c, int64_t xs_h,
                     __nv_bfloat16* __restrict__ xs, int C, int H, int W, int Cp) {
  __shared__ __align__(16) __nv_bfloat16 s_hi[32][64 + 8];
  __shared__ __align__(16) __nv_bfloat16 s_lo[32][64 + 8];
  const int cblocks = Cp / 64;
  const int cb = blockIdx.z % cblocks, b = blockIdx.z / cblocks;
  const int h = blockIdx.y, w0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const bool pix_ok = (w0 + tx) < W;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int cl = ty + 8 * i, c = cb * 64 + cl;
    float v = (pix_ok && c < C) ? __ldg(x + b * xs_b + c * xs_c + h * xs_h + w0 + tx) : 0.f;
    split_bf16(v, s_hi[tx][cl], s_lo[tx][cl]);
  }
  __syncthreads();
  const int pix = threadIdx.x >> 3, chunk = threadIdx.x & 7;
  if (w0 + pix < W) {
    const int64_t plane = (int64_t)H * W * Cp;
    int64_t off = ((int64_t)h * W + w0 + pix) * Cp + cb * 64 + chunk * 8;
    *reinterpret_cast<uint4*>(xs + ((int64_t)b * 2 + 0) * plane + off) = *reinterpret_cast<const uint4*>(&s_hi[pix][chunk * 8]);
    *reinterpret_cast<uint4*>(xs + ((int64_t)b * 2 + 1) * plane + off) = *reinterpret_cast<const uint4*>(&s_lo[pix][chunk * 8]);
  }
}

// CL2 -> NCHW fp32 (hi + lo).  grid: (ceil(W/32), H, B*(Cp/64)); block 256
__global__ void __launch_bounds__(256)
cl2_to_nchw_kernel(const __nv_bfloat16* __restrict__ xs, float* __restrict__ y, int C, int H, int W, int Cp) {
  __shared__ float s[32][64 + 1];
  const int cblocks = Cp / 64;
  const int cb = blockIdx.z % cblocks, b = blockIdx.z / cblocks;
  const int h = blockIdx.y, w0 = blockIdx.x * 32;
  const int pix = threadIdx.x >> 3, chunk = threadIdx.x & 7;
  if (w0 + pix < W) {
    const int64_t plane = (int64_t)H * W * Cp;
    int64_t off = ((int64_t)h * W + w0 + pix) * Cp + cb * 64 + chunk * 8;
    uint4 hi = *reinterpret_cast<const uint4*>(xs + ((int64_t)b * 2 + 0) * plane + off);
    uint4 lo = *reinterpret_cast<const uint4*>(xs + ((int64_t)b * 2 + 1) * plane + off);
    const __nv_bfloat16* ph = reinterpret_cast<const __nv_bfloat16*>(&hi);
    const __nv_bfloat16* pl = reinterpret_cast<const __nv_bfloat16*>(&lo);
#pragma unroll
    for (int j = 0; j < 8; ++j) s[pix][chunk * 8 + j] = __bfloat162float(ph[j]) + __bfloat162float(pl[j]);
  }
  __syncthreads();
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  if (w0 + tx < W) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      int cl = ty + 8 * i, c = cb * 64 + cl;
      if (c < C) y[(((int64_t)b * C + c) * H + h) * W + w0 + tx] = s[tx][cl];
    }
  }
}

// packed [Cout2][Cin2][taps] fp32 -> wq[(tap*2 + s)][npad][Cpt] bf16; chan_map[col] = input channel
// of GEMM column `col` (or -1 for padding); NULL = identity.
__global__ void cconv_prep_w_kernel(const float* __restrict__ packed, const int* __restrict__ chan_map,
                                    __nv_bfloat16* __restrict__ wq, int Cout2, int Cin2, int taps, int npad,
                                    int Cpt) {
  int64_t total = (int64_t)taps * npad * Cpt;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  int col = (int)(i % Cpt); int64_t r = i / Cpt;
  int co = (int)(r % npad); int tap = (int)(r / npad);
  int c = chan_map ? chan_map[col] : col;
  float v = (co < Cout2 && c >= 0 && c < Cin2) ? __ldg(packed + ((int64_t)co * Cin2 + c) * taps + tap) : 0.f;
  __nv_bfloat16 hi, lo;
  split_bf16(v, hi, lo);
  int64_t slab = (int64_t)npad * Cpt;
  wq[((int64_t)tap * 2 + 0) * slab + (int64_t)co * Cpt + col] = hi;
  wq[((int64_t)tap * 2 + 1) * slab + (int64_t)co * Cpt + col] = lo;
}

// ------------------------------------------------------------------------------ GEMM kernel
// K-major operand tile, 128-byte swizzle: rows of 128 B (64 bf16), 8-row atoms of 1024 B back to
// back (SBO).  (Measured on B200, scripts/mma_rate.cu: with the 64-byte swizzle every MMA costs
// >= 140 cycles, with the 128-byte swizzle max(N/2, ~84) cycles -- so kb is always 64.)
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

constexpr int kGemmThreads = 320;          // warp 0 TMA producer, warp 1 MMA issuer, warps 2..9 epilogue

struct TileCoord { int ci, b, h0, w0, nacc; };

// mbarrier wait that charges its duration to `acc` when profiling is on
__device__ __forceinline__ void timed_wait(uint64_t* bar, uint32_t parity, bool prof, unsigned long long& acc) {
  if (!prof) { ptx::mbar_wait(bar, parity); return; }
  const long long t0 = clock64();
  ptx::mbar_wait(bar, parity);
  acc += (unsigned long long)(clock64() - t0);
}

__device__ __forceinline__ TileCoord decode_tile(const GemmParams& p, int tile) {
  // class-major (heaviest class first), then batch, then row tile, then column tile
  int ci = 0;
#pragma unroll
  for (int c = 1; c < kMaxClasses; ++c)
    if (c < p.ncls && tile >= p.cls[c].tile_begin) ci = c;
  const GClass& C = p.cls[ci];
  int t = tile - C.tile_begin;
  const int tw = t % C.tiles_w; t /= C.tiles_w;
  const int th = t % C.tiles_h;
  TileCoord tc;
  tc.ci = ci; tc.b = t / C.tiles_h;
  tc.h0 = th * p.TH * p.nacc; tc.w0 = tw * p.TW;
  int nacc = (C.Hq_out - tc.h0 + p.TH - 1) / p.TH;
  tc.nacc = nacc > p.nacc ? p.nacc : nacc;
  return tc;
}

// Persistent: grid = #SMs, every CTA walks tiles blockIdx.x, += gridDim.x.  The TMA producer runs
// ahead across tile boundaries, and with two TMEM accumulator buffers (when nacc * acc_cols <= 256)
// the epilogue of tile i overlaps the MMAs of tile i+1.
template <bool kConcat>
__global__ void __launch_bounds__(kGemmThreads, 1)
cconv_gemm_kernel(const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmX1,
                  const __grid_constant__ CUtensorMap tmW, const __grid_constant__ GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t a_full[2], a_empty[2], b_full[8], b_empty[8], t_full[2], t_empty[2];
  __shared__ uint32_t tmem_base_slot;
  __shared__ __align__(16) float s_bias[320];          // folded bias per accumulator column, 0 beyond Cout2

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_ring = smem;
  uint8_t* b_ring = smem + (size_t)p.SA * p.a_slot;

  if (threadIdx.x == 0) {
    ptx::prefetch_tmap(&tmX0);
    ptx::prefetch_tmap(&tmX1);
    ptx::prefetch_tmap(&tmW);
    for (int s = 0; s < p.SA; ++s) { ptx::mbar_init(&a_full[s], 1); ptx::mbar_init(&a_empty[s], 1); }
    for (int s = 0; s < p.SB; ++s) { ptx::mbar_init(&b_full[s], 1); ptx::mbar_init(&b_empty[s], 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(&t_full[s], 1); ptx::mbar_init(&t_empty[s], 8); }
    ptx::fence_mbar_init();
  }
  for (int c = threadIdx.x; c < 320; c += blockDim.x) s_bias[c] = (c < p.Cout2 && p.bias != nullptr) ? __ldg(p.bias + c) : 0.f;
  if (warp == 1) {
    ptx::tmem_alloc(&tmem_base_slot, (uint32_t)p.tmem_cols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_acc = tmem_base_slot;
  const int units = kConcat ? 1 : 2 * p.nsplit;      // B ring units per tap

  if (warp == 0) {
    if (ptx::elect_one()) {
      // ===== TMA producer =====
      uint32_t sa = 0, pa = 1, sb = 0, pb = 1;            // ring slot + the parity its "empty" barrier must show
      const uint32_t SB = (uint32_t)p.SB;
      const bool prof = p.prof != nullptr;
      unsigned long long w_a = 0, w_b = 0;
      const long long t_start = prof ? clock64() : 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
        const TileCoord tc = decode_tile(p, tile);
        const GClass& C = p.cls[tc.ci];
        for (int g = C.g0; g < C.g0 + C.ng; ++g) {
          const GGroup G = p.groups[g];
          const CUtensorMap* tm = G.src ? &tmX1 : &tmX0;
          const int wc = p.col_mul * tc.w0 + G.col_add, hc = p.row_mul * tc.h0 + G.row_add;
          const int nkb = p.kblocks[G.src], kcol0 = p.kcol0[G.src];
          for (int kbi = 0; kbi < nkb; ++kbi) {
            timed_wait(&a_empty[sa], pa, prof, w_a);
            uint8_t* as = a_ring + (size_t)sa * p.a_slot;
            ptx::mbar_arrive_expect_tx(&a_full[sa], (uint32_t)p.a_slot);
            ptx::tma_load_4d(as, tm, &a_full[sa], kbi * 64, wc, hc, tc.b * 2 + 0);
            ptx::tma_load_4d(as + p.a_half, tm, &a_full[sa], kbi * 64, wc, hc, tc.b * 2 + 1);
            if (++sa == 2) { sa = 0; pa ^= 1u; }
            const int kcol = kcol0 + kbi * 64;
            for (int ti = 0; ti < G.ntaps; ++ti) {
              const int slab0 = p.taps[G.tap0 + ti].wslab * units;
              for (int u = 0; u < units; ++u) {
                timed_wait(&b_empty[sb], pb, prof, w_b);
                uint8_t* bs = b_ring + (size_t)sb * p.b_unit;
                ptx::mbar_arrive_expect_tx(&b_full[sb], (uint32_t)p.b_unit);
                // tmW views wq as [taps * units][unit_rows][Cpt]; unit u = (hi|lo) * nsplit + column slice
                ptx::tma_load_3d(bs, &tmW, &b_full[sb], kcol, 0, slab0 + u);
                if (++sb == SB) { sb = 0; pb ^= 1u; }
              }
            }
          }
        }
      }
      if (prof) {
        atomicAdd(p.prof + 0, (unsigned long long)(clock64() - t_start));
        atomicAdd(p.prof + 1, w_a);
        atomicAdd(p.prof + 2, w_b);
      }
    }
  } else if (warp == 1) {
    if (ptx::elect_one()) {
      // ===== MMA issuer =====
      // One thread issues every MMA, so its own instruction stream is on the critical path: a 128x256x16
      // MMA retires in 128 cycles, and the first version of this loop (64-bit descriptor arithmetic,
      // div/mod ring indices, mode branches inside the k loop) spent ~230 cycles of scalar code per
      // MMA (SE_GEMM_PROF).  Everything below is 32-bit descriptor-low-word arithmetic on running
      // counters; the descriptor high word is the constant kDescHi.
      const uint32_t idesc1 = ptx::make_idesc_bf16_f32(128, kConcat ? p.npad : p.unit_rows);
      const uint32_t idesc2 = ptx::make_idesc_bf16_f32(128, 2 * p.npad);     // concat mode only
      const uint32_t sub16 = (uint32_t)(p.TH * p.TWb * 128) >> 4;  // one 128-pixel sub-tile of the halo box
      const uint32_t tw16 = (uint32_t)(p.TWb * 128) >> 4;
      const uint32_t half16 = (uint32_t)p.a_half >> 4;
      const uint32_t a_ring32 = (ptx::smem_u32(a_ring) >> 4) | 0x10000u, a_slot16 = (uint32_t)p.a_slot >> 4;
      const uint32_t b_ring32 = (ptx::smem_u32(b_ring) >> 4) | 0x10000u, b_unit16 = (uint32_t)p.b_unit >> 4;
      const uint32_t SB = (uint32_t)p.SB, acc_cols = (uint32_t)p.acc_cols, unit_rows = (uint32_t)p.unit_rows;
      const int nsplit = p.nsplit;
      uint32_t sa = 0, pa = 0, sb = 0, pb = 0, lt = 0;
      const bool prof = p.prof != nullptr;
      unsigned long long w_t = 0, w_a = 0, w_b = 0;
      const long long t_start = prof ? clock64() : 0;
      for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++lt) {
        const TileCoord tc = decode_tile(p, tile);
        const GClass& C = p.cls[tc.ci];
        const uint32_t buf = lt % (uint32_t)p.nbuf, tph = (lt / (uint32_t)p.nbuf) & 1u;
        timed_wait(&t_empty[buf], tph ^ 1u, prof, w_t);    // epilogue has drained this accumulator buffer
        ptx::tc_fence_after();
        const uint32_t tmem_buf = tmem_acc + buf * (uint32_t)p.buf_cols;
        const uint32_t nacc = (uint32_t)tc.nacc;
        uint32_t keep = 0;                                  // 0 until the first tap of the tile has been issued
        for (int g = C.g0; g < C.g0 + C.ng; ++g) {
          const GGroup G = p.groups[g];
          const int nkb = p.kblocks[G.src];
          for (int kbi = 0; kbi < nkb; ++kbi) {
            // k-steps of 16 channels; the pad channels at the end of a source are zeros on both operands: skipped
            const uint32_t kend = kbi == nkb - 1 ? (uint32_t)p.klast[G.src] : 8u;
            timed_wait(&a_full[sa], pa, prof, w_a);
            ptx::tc_fence_after();
            const uint32_t a_base32 = a_ring32 + sa * a_slot16;
            for (int ti = 0; ti < G.ntaps; ++ti) {
              const GTap tp = p.taps[G.tap0 + ti];
              // row-shifted (multiples of 1024 B) and, for single-row kernels, column-shifted (128 B per pixel) view
              // of the group's halo box.  A column shift moves the start address INSIDE a 1024-byte swizzle atom; the
              // tensor core applies the 128-byte swizzle to absolute shared-memory address bits (as the TMA did when it
              // wrote the box), so the shifted start needs nothing else -- measured: with the descriptor's matrix
              // base-offset field set to (shift & 7) every result is wrong, with 0 every result is right
              // (tests/test_gpu_fold.py::test_single_row_kernels_share_one_halo_box).
              const uint32_t a_tap32 = a_base32 + (uint32_t)tp.arow * tw16 + (uint32_t)tp.acol * 8u;
              if (kConcat) {
                timed_wait(&b_full[sb], pb, prof, w_b);
                ptx::tc_fence_after();
                const uint32_t b32 = b_ring32 + sb * b_unit16;
                uint32_t ah = a_tap32, d = tmem_buf;
                for (uint32_t a = 0; a < nacc; ++a, ah += sub16, d += acc_cols) {
                  const uint32_t al = ah + half16;
                  // D[:, 0:2N) += a_hi * [b_hi | b_lo]   ;   D[:, 0:N) += a_lo * b_hi
                  ptx::umma_bf16_lo32(d, ah, b32, idesc2, keep);
                  ptx::umma_bf16_lo32(d, al, b32, idesc1, 1);
#pragma unroll
                  for (uint32_t k = 2; k < 8; k += 2) {     // 16 bf16 = 32 B = 2 descriptor units
                    if (k < kend) {
                      ptx::umma_bf16_lo32(d, ah + k, b32 + k, idesc2, 1);
                      ptx::umma_bf16_lo32(d, al + k, b32 + k, idesc1, 1);
                    }
                  }
                }
                ptx::umma_commit(&b_empty[sb]);
                if (++sb == SB) { sb = 0; pb ^= 1u; }
              } else {
                // b_hi units (one per column slice): D += a_lo * b_hi + a_hi * b_hi
                for (int nh = 0; nh < nsplit; ++nh) {
                  timed_wait(&b_full[sb], pb, prof, w_b);
                  ptx::tc_fence_after();
                  const uint32_t b32 = b_ring32 + sb * b_unit16;
                  uint32_t ah = a_tap32, d = tmem_buf + (uint32_t)nh * unit_rows;
                  for (uint32_t a = 0; a < nacc; ++a, ah += sub16, d += acc_cols) {
                    const uint32_t al = ah + half16;
                    ptx::umma_bf16_lo32(d, al, b32, idesc1, keep);
                    ptx::umma_bf16_lo32(d, ah, b32, idesc1, 1);
#pragma unroll
                    for (uint32_t k = 2; k < 8; k += 2) {
                      if (k < kend) {
                        ptx::umma_bf16_lo32(d, al + k, b32 + k, idesc1, 1);
                        ptx::umma_bf16_lo32(d, ah + k, b32 + k, idesc1, 1);
                      }
                    }
                  }
                  ptx::umma_commit(&b_empty[sb]);
                  if (++sb == SB) { sb = 0; pb ^= 1u; }
                }
                // b_lo units: D += a_hi * b_lo
                for (int nh = 0; nh < nsplit; ++nh) {
                  timed_wait(&b_full[sb], pb, prof, w_b);
                  ptx::tc_fence_after();
                  const uint32_t b32 = b_ring32 + sb * b_unit16;
                  uint32_t ah = a_tap32, d = tmem_buf + (uint32_t)nh * unit_rows;
                  for (uint32_t a = 0; a < nacc; ++a, ah += sub16, d += acc_cols) {
#pragma unroll
                    for (uint32_t k = 0; k < 8; k += 2)
                      if (k < kend) ptx::umma_bf16_lo32(d, ah + k, b32 + k, idesc1, 1);
                  }
                  ptx::umma_commit(&b_empty[sb]);
                  if (++sb == SB) { sb = 0; pb ^= 1u; }
                }
              }
              keep = 1;
            }
            ptx::umma_commit(&a_empty[sa]);
            if (++sa == 2) { sa = 0; pa ^= 1u; }
          }
        }
        ptx::umma_commit(&t_full[buf]);
      }
      if (prof) {
        atomicAdd(p.prof + 3, (unsigned long long)(clock64() - t_start));
        atomicAdd(p.prof + 4, w_t);
        atomicAdd(p.prof + 5, w_a);
        atomicAdd(p.prof + 6, w_b);
      }
    }
  } else {
    // ===== epilogue: TMEM -> registers -> (+bias, activation) -> fp32 and/or CL2 bf16 =====
    // Eight warps: warp w may only touch TMEM lanes 32*(w%4)..+31, so two warps share each lane
    // quarter and take alternate 32-column chunks.  All per-element decisions (which outputs, limits,
    // scale, vector stores) are warp-uniform and hoisted out of the column loops.
    const int q = warp & 3;                       // TMEM lane quarter this warp may access
    const int chalf = (warp - 2) >> 2;            // 0: even chunks, 1: odd chunks
    const int m = q * 32 + lane;
    const int hl = m / p.TW, wl = m - hl * p.TW;
    const int64_t plane = (int64_t)p.Ho * p.Wo;
    const bool do_f32 = p.y != nullptr, do_cl2 = p.y_cl2 != nullptr;
    const bool checked = p.y_lim != 0 || p.y_scale != nullptr;
    const bool act = p.slope != 1.f;
    const float slope = p.slope;
    const int64_t lo_off = plane * p.y_Cp;
    const int nph = p.fold_h * p.fold_w, Cg = p.Cg;
    // 32-column chunks per phase: the accumulator columns, then (CL2 output only) the zero pad channels up to y_Cp
    const int cpp = ((do_cl2 && p.y_Cp > Cg ? p.y_Cp : Cg) + 31) >> 5;
    uint32_t lt = 0;
    const bool prof = p.prof != nullptr;
    unsigned long long w_f = 0;
    const long long t_start = prof ? clock64() : 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++lt) {
      const TileCoord tc = decode_tile(p, tile);
      const GClass& C = p.cls[tc.ci];
      const uint32_t buf = lt % (uint32_t)p.nbuf, tph = (lt / (uint32_t)p.nbuf) & 1u;
      timed_wait(&t_full[buf], tph, prof, w_f);
      ptx::tc_fence_after();
      const bool have = C.ng > 0;                 // a class no tap reaches is bias-only
      const int64_t b = tc.b;
      for (int a = 0; a < tc.nacc; ++a) {
        const int qh = tc.h0 + a * p.TH + hl, qw = tc.w0 + wl;
        const bool q_ok = qh < C.Hq_out && qw < C.Wq_out;
        const int ho0 = qh * C.oh_mul + C.oh_add, wo0 = qw * C.ow_mul + C.ow_add;
        const uint32_t trow = tmem_acc + ((uint32_t)(q * 32) << 16) + buf * (uint32_t)p.buf_cols + (uint32_t)(a * p.acc_cols);
        // output coordinates of phase 0 (the only phase unless the launch is folded)
        int ho = ho0, wo = wo0;
        bool ok = q_ok && ho < p.Ho && wo < p.Wo;
        int64_t ypos = (int64_t)wo * p.y_ws;                 // + co * y_cs: position inside the (b, ho) row block
        float* yb = do_f32 ? p.y + b * p.y_bs + (int64_t)ho * p.y_hs + p.y_off + ypos : nullptr;
        __nv_bfloat16* cb = do_cl2 ? p.y_cl2 + (b * 2 * plane + (int64_t)ho * p.Wo + wo) * p.y_Cp : nullptr;
        for (int g = chalf; g < nph * cpp; g += 2) {
          int cc = g << 5, tcol = cc;                        // channel offset inside the phase, accumulator column
          if (nph > 1) {
            const int ph = g / cpp;
            cc = (g - ph * cpp) << 5;
            tcol = ph * Cg + cc;
            const int r = ph / p.fold_w;
            ho = ho0 + r * p.ph_h_mul; wo = wo0 + (ph - r * p.fold_w) * p.ph_w_mul;
            ok = q_ok && ho < p.Ho && wo < p.Wo;
            ypos = (int64_t)wo * p.y_ws;
            yb = do_f32 ? p.y + b * p.y_bs + (int64_t)ho * p.y_hs + p.y_off + ypos : nullptr;
            cb = do_cl2 ? p.y_cl2 + (b * 2 * plane + (int64_t)ho * p.Wo + wo) * p.y_Cp : nullptr;
          }
          const int c0 = cc;
          float v[32];
          if (have && cc < Cg) {
            uint32_t r[32];
            ptx::tmem_ld_32x32(trow + (uint32_t)tcol, r);
            if (kConcat) {
              uint32_t r2[32];
              ptx::tmem_ld_32x32(trow + (uint32_t)(p.npad + tcol), r2);
              ptx::tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) + __uint_as_float(r2[j]);
            } else {
              ptx::tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
            }
            if (cc + 32 > Cg) {
              // accumulator width is a multiple of 16, chunks are 32 wide: the upper half of the last chunk is not
              // part of this accumulator -- those columns are pad channels of the output and must read as zero
#pragma unroll
              for (int j = 16; j < 32; ++j) v[j] = 0.f;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = 0.f;
          }
          if (!ok) continue;
          if (cc < Cg) {
            const float4* b4 = reinterpret_cast<const float4*>(s_bias + tcol);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 bb = b4[j];
              v[4 * j] += bb.x; v[4 * j + 1] += bb.y; v[4 * j + 2] += bb.z; v[4 * j + 3] += bb.w;
            }
          }
          if (act) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f) + slope * fminf(v[j], 0.f);
          }
          if (do_f32 && c0 < p.Cout2g) {
            const int nv = p.Cout2g - c0;                     // valid columns in this chunk (may exceed 32)
            if (p.y_vec4 == 2) {
              // 256-bit stores: a whole 32-byte sector per lane and instruction (lanes are a row apart)
#pragma unroll
              for (int j = 0; j < 4; ++j)
                if (8 * j + 7 < nv) {
                  uint32_t w[8];
#pragma unroll
                  for (int i = 0; i < 8; ++i) w[i] = __float_as_uint(v[8 * j + i]);
                  ptx::st_global_v8(yb + c0 + 8 * j, w);
                }
            } else if (p.y_vec4) {
              float4* y4 = reinterpret_cast<float4*>(yb + c0);
#pragma unroll
              for (int j = 0; j < 8; ++j)
                if (4 * j + 3 < nv) y4[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            } else if (!checked) {
              float* yp = yb + (int64_t)c0 * p.y_cs;
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < nv) yp[(int64_t)j * p.y_cs] = v[j];
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                if (j < nv) {
                  const int64_t pos = ypos + (int64_t)(c0 + j) * p.y_cs;
                  const int64_t dst = pos + p.y_off;
                  if (p.y_lim == 0 || (dst >= 0 && dst < p.y_lim))
                    yb[(int64_t)(c0 + j) * p.y_cs] = p.y_scale ? v[j] * __ldg(p.y_scale + pos) : v[j];
                }
              }
            }
          }
          if (do_cl2 && c0 < p.y_Cp) {
            uint32_t hi[16], lo[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float f0 = v[2 * j], f1 = v[2 * j + 1];
              const __nv_bfloat162 h2 = __floats2bfloat162_rn(f0, f1);
              const uint32_t hb = *reinterpret_cast<const uint32_t*>(&h2);
              const float r0 = f0 - __uint_as_float(hb << 16), r1 = f1 - __uint_as_float(hb & 0xffff0000u);
              const __nv_bfloat162 l2 = __floats2bfloat162_rn(r0, r1);
              hi[j] = hb;
              lo[j] = *reinterpret_cast<const uint32_t*>(&l2);
            }
            // 256-bit stores: a full 32-byte sector per lane and instruction (lanes are a pixel >= 128 B apart)
            __nv_bfloat16* dh = cb + c0;
            __nv_bfloat16* dl = cb + lo_off + c0;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              ptx::st_global_v8(dh + 16 * j, hi + 8 * j);
              ptx::st_global_v8(dl + 16 * j, lo + 8 * j);
            }
          }
        }
      }
      // this warp is done reading the buffer: hand it back to the MMA issuer
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&t_empty[buf]);
    }
    if (prof && warp == 2 && lane == 0) {
      atomicAdd(p.prof + 7, (unsigned long long)(clock64() - t_start));
      atomicAdd(p.prof + 8, w_f);
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

// bf16 tensor; dims / strides (bytes, for dims 1..rank-1) innermost-first; box = bounding box in
// elements; estr = element (traversal) strides; zero OOB fill
static int make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
                          const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* estr, bool sw128) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr) return fail(SE_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t gdim[5], gstride[4];
  cuuint32_t bdim[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i]; bdim[i] = box[i]; es[i] = estr[i];
    if (i > 0) gstride[i - 1] = strides_bytes[i - 1];
  }
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstride,
                  bdim, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  sw128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(SE_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return SE_OK;
}

static inline int round_up(int a, int b) { return (a + b - 1) / b * b; }
static inline int pymod(int a, int b) { int r = a % b; return r < 0 ? r + b : r; }
static inline int floordiv(int a, int b) { return (a - pymod(a, b)) / b; }

// classes / groups / taps of a layer (tile-shape independent part of GemmParams)
static int build_structure(const se_cconv_args* a, GemmParams& p, int& span_max, bool wshare, int& span_w) {
  span_w = 0;
  const bool tr = a->transposed != 0;
  const int ph = a->ph + (tr ? 0 : a->pad_top), pw = a->pw + (tr ? 0 : a->pad_left);
  p.ncls = 0;
  int ng = 0, nt = 0;
  span_max = 0;
  const int nch = tr ? a->sh : 1, ncw = tr ? a->sw : 1;
  p.row_mul = tr ? 1 : a->sh;
  p.col_mul = tr ? 1 : a->sw;
  for (int ch = 0; ch < nch; ++ch) {
    for (int cw = 0; cw < ncw; ++cw) {
      GClass c;
      if (tr) {
        c.Hq_out = (a->Ho - ch + a->sh - 1) / a->sh; c.Wq_out = (a->Wo - cw + a->sw - 1) / a->sw;
        c.oh_mul = a->sh; c.oh_add = ch; c.ow_mul = a->sw; c.ow_add = cw;
      } else {
        // folded launch: the M dimension walks the quotient grid, the column groups are the output phases
        const int fh = a->fold_h > 1 ? a->fold_h : 1, fw = a->fold_w > 1 ? a->fold_w : 1;
        const int hm = a->out_h_mul > 1 ? a->out_h_mul : 1, wm = a->out_w_mul > 1 ? a->out_w_mul : 1;
        c.oh_mul = fh * hm; c.oh_add = a->out_h_add; c.ow_mul = fw * wm; c.ow_add = a->out_w_add;
        c.Hq_out = a->Ho > c.oh_add ? (a->Ho - c.oh_add + c.oh_mul - 1) / c.oh_mul : 0;
        c.Wq_out = a->Wo > c.ow_add ? (a->Wo - c.ow_add + c.ow_mul - 1) / c.ow_mul : 0;
      }
      if (c.Hq_out <= 0 || c.Wq_out <= 0) continue;
      c.g0 = ng; c.ng = 0;
      const int first_group = ng;
      // groups of source 0: one per (row phase r, kernel column dw); taps ordered by row offset
      const int nphase = tr ? 1 : a->sh;
      for (int r = 0; r < nphase; ++r) {
        for (int dw = 0; dw < a->kw; ++dw) {
          int col_add;
          if (tr) {
            int nw = cw + pw - dw;
            if (pymod(nw, a->sw) != 0) continue;
            col_add = nw / a->sw;
          } else {
            col_add = dw - pw;
          }
          int offs[kMaxTaps], slabs[kMaxTaps], n = 0;
          for (int dh = 0; dh < a->kh; ++dh) {
            int off;
            if (tr) {
              int nh = ch + ph - dh;
              if (pymod(nh, a->sh) != 0) continue;
              off = nh / a->sh;
            } else {
              if (pymod(dh - ph, a->sh) != r) continue;
              off = floordiv(dh - ph, a->sh);
            }
            offs[n] = off; slabs[n] = dh * a->kw + dw; ++n;
          }
          if (n == 0) continue;
          int omin = offs[0], omax = offs[0];
          for (int i = 1; i < n; ++i) { if (offs[i] < omin) omin = offs[i]; if (offs[i] > omax) omax = offs[i]; }
          if (ng >= kMaxGroups || nt + n > kMaxTaps) return fail(SE_ERR_UNSUPPORTED, "se_cconv_gemm: too many taps");
          GGroup G;
          G.row_add = (int16_t)(tr ? omin : a->sh * omin + r);
          G.col_add = (int16_t)col_add;
          G.src = 0; G.ntaps = (uint8_t)n; G.tap0 = (uint8_t)nt; G.pad_ = 0;
          for (int i = 0; i < n; ++i) {
            p.taps[nt].arow = (uint16_t)(offs[i] - omin); p.taps[nt].wslab = (uint16_t)slabs[i];
            p.taps[nt].acol = 0; p.taps[nt].pad_ = 0; ++nt;
          }
          if (omax - omin > span_max) span_max = omax - omin;
          p.groups[ng++] = G;
        }
      }
      if (wshare && ng - first_group > 1) {
        // single-row kernel: the kernel columns differ only by a column shift of the SAME input rows, so they share
        // ONE staged halo box [rows][TW + span][64] and reach it through column-shifted descriptors (128 B per pixel,
        // matrix base offset = shift & 7) instead of one A load per column: kw x less L2 -> shared-memory fill
        int cmin = p.groups[first_group].col_add, cmax = cmin;
        bool ok = true;
        for (int g = first_group; g < ng; ++g) {
          ok = ok && p.groups[g].ntaps == 1 && p.groups[g].row_add == p.groups[first_group].row_add;
          if (p.groups[g].col_add < cmin) cmin = p.groups[g].col_add;
          if (p.groups[g].col_add > cmax) cmax = p.groups[g].col_add;
        }
        if (!ok) return fail(SE_ERR_UNSUPPORTED, "se_cconv_gemm: column sharing needs a single-row kernel");
        for (int g = first_group; g < ng; ++g)
          p.taps[p.groups[g].tap0].acol = (uint16_t)(p.groups[g].col_add - cmin);
        p.groups[first_group].col_add = (int16_t)cmin;
        p.groups[first_group].ntaps = (uint8_t)(ng - first_group);
        ng = first_group + 1;
        if (cmax - cmin > span_w) span_w = cmax - cmin;
      }
      // source 1 re-uses the same taps
      if (a->nsrc == 2) {
        const int n0 = ng - first_group;
        if (ng + n0 > kMaxGroups) return fail(SE_ERR_UNSUPPORTED, "se_cconv_gemm: too many tap groups");
        for (int i = 0; i < n0; ++i) { GGroup G = p.groups[first_group + i]; G.src = 1; p.groups[ng++] = G; }
      }
      c.ng = ng - c.g0;
      if (c.ng == 0) {
        // a class no kernel tap reaches (kernel smaller than stride): output is bias only; keep one empty class
      }
      c.tiles_h = c.tiles_w = c.tile_begin = 0;
      p.cls[p.ncls++] = c;
    }
  }
  return SE_OK;
}

struct TileChoice { int TH, TW, nacc, SA, SB, concat, nsplit; double cost; };

// MMA cost model measured with scripts/mma_rate.cu (cta_group::1, SS operands, 128B swizzle):
//   cycles(128 x N x 16) = max(N / 2, ~84)
// cycles of one 128 x n x 16 bf16 MMA.  model 0: max(n / 2, 84) (scripts/mma_rate.cu, MMAs alone on the SM);
// model 1: what the issuing thread sees inside the pipeline (SE_GEMM_PROF: 135 at n = 256, 90 at 128, ~75 at <= 64)
static inline double mma_cycles(int n) {
  static const char* env = getenv("SE_GEMM_MMA_MODEL");
  static const int model = env ? atoi(env) : 0;
  if (model == 0) return fmax(n / 2.0, 84.0);
  if (n <= 64) return 75.0;
  if (n <= 128) return 75.0 + (n - 64) * (15.0 / 64.0);
  return 90.0 + (n - 128) * (45.0 / 128.0);
}
constexpr double kFillBytesPerCycle = 36.0;      // sustained L2 -> SMEM fill per SM
constexpr double kFillLatency = 2500.0;          // request -> data landed (cycles), covers one B unit

static TileChoice choose_tile(const se_cconv_args* a, const GemmParams& p, int span_max, int npad, int span_w) {
  TileChoice best{0, 0, 0, 0, 0, 0, 1, 1e300};
  const int nsm = num_sms();
  static const char* ovh_env = getenv("SE_GEMM_TILE_OVH");   // tuning aid: fixed cycles per tile in the cost model
  // measured with SE_GEMM_PROF on the short-K layers: ~2-2.8 kcycles per tile that no MMA covers (DCUNet 7x1: 800 -> 2500
  // moves the plan from 4x32 / 1 accumulator to 8x16 / 2 accumulators, 1.98 -> 1.56 ms)
  static const char* epi_env = getenv("SE_GEMM_EPI_COL");    // tuning aid: epilogue cycles per accumulator column
  // measured (SE_GEMM_PROF, 8 epilogue warps, 256-bit stores): 41-52 cycles per accumulator column
  const double epi_col = epi_env ? atof(epi_env) : 50.0;
  // Split mode under-delivers against this model wherever the concat form exists (three 128-row operand reads per
  // k-step instead of two, two B units per tap in a 4-5 slot ring).  Forced-plan sweeps (SE_GEMM_FORCE) over DCCRN,
  // DCUNet and FRCRN put the concat form ahead on every N = 128 layer, by 6-20 %; penalties of 1.15 / 1.3 / 1.6 gave
  // DCUNet conv stacks of 30.3 / 29.9 / 29.6 ms (31.3 with none) and FRCRN 19.0 / 18.8 / 17.9 ms (18.8): at 1.6 the
  // planner takes the concat form whenever npad <= 128.
  static const char* pen_env = getenv("SE_GEMM_SPLIT_PENALTY");
  const double split_penalty = pen_env ? atof(pen_env) : 1.6;
  const double tile_ovh = ovh_env ? atof(ovh_env) : 2500.0;
  for (int mode = 0; mode < 3; ++mode) {
    // mode 0: concat (a_hi x [b_hi|b_lo], a_lo x b_hi); 1: split hi / lo; 2: split + two column slices
    const bool concat = mode == 0;
    const int nsplit = mode == 2 ? 2 : 1;
    if (concat && npad > 128) continue;
    if (nsplit == 2 && (npad % 32 != 0 || npad < 64)) continue;
    const int acc_cols = concat ? 2 * npad : npad;
    const int unit_rows = concat ? 2 * npad : npad / nsplit;
    const int64_t b_unit = (int64_t)unit_rows * 128;
    const int units = concat ? 1 : 2 * nsplit;
    for (int TW = 128; TW >= 8; TW >>= 1) {
      const int TH = 128 / TW;
      if (span_w > 0 && TW != 128) continue;               // shared column halo: the 128 GEMM rows are ONE image row
      const int TWb = span_w > 0 ? round_up(TW + span_w, 8) : TW;
      for (int nacc = 1; nacc <= 4; nacc <<= 1) {
        const int tm_need = (nacc - 1) * acc_cols + (concat ? npad : 0) + round_up(npad, 32);
        if (tm_need > 512) continue;
        const int rows = nacc * TH + span_max;
        if (rows * p.row_mul > 256 || TWb * p.col_mul > 256) continue;
        const int64_t a_slot = 2LL * rows * TWb * 128;
        const int SA = 2;
        int64_t left = kSmemBudget - 1024 - SA * a_slot;
        if (left < 2 * b_unit) continue;
        int SB = (int)(left / b_unit);
        if (SB > 8) SB = 8;
        double total = 0.0; int64_t tiles = 0;
        for (int ci = 0; ci < p.ncls; ++ci) {
          const GClass& c = p.cls[ci];
          const int64_t th = (c.Hq_out + nacc * TH - 1) / (nacc * TH), tw = (c.Wq_out + TW - 1) / TW;
          const int64_t n = th * tw * a->B;
          const double acc_avg = (double)((c.Hq_out + TH - 1) / TH) / (double)th;   // live accumulators
          // per-unit MMA time (4 k-steps); a unit shorter than the fill latency / lookahead stalls the pipe
          const double floor_u = kFillLatency / (double)(SB - 1);
          double t_tap;
          if (concat) t_tap = fmax(acc_avg * 4.0 * (mma_cycles(2 * npad) + mma_cycles(npad)), floor_u);
          else t_tap = nsplit * (fmax(acc_avg * 4.0 * 2.0 * mma_cycles(unit_rows), floor_u) +
                                 fmax(acc_avg * 4.0 * mma_cycles(unit_rows), floor_u));
          double per_tile = 0.0;
          for (int g = c.g0; g < c.g0 + c.ng; ++g) {
            const GGroup& G = p.groups[g];
            const int kblocks = (G.src ? a->src_Cp[1] : a->src_Cp[0]) / 64;
            const double bytes = (double)a_slot + (double)G.ntaps * units * b_unit;
            per_tile += kblocks * fmax(G.ntaps * t_tap, bytes / kFillBytesPerCycle);
          }
          // epilogue: measured ~180 cycles per accumulator column (4 warps, one per SM sub-partition);
          // it hides behind the next tile's MMAs only with two TMEM buffers
          const int buf_cols = round_up(nacc * acc_cols, 32);
          const bool two_bufs = buf_cols + tm_need <= 512;
          const double epi = acc_avg * (epi_col * npad + 1500.0);
          per_tile = two_bufs ? fmax(per_tile, epi) + tile_ovh : per_tile + epi + tile_ovh;
          total += per_tile * n;
          tiles += n;
        }
        const double waves = ceil((double)tiles / nsm);
        double cost = total / (double)tiles * waves;                 // ~cycles, 1 CTA per SM
        if (TW < 32) cost *= 1.0 + 0.03 * (32 / TW);                 // narrow rows: poorly coalesced epilogue
        if (!concat && npad <= 128) cost *= split_penalty;
        if (cost < best.cost) best = TileChoice{TH, TW, nacc, SA, SB, concat ? 1 : 0, nsplit, cost};
      }
    }
  }
  return best;
}

int cconv_gemm_launch(const se_cconv_args* a, cudaStream_t s) {
  const char* nm = "se_cconv_gemm";
  if (a->Cout2 > 256 || a->kh * a->kw > kMaxTaps)
    return fail(SE_ERR_UNSUPPORTED, "%s: needs Cout2 <= 256 and kh*kw <= %d", nm, kMaxTaps);
  if (a->B > 32767) return fail(SE_ERR_UNSUPPORTED, "%s: batch too large", nm);
  if (a->transposed && (a->pad_top || a->pad_left))
    return fail(SE_ERR_UNSUPPORTED, "%s: pad_top/pad_left only for the forward convolution", nm);
  if ((a->out_h_mul > 1 || a->out_w_mul > 1 || a->out_h_add || a->out_w_add) &&
      (a->transposed || a->out_h_add < 0 || a->out_w_add < 0 || a->out_h_add >= (a->out_h_mul > 1 ? a->out_h_mul : 1) ||
       a->out_w_add >= (a->out_w_mul > 1 ? a->out_w_mul : 1)))
    return fail(SE_ERR_BAD_ARG, "%s: out_*_mul / out_*_add describe a sub-grid of a forward launch, 0 <= add < mul", nm);
  if (a->fold_h > 1 || a->fold_w > 1) {
    const int nph = (a->fold_h > 1 ? a->fold_h : 1) * (a->fold_w > 1 ? a->fold_w : 1);
    if (a->transposed || a->fold_cg <= 0 || a->fold_cg % 32 != 0 ||
        a->Cout2 != nph * a->fold_cg || a->fold_cout2 <= 0 || a->fold_cout2 > a->fold_cg)
      return fail(SE_ERR_BAD_ARG, "%s: a folded launch is a forward convolution over the quotient grid with Cout2 = phases "
                                  "* fold_cg columns, fold_cg a multiple of 32 and fold_cout2 <= fold_cg", nm);
  }
  for (int i = 0; i < a->nsrc; ++i)
    if (a->src_Cp[i] <= 0 || a->src_Cp[i] % 64 != 0)
      return fail(SE_ERR_BAD_ARG, "%s: source channel pitch must be a positive multiple of 64", nm);

  GemmParams p;
  int span_max = 0, span_w = 0;
  // single-row kernels (1 x k, column stride 1): the k kernel columns share one staged halo box (SE_GEMM_WSHARE=0: off)
  static const char* ws_env = getenv("SE_GEMM_WSHARE");
  const bool wshare = !(ws_env && atoi(ws_env) == 0) && a->kh == 1 && a->kw > 1 && a->sw == 1 &&
                      (!a->transposed || a->sh == 1);
  int rc = build_structure(a, p, span_max, wshare, span_w);
  if (rc != SE_OK) return rc;
  const int npad = round_up(a->Cout2, 16);
  static const char* force = getenv("SE_GEMM_FORCE");       // "TW,nacc,mode" (tuning aid; mode 0 concat, 1 split, 2 split x2)
  static const char* dbg_env = getenv("SE_GEMM_DEBUG");     // bit2: print the plan of every launch
  const int dbg = dbg_env ? atoi(dbg_env) : 0;
  TileChoice tc = choose_tile(a, p, span_max, npad, span_w);
  if (force) {
    int fTW = 0, fn = 0, fm = 0;
    if (span_w == 0 && sscanf(force, "%d,%d,%d", &fTW, &fn, &fm) == 3 && fTW >= 8 && fTW <= 128 && fm >= 0 && fm <= 2 &&
        (fm != 0 || npad <= 128) && (fm != 2 || (npad % 32 == 0 && npad >= 64))) {
      const int ac = fm == 0 ? 2 * npad : npad, ur = fm == 0 ? 2 * npad : npad / (fm == 2 ? 2 : 1);
      const int64_t as = 2LL * (fn * (128 / fTW) + span_max) * fTW * 128, bu = (int64_t)ur * 128;
      int64_t left = kSmemBudget - 1024 - 2 * as;
      int SB = left > 0 ? (int)(left / bu) : 0;
      if (SB > 8) SB = 8;
      if (SB >= 2 && (fn - 1) * ac + (fm == 0 ? npad : 0) + round_up(npad, 32) <= 512)
        tc = TileChoice{128 / fTW, fTW, fn, 2, SB, fm == 0 ? 1 : 0, fm == 2 ? 2 : 1, 0.0};
    }
  }
  const bool concat = tc.concat != 0;
  if (tc.TW == 0) return fail(SE_ERR_UNSUPPORTED, "%s: no tile shape fits shared memory", nm);
  p.TH = tc.TH; p.TW = tc.TW; p.nacc = tc.nacc;
  p.concat = concat ? 1 : 0; p.acc_cols = concat ? 2 * npad : npad;
  p.nsplit = tc.nsplit; p.unit_rows = concat ? 2 * npad : npad / tc.nsplit;
  const int rows = tc.nacc * tc.TH + span_max;
  p.TWb = span_w > 0 ? round_up(tc.TW + span_w, 8) : tc.TW;
  p.a_half = rows * p.TWb * 128; p.a_slot = 2 * p.a_half;
  p.b_unit = p.unit_rows * 128;
  p.SA = tc.SA; p.SB = tc.SB;
  p.Ho = a->Ho; p.Wo = a->Wo; p.Cout2 = a->Cout2; p.npad = npad; p.y_Cp = a->y_Cp;
  const bool folded = a->fold_h > 1 || a->fold_w > 1;
  p.fold_h = a->fold_h > 1 ? a->fold_h : 1; p.fold_w = a->fold_w > 1 ? a->fold_w : 1;
  p.Cg = folded ? a->fold_cg : npad;
  p.Cout2g = folded ? a->fold_cout2 : a->Cout2;
  p.ph_h_mul = a->out_h_mul > 1 ? a->out_h_mul : 1; p.ph_w_mul = a->out_w_mul > 1 ? a->out_w_mul : 1;
  // TMEM: nacc accumulators per buffer; two buffers when they fit (epilogue overlaps the next tile's MMAs)
  const int tm_one = (tc.nacc - 1) * p.acc_cols + (concat ? npad : 0) + round_up(npad, 32);
  p.buf_cols = round_up(tc.nacc * p.acc_cols, 32);
  p.nbuf = (p.buf_cols + tm_one <= 512) ? 2 : 1;
  const int tm_need = (p.nbuf - 1) * p.buf_cols + tm_one;
  p.tmem_cols = 32; while (p.tmem_cols < tm_need) p.tmem_cols <<= 1;
  p.bias = a->bias; p.slope = a->slope; p.y = a->y_nchw; p.y_cl2 = reinterpret_cast<__nv_bfloat16*>(a->y_cl2);
  if (a->y_cs != 0 || a->y_ws != 0) {
    p.y_bs = a->y_bs; p.y_cs = a->y_cs; p.y_hs = a->y_hs; p.y_ws = a->y_ws; p.y_off = a->y_off; p.y_lim = a->y_lim;
  } else {
    const int64_t plane = (int64_t)a->Ho * a->Wo;
    p.y_bs = (int64_t)a->Cout2 * plane; p.y_cs = plane; p.y_hs = a->Wo; p.y_ws = 1; p.y_off = 0; p.y_lim = 0;
  }
  p.y_scale = a->y_scale;
  p.y_vec4 = (p.y != nullptr && p.y_cs == 1 && p.y_lim == 0 && p.y_scale == nullptr && a->Cout2 % 4 == 0 &&
              p.y_bs % 4 == 0 && p.y_hs % 4 == 0 && p.y_ws % 4 == 0 && p.y_off % 4 == 0 &&
              reinterpret_cast<uintptr_t>(p.y) % 16 == 0) ? 1 : 0;
  if (p.y_vec4 && a->Cout2 % 8 == 0 && p.y_bs % 8 == 0 && p.y_hs % 8 == 0 && p.y_ws % 8 == 0 && p.y_off % 8 == 0 &&
      reinterpret_cast<uintptr_t>(p.y) % 32 == 0)
    p.y_vec4 = 2;
  p.kblocks[0] = a->src_Cp[0] / 64; p.kcol0[0] = 0;
  p.kblocks[1] = a->nsrc == 2 ? a->src_Cp[1] / 64 : 0; p.kcol0[1] = a->src_Cp[0];
  for (int i = 0; i < 2; ++i) {
    // real channels of source i (0 = unknown: every channel of the pitch): pad channels are zero in the activations
    // AND in the prepared weights, so whole 16-channel k-steps of padding are not issued
    const int Cp = i < a->nsrc ? a->src_Cp[i] : 0, Cr = i < a->nsrc ? a->src_C[i] : 0;
    int last = 64;
    if (Cr > 0 && Cr <= Cp && Cr > Cp - 64) last = Cr - (Cp - 64);
    p.klast[i] = 2 * ((last + 15) / 16);
  }
  int64_t ntiles = 0;
  for (int ci = 0; ci < p.ncls; ++ci) {
    GClass& c = p.cls[ci];
    c.tiles_h = (c.Hq_out + tc.nacc * tc.TH - 1) / (tc.nacc * tc.TH);
    c.tiles_w = (c.Wq_out + tc.TW - 1) / tc.TW;
    c.tile_begin = (int)ntiles;
    ntiles += (int64_t)c.tiles_h * c.tiles_w * a->B;
  }
  if (ntiles <= 0) return SE_OK;
  if (ntiles > 2147483647LL) return fail(SE_ERR_UNSUPPORTED, "%s: too many tiles", nm);
  p.ntiles = (int)ntiles;

  CUtensorMap tmX[2], tmW;
  for (int i = 0; i < 2; ++i) {
    const int j = i < a->nsrc ? i : 0;
    const uint64_t Cp = (uint64_t)a->src_Cp[j];
    // a source smaller than the logical extent reads as zero beyond its own edge (TMA out-of-bounds fill):
    // this is the F.pad-to-the-skip's-size of the DCUNet / FRCRN decoders
    const int sH = a->src_H[j] > 0 ? a->src_H[j] : a->H, sW = a->src_W[j] > 0 ? a->src_W[j] : a->W;
    uint64_t dims[4] = {Cp, (uint64_t)sW, (uint64_t)sH, (uint64_t)(a->B * 2)};
    uint64_t str[3] = {Cp * 2, (uint64_t)a->src_Wpitch[j] * Cp * 2,
                       (uint64_t)a->src_Hpitch[j] * a->src_Wpitch[j] * Cp * 2};
    uint32_t box[4] = {64, (uint32_t)(p.TWb * p.col_mul), (uint32_t)(rows * p.row_mul), 1};
    uint32_t es[4] = {1, (uint32_t)p.col_mul, (uint32_t)p.row_mul, 1};
    rc = make_tmap_bf16(&tmX[i], a->src[j], 4, dims, str, box, es, true);
    if (rc != SE_OK) return rc;
  }
  {
    // wq is [taps][2][npad][Cpt], viewed as [taps * units][unit_rows][Cpt]
    const uint64_t Cpt = (uint64_t)a->src_Cp[0] + (a->nsrc == 2 ? a->src_Cp[1] : 0);
    const uint64_t rows_u = (uint64_t)p.unit_rows, nu = (uint64_t)a->kh * a->kw * (concat ? 1 : 2 * tc.nsplit);
    uint64_t dims[3] = {Cpt, rows_u, nu};
    uint64_t str[2] = {Cpt * 2, Cpt * 2 * rows_u};
    uint32_t box[3] = {64, (uint32_t)rows_u, 1};
    uint32_t es[3] = {1, 1, 1};
    rc = make_tmap_bf16(&tmW, a->wq, 3, dims, str, box, es, true);
    if (rc != SE_OK) return rc;
  }
  const int smem = p.SA * p.a_slot + p.SB * p.b_unit + 1024;
  static se::PerDevice<int> smem_attr_pd(0);
  int& smem_attr = smem_attr_pd.get();
  if (smem > smem_attr) {
    SE_CUDA(cudaFuncSetAttribute(cconv_gemm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    SE_CUDA(cudaFuncSetAttribute(cconv_gemm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    smem_attr = smem;
  }
  if (dbg & 4)
    fprintf(stderr, "se_cconv_gemm: tr=%d Cp=%d+%d N=%d k=%dx%d s=%dx%d -> TH=%d TW=%d nacc=%d concat=%d nsplit=%d SA=%d SB=%d rows=%d smem=%d tmem=%d nbuf=%d tiles=%lld ncls=%d\n",
            a->transposed, a->src_Cp[0], a->nsrc == 2 ? a->src_Cp[1] : 0, npad, a->kh, a->kw, a->sh, a->sw, tc.TH, tc.TW,
            tc.nacc, p.concat, p.nsplit, tc.SA, tc.SB, rows, smem, p.tmem_cols, p.nbuf, (long long)ntiles, p.ncls);
  if ((dbg & 4) && span_w > 0) fprintf(stderr, "se_cconv_gemm:   column-shared halo box: TWb=%d (span %d)\n", p.TWb, span_w);
  const unsigned grid = (unsigned)(ntiles < num_sms() ? ntiles : num_sms());
  static const char* prof_env = getenv("SE_GEMM_PROF");      // tuning aid: synchronous, prints per-role cycle counters
  p.prof = nullptr;
  if (prof_env && atoi(prof_env) != 0) {
    static unsigned long long* d_prof = nullptr;
    if (d_prof == nullptr) SE_CUDA(cudaMalloc(&d_prof, 16 * sizeof(unsigned long long)));
    SE_CUDA(cudaMemsetAsync(d_prof, 0, 16 * sizeof(unsigned long long), s));
    p.prof = d_prof;
  }
  if (concat) cconv_gemm_kernel<true><<<grid, kGemmThreads, smem, s>>>(tmX[0], tmX[1], tmW, p);
  else cconv_gemm_kernel<false><<<grid, kGemmThreads, smem, s>>>(tmX[0], tmX[1], tmW, p);
  if (p.prof != nullptr) {
    unsigned long long h[16];
    SE_CUDA(cudaStreamSynchronize(s));
    SE_CUDA(cudaMemcpy(h, p.prof, sizeof(h), cudaMemcpyDeviceToHost));
    const double g = (double)grid, tl = (double)ntiles / g;
    fprintf(stderr, "se_cconv_gemm prof: N=%d Cp=%d k=%dx%d tiles/cta=%.1f | per tile kcyc: producer %.1f (wait a_empty %.1f, b_empty %.1f) | "
            "mma %.1f (wait t_empty %.1f, a_full %.1f, b_full %.1f) | epilogue %.1f (wait t_full %.1f)\n",
            npad, a->src_Cp[0] + (a->nsrc == 2 ? a->src_Cp[1] : 0), a->kh, a->kw, tl,
            h[0] / g / tl / 1e3, h[1] / g / tl / 1e3, h[2] / g / tl / 1e3, h[3] / g / tl / 1e3, h[4] / g / tl / 1e3,
            h[5] / g / tl / 1e3, h[6] / g / tl / 1e3, h[7] / g / tl / 1e3, h[8] / g / tl / 1e3);
  }
  return check_launch("cconv_gemm_kernel");
}

}  // namespace se

// ---- nn.Module path: NCHW fp32 in / out through a caller-provided workspace -------------------
namespace se {

static void module_plan(int64_t B, int Cin2, int H, int W, int Cout2, int taps, int& Cp, int64_t& xs_bytes,
                        int64_t& wq_bytes) {
  Cp = round_up(Cin2, 64);
  xs_bytes = (B * 2 * (int64_t)H * W * Cp * 2 + 1023) / 1024 * 1024;
  wq_bytes = ((int64_t)taps * 2 * round_up(Cout2, 16) * Cp * 2 + 1023) / 1024 * 1024;
}

bool cconv_umma_supported(bool transposed, int64_t B, const ConvGeom& g) {
  (void)transposed;
  return g.Cout2 <= 256 && g.kh * g.kw <= kMaxTaps && B <= 32767 && g.Cin2 >= 16;
}

int64_t cconv_umma_workspace_bytes(bool transposed, int64_t B, int Cin2, int H, int W, int Cout2, int kh,
                                   int kw, int sh, int sw) {
  (void)transposed; (void)sh; (void)sw;
  int Cp; int64_t xs, wq;
  module_plan(B, Cin2, H, W, Cout2, kh * kw, Cp, xs, wq);
  return xs + wq + 1024;
}

int cconv_fwd_umma(bool transposed, const float* x, const float* packed, const float* bias, int64_t B,
                   const ConvGeom& g, float* y, void* ws, int64_t ws_bytes, cudaStream_t s) {
  int Cp; int64_t xs_bytes, wq_bytes;
  module_plan(B, g.Cin2, g.H, g.W, g.Cout2, g.kh * g.kw, Cp, xs_bytes, wq_bytes);
  uintptr_t base = (reinterpret_cast<uintptr_t>(ws) + 1023) & ~uintptr_t(1023);
  if (ws == nullptr || (int64_t)(base - reinterpret_cast<uintptr_t>(ws)) + xs_bytes + wq_bytes > ws_bytes)
    return fail(SE_ERR_WORKSPACE, "se_cconv*_fwd: workspace of %lld bytes needed (se_cconv_workspace_bytes)",
                (long long)(xs_bytes + wq_bytes + 1024));
  void* xs = reinterpret_cast<void*>(base);
  void* wq = reinterpret_cast<void*>(base + xs_bytes);
  int rc = se_cl2_from_nchw(x, B, g.Cin2, g.H, g.W, xs, Cp, s);
  if (rc != SE_OK) return rc;
  rc = se_cconv_prep_weight(packed, g.Cout2, g.Cin2, g.kh * g.kw, nullptr, Cp, wq, s);
  if (rc != SE_OK) return rc;
  se_cconv_args a{};
  a.transposed = transposed ? 1 : 0; a.B = B; a.nsrc = 1;
  a.src[0] = xs; a.src_Cp[0] = Cp; a.H = g.H; a.W = g.W; a.src_Hpitch[0] = g.H; a.src_Wpitch[0] = g.W;
  a.Cout2 = g.Cout2; a.kh = g.kh; a.kw = g.kw; a.sh = g.sh; a.sw = g.sw; a.ph = g.ph; a.pw = g.pw;
  a.Ho = g.Ho; a.Wo = g.Wo; a.wq = wq; a.bias = bias; a.slope = 1.f; a.y_nchw = y;
  return se_cconv_gemm(&a, s);
}

}  // namespace se

// =================================================================================== C ABI
extern "C" int se_cconv_gemm(const se_cconv_args* a, void* stream) {
  SE_REQUIRE(a != nullptr, "se_cconv_gemm: null args");
  SE_REQUIRE(a->nsrc == 1 || a->nsrc == 2, "se_cconv_gemm: nsrc must be 1 or 2");
  SE_REQUIRE(a->src[0] && (a->nsrc == 1 || a->src[1]) && a->wq, "se_cconv_gemm: null pointer");
  SE_REQUIRE(a->y_nchw || a->y_cl2, "se_cconv_gemm: no output requested");
  SE_REQUIRE(a->B >= 0 && a->H > 0 && a->W > 0 && a->Cout2 > 0 && a->kh > 0 && a->kw > 0 && a->sh > 0 &&
             a->sw > 0 && a->Ho > 0 && a->Wo > 0, "se_cconv_gemm: bad geometry");
  SE_REQUIRE(a->y_cl2 == nullptr ||
             (a->y_Cp >= ((a->fold_h > 1 || a->fold_w > 1) ? a->fold_cout2 : a->Cout2) && a->y_Cp % 64 == 0),
             "se_cconv_gemm: y_Cp must be a multiple of 64 >= Cout2");
  for (int i = 0; i < a->nsrc; ++i)
  {
    const int sH = a->src_H[i] > 0 ? a->src_H[i] : a->H, sW = a->src_W[i] > 0 ? a->src_W[i] : a->W;
    SE_REQUIRE(sH <= a->H && sW <= a->W, "se_cconv_gemm: a source is larger than the logical extent");
    SE_REQUIRE(a->src_Hpitch[i] >= sH && a->src_Wpitch[i] >= sW, "se_cconv_gemm: source pitch < extent");
  }
  if (a->B == 0) return SE_OK;
  return se::cconv_gemm_launch(a, (cudaStream_t)stream);
}

extern "C" int se_cl2_from_nchw(const float* x, int64_t B, int C, int H, int W, void* cl2, int Cp, void* stream) {
  SE_REQUIRE(x && cl2, "se_cl2_from_nchw: null pointer");
  SE_REQUIRE(B >= 0 && C > 0 && H > 0 && W > 0 && Cp >= C && Cp % 64 == 0, "se_cl2_from_nchw: bad dims");
  if (B == 0) return SE_OK;
  dim3 grid((unsigned)((W + 31) / 32), (unsigned)H, (unsigned)(B * (Cp / 64)));
  if (grid.y > 65535 || grid.z > 65535) return se::fail(SE_ERR_UNSUPPORTED, "se_cl2_from_nchw: tensor too large");
  se::cl2_from_nchw_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      x, (int64_t)C * H * W, (int64_t)H * W, (int64_t)W, reinterpret_cast<__nv_bfloat16*>(cl2), C, H, W, Cp);
  return se::check_launch("se_cl2_from_nchw");
}

// Same conversion from an fp32 tensor addressed as x[b*xs_b + c*xs_c + h*xs_h + w] (unit stride
// along w).  The iSTFT uses it to view a [B, 2F, T] spectrogram as one image whose rows are the
// utterances: B=1, C=2F (xs_c = T), H=batch (xs_h = 2F*T).
extern "C" int se_cl2_from_strided(const float* x, int64_t xs_b, int64_t xs_c, int64_t xs_h, int64_t B, int C,
                                   int H, int W, void* cl2, int Cp, void* stream) {
  SE_REQUIRE(x && cl2, "se_cl2_from_strided: null pointer");
  SE_REQUIRE(B >= 0 && C > 0 && H > 0 && W > 0 && Cp >= C && Cp % 64 == 0, "se_cl2_from_strided: bad dims");
  if (B == 0) return SE_OK;
  dim3 grid((W + 31) / 32, H, (unsigned)(B * (Cp / 64)));
  if (grid.y > 65535 || grid.z > 65535) return se::fail(SE_ERR_UNSUPPORTED, "se_cl2_from_strided: tensor too large");
  se::cl2_from_nchw_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      x, xs_b, xs_c, xs_h, reinterpret_cast<__nv_bfloat16*>(cl2), C, H, W, Cp);
  return se::check_launch("se_cl2_from_strided");
}

// STFT framing for the tensor-core path: reflect-pad (conv_stft.py:60-62) and cut the signal into
// hop-sample pixels, split into bf16 hi|lo planes: cl2[s][b][q][col] = xpad[b][q*hop + col]
// (col < hop, else 0) -- a CL2 image with one row per utterance and `hop` channels, so that the
// framed convolution (conv_stft.py:63) is a 1 x ceil(win/hop) convolution over it.
namespace se {
__global__ void __launch_bounds__(256)
stft_frames_cl2_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int64_t B, int64_t L, int hop,
                       int pad, int Q, int Cp) {
  const int vec = Cp / 8;
  const int64_t total = B * Q * vec;
  const int64_t plane = B * Q * (int64_t)Cp;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int v8 = (int)(i % vec);
    const int64_t r = i / vec;
    const int q = (int)(r % Q);
    const int64_t b = r / Q;
    const float* xb = x + b * L;
    __align__(16) __nv_bfloat16 hi[8], lo[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int col = v8 * 8 + j;
      float val = 0.f;
      if (col < hop) {
        int64_t n = (int64_t)q * hop + col - pad;
        if (n < 0) n = -n;
        if (n >= L) n = 2 * (L - 1) - n;
        if (n >= 0 && n < L && (int64_t)q * hop + col < L + 2 * (int64_t)pad) val = __ldg(xb + n);
      }
      split_bf16(val, hi[j], lo[j]);
    }
    const int64_t off = (b * Q + q) * Cp + v8 * 8;
    *reinterpret_cast<uint4*>(out + off) = *reinterpret_cast<const uint4*>(hi);
    *reinterpret_cast<uint4*>(out + plane + off) = *reinterpret_cast<const uint4*>(lo);
  }
}
}  // namespace se

// Row-major fp32 matrix x[M][ld] (K <= ld columns used) -> CL2 image with one pixel per row:
// cl2[s][m][col] (col < K, zero above).  Feeds plain linear layers (LSTM input projections,
// ComplexLinear: /root/reference/models/modules/complex_nn.py:93-145) to the GEMM engine.
namespace se {
__global__ void __launch_bounds__(256)
cl2_from_rows_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int64_t M, int K, int64_t ld, int Cp) {
  const int vec = Cp / 8;
  const int64_t total = M * vec, plane = M * (int64_t)Cp;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int v8 = (int)(i % vec);
    const int64_t m = i / vec;
    const float* xr = x + m * ld + v8 * 8;
    __align__(16) __nv_bfloat16 hi[8], lo[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) split_bf16(v8 * 8 + j < K ? __ldg(xr + j) : 0.f, hi[j], lo[j]);
    const int64_t off = m * Cp + v8 * 8;
    *reinterpret_cast<uint4*>(out + off) = *reinterpret_cast<const uint4*>(hi);
    *reinterpret_cast<uint4*>(out + plane + off) = *reinterpret_cast<const uint4*>(lo);
  }
}
}  // namespace se

extern "C" int se_cl2_from_rows(const float* x, int64_t M, int K, int64_t ld, void* cl2, int Cp, void* stream) {
  SE_REQUIRE(x && cl2, "se_cl2_from_rows: null pointer");
  SE_REQUIRE(M >= 0 && K > 0 && ld >= K && Cp >= K && Cp % 64 == 0, "se_cl2_from_rows: bad dims");
  if (M == 0) return SE_OK;
  const int64_t total = M * (Cp / 8);
  const unsigned blocks = (unsigned)std::min<int64_t>((total + 255) / 256, 148 * 16);
  se::cl2_from_rows_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, reinterpret_cast<__nv_bfloat16*>(cl2), M, K, ld, Cp);
  return se::check_launch("se_cl2_from_rows");
}

extern "C" int se_stft_frames_cl2(const float* x, int64_t B, int64_t L, int hop, int pad, int Q, int Cp, void* cl2,
                                  void* stream) {
  SE_REQUIRE(x && cl2, "se_stft_frames_cl2: null pointer");
  SE_REQUIRE(B >= 0 && L > 0 && hop > 0 && pad >= 0 && pad < L && Q > 0 && Cp >= hop && Cp % 64 == 0,
             "se_stft_frames_cl2: bad dims");
  if (B == 0) return SE_OK;
  const int64_t total = B * Q * (Cp / 8);
  const unsigned blocks = (unsigned)std::min<int64_t>((total + 255) / 256, 148 * 16);
  se::stft_frames_cl2_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(
      x, reinterpret_cast<__nv_bfloat16*>(cl2), B, L, hop, pad, Q, Cp);
  return se::check_launch("se_stft_frames_cl2");
}

extern "C" int se_cl2_to_nchw(const void* cl2, int64_t B, int C, int H, int W, int Cp, float* y, void* stream) {
  SE_REQUIRE(y && cl2, "se_cl2_to_nchw: null pointer");
  SE_REQUIRE(B >= 0 && C > 0 && H > 0 && W > 0 && Cp >= C && Cp % 64 == 0, "se_cl2_to_nchw: bad dims");
  if (B == 0) return SE_OK;
  dim3 grid((unsigned)((W + 31) / 32), (unsigned)H, (unsigned)(B * (Cp / 64)));
  if (grid.y > 65535 || grid.z > 65535) return se::fail(SE_ERR_UNSUPPORTED, "se_cl2_to_nchw: tensor too large");
  se::cl2_to_nchw_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const __nv_bfloat16*>(cl2), y, C, H, W, Cp);
  return se::check_launch("se_cl2_to_nchw");
}

extern "C" int64_t se_cconv_wq_elems(int Cout2, int Cp_total, int taps) {
  return (int64_t)taps * 2 * ((Cout2 + 15) / 16 * 16) * Cp_total;
}

extern "C" int se_cconv_prep_weight(const float* packed, int Cout2, int Cin2, int taps, const int* chan_map,
                                    int Cp_total, void* wq, void* stream) {
  SE_REQUIRE(packed && wq, "se_cconv_prep_weight: null pointer");
  SE_REQUIRE(Cout2 > 0 && Cin2 > 0 && taps > 0 && Cp_total > 0 && Cp_total % 64 == 0,
             "se_cconv_prep_weight: bad dims");
  const int npad = (Cout2 + 15) / 16 * 16;
  int64_t total = (int64_t)taps * npad * Cp_total;
  se::cconv_prep_w_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      packed, chan_map, reinterpret_cast<__nv_bfloat16*>(wq), Cout2, Cin2, taps, npad, Cp_total);
  return se::check_launch("se_cconv_prep_weight");
}

// ---- allocation-free variant of se_cconv2d_fwd (no counterpart in the reference; see include/se_b200.h) ----
// The tensor engine's module path stages the input as CL2 and the weight as bf16 hi|lo slabs in scratch memory.
// Here that scratch is a file-scope device array instead of a caller-provided workspace, so that callers which
// may not allocate (CUDA-graph capture, serving loops) have an entry that never calls cudaMalloc / cudaFree.
// `g_static_ws` below is, in HOST code, only a handle for the symbol: the device address is looked up once per
// device with cudaGetSymbolAddress (that device current) and that pointer is what the kernels receive.
// Both staging kernels overwrite every byte the GEMM later reads (pad channels included), so nothing has to be
// cleared between calls; calls that overlap on the GPU (two streams, two threads) share the array and race.
static __device__ __align__(256) unsigned char g_static_ws[SE_STATIC_WS_BYTES];

namespace se {
static int static_ws_ptr(void** out) {
  constexpr int kMaxDev = 64;
  static std::atomic<void*> ptrs[kMaxDev];
  int dev = -1;
  SE_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDev) return fail(SE_ERR_UNSUPPORTED, "se_cconv2d_fwd_static: device ordinal %d >= %d", dev, kMaxDev);
  void* p = ptrs[dev].load(std::memory_order_acquire);
  if (p == nullptr) {
    SE_CUDA(cudaGetSymbolAddress(&p, g_static_ws));      // same value from every thread: a benign race
    if (p == nullptr) return fail(SE_ERR_CUDA, "se_cconv2d_fwd_static: cudaGetSymbolAddress returned NULL");
    ptrs[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return SE_OK;
}
}  // namespace se

extern "C" int se_cconv2d_fwd_static(const float* x, const float* packed, const float* bias, int64_t B, int Cin2,
                                     int H, int W, int Cout2, int kh, int kw, int sh, int sw, int ph, int pw,
                                     float* y, int Ho, int Wo, void* stream) {
  const char* nm = "se_cconv2d_fwd_static";
  SE_REQUIRE(x && packed && y, "%s: null pointer", nm);
  SE_REQUIRE(B >= 0 && Cin2 > 0 && Cout2 > 0 && (Cin2 % 2) == 0 && (Cout2 % 2) == 0,
             "%s: channel counts must be positive and even", nm);
  SE_REQUIRE(H > 0 && W > 0 && kh > 0 && kw > 0 && sh > 0 && sw > 0 && ph >= 0 && pw >= 0 && Ho > 0 && Wo > 0,
             "%s: bad geometry", nm);
  SE_REQUIRE(Ho == (H + 2 * ph - kh) / sh + 1 && Wo == (W + 2 * pw - kw) / sw + 1,
             "%s: output size does not match geometry", nm);
  if (B == 0) return SE_OK;
  se::ConvGeom g{Cin2, H, W, Cout2, kh, kw, sh, sw, ph, pw, Ho, Wo};
  if (!se::cconv_umma_supported(false, B, g))
    return se::fail(SE_ERR_UNSUPPORTED, "%s: shape outside the tensor engine (Cin2 >= 16, Cout2 <= 256, taps <= %d)", nm,
                    se::kMaxTaps);
  const int64_t need = se::cconv_umma_workspace_bytes(false, B, Cin2, H, W, Cout2, kh, kw, sh, sw);
  if (need > (int64_t)SE_STATIC_WS_BYTES)
    return se::fail(SE_ERR_WORKSPACE, "%s: this problem needs %lld bytes of scratch, the static array holds %lld "
                    "(SE_STATIC_WS_BYTES); use se_cconv2d_fwd with a caller-provided workspace", nm, (long long)need,
                    (long long)SE_STATIC_WS_BYTES);
  const char* prof = getenv("SE_GEMM_PROF");              // the profiling aid allocates a counter buffer
  if (prof && atoi(prof) != 0) return se::fail(SE_ERR_UNSUPPORTED, "%s: unset SE_GEMM_PROF (it allocates)", nm);
  void* ws = nullptr;
  int rc = se::static_ws_ptr(&ws);
  if (rc != SE_OK) return rc;
  return se::cconv_fwd_umma(false, x, packed, bias, B, g, y, ws, (int64_t)SE_STATIC_WS_BYTES, (cudaStream_t)stream);
}

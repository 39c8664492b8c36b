// Recurrent half of an LSTM layer as one persistent kernel.
//
// The reference runs ComplexLSTM as four nn.LSTM passes (/root/reference/models/modules/
// complex_nn.py:115-145); on a GPU the time-step loop of each pass is a chain of ~T tiny launches /
// grid syncs.  Here the input projection  P[n,t,:] = W_ih x[n,t] + b_ih + b_hh  is hoisted into one
// GEMM by the caller, and this kernel walks the T steps of a few batch rows entirely on one SM:
//   gates = P[n,t] + W_hh h_{t-1};  i,f,o = sigmoid, g = tanh (gate order i|f|g|o, torch.nn.LSTM);
//   c_t = f c_{t-1} + i g;  h_t = o tanh(c_t)
// Batch rows never interact, so there is no inter-CTA synchronisation at all.  Three kernels:
//   lstm_rec_kernel          one thread per gate column (4H threads) keeps its W_hh row on chip for the whole
//                            sequence -- KR values in registers, the other H-KR in shared memory, column-major so a
//                            warp reads consecutive words -- h_{t-1} broadcast from shared memory, two
//                            __syncthreads per step.  Used for H = 64.
//   lstm_rec2_kernel         two gate columns per thread, the four gates of a unit in one warp, one barrier per
//                            step: the production kernel for H = 128 at more than 74 row jobs.
//   lstm_rec_cluster_kernel  2-CTA cluster, all of W_hh in registers, h halves exchanged through DSMEM: H = 128 with
//                            at most (#SMs / 2) row jobs.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdlib>
#include "common.cuh"
#include "../../include/se_b200.h"

namespace se {

// full-precision libm forms: the state feeds back T times, so the fast intrinsics' few-ulp errors (and
// the cancellation of 1 - 2/(e^2x + 1) near 0) would show up at the 1e-5 parity bar
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }
__device__ __forceinline__ float tanhf_(float x) { return tanhf(x); }

template <int H, int KR, int R>
__global__ void __launch_bounds__(4 * H, 1)
lstm_rec_kernel(const float* __restrict__ pre, const float* __restrict__ w_hh0, const float* __restrict__ w_hh1,
                float* __restrict__ h_out0, float* __restrict__ h_out1, int64_t N, int64_t T, int64_t pre_ld) {
  constexpr int G = 4 * H, KS = H - KR;
  // blockIdx.y selects one of two LSTMs that share the pre-gate matrix (columns [0,4H) and [4H,8H)):
  // the real / imaginary LSTMs of a ComplexLSTM layer run side by side in one launch
  const float* __restrict__ w_hh = blockIdx.y ? w_hh1 : w_hh0;
  float* __restrict__ h_out = blockIdx.y ? h_out1 : h_out0;
  pre += blockIdx.y * G;
  extern __shared__ __align__(16) float smem[];
  float* w_s = smem;                 // [KS][G]
  float* h_s = w_s + KS * G;         // [R][H]
  float* g_s = h_s + R * H;          // [R][G]
  const int j = threadIdx.x;
  const int gate = j / H;            // warp-uniform (H is a multiple of 32)

  float w[KR];
#pragma unroll
  for (int k = 0; k < KR; ++k) w[k] = __ldg(w_hh + (int64_t)j * H + k);
  for (int k = 0; k < KS; ++k) w_s[k * G + j] = __ldg(w_hh + (int64_t)j * H + KR + k);

  for (int64_t row0 = (int64_t)blockIdx.x * R; row0 < N; row0 += (int64_t)gridDim.x * R) {
    float c[R];
    const float* prow[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      c[r] = 0.f;
      const int64_t row = row0 + r < N ? row0 + r : N - 1;       // tail rows recompute the last row (never stored)
      prow[r] = pre + row * T * pre_ld + j;
    }
    for (int i = j; i < R * H; i += G) h_s[i] = 0.f;
    float nxt[R];
#pragma unroll
    for (int r = 0; r < R; ++r) nxt[r] = __ldg(prow[r]);
    __syncthreads();
    for (int64_t t = 0; t < T; ++t) {
      float acc[R][2];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        acc[r][0] = nxt[r];
        acc[r][1] = 0.f;
        if (t + 1 < T) nxt[r] = __ldg(prow[r] + (t + 1) * pre_ld);
      }
#pragma unroll
      for (int k = 0; k < KR; k += 4) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const float4 hv = *reinterpret_cast<const float4*>(h_s + r * H + k);
          acc[r][0] = fmaf(w[k], hv.x, acc[r][0]);
          acc[r][1] = fmaf(w[k + 1], hv.y, acc[r][1]);
          acc[r][0] = fmaf(w[k + 2], hv.z, acc[r][0]);
          acc[r][1] = fmaf(w[k + 3], hv.w, acc[r][1]);
        }
      }
#pragma unroll 4
      for (int k = 0; k < KS; k += 4) {
        const float w0 = w_s[(k + 0) * G + j], w1 = w_s[(k + 1) * G + j];
        const float w2 = w_s[(k + 2) * G + j], w3 = w_s[(k + 3) * G + j];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const float4 hv = *reinterpret_cast<const float4*>(h_s + r * H + KR + k);
          acc[r][0] = fmaf(w0, hv.x, acc[r][0]);
          acc[r][1] = fmaf(w1, hv.y, acc[r][1]);
          acc[r][0] = fmaf(w2, hv.z, acc[r][0]);
          acc[r][1] = fmaf(w3, hv.w, acc[r][1]);
        }
      }
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float a = acc[r][0] + acc[r][1];
        g_s[r * G + j] = gate == 2 ? tanhf_(a) : sigmoidf_(a);
      }
      __syncthreads();
      if (j < H) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const float* g = g_s + r * G;
          c[r] = g[H + j] * c[r] + g[j] * g[2 * H + j];
          const float h = g[3 * H + j] * tanhf_(c[r]);
          h_s[r * H + j] = h;
          if (row0 + r < N) h_out[((row0 + r) * T + t) * H + j] = h;
        }
      }
      __syncthreads();
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Two gate columns per thread (2H threads), the production kernel for H = 128.  ncu on the one-column kernel above
// (profiles/r1f_lstm_full_summary.txt): the shared-memory wavefront pipe is the limiter (79 % busy on the occupied
// SMs) -- per step and CTA 16 warps x 64 broadcast LDS.128 of h (two wavefronts each: a broadcast wavefront carries
// 8 bytes) + the 128 KB of weights that do not fit the register file.  Halving the thread count halves the h
// wavefronts (every h value now feeds two columns) and leaves 2 x KR = 192 weight registers per thread, so only
// (H - KR) / H = 1/4 of the matrix is re-read from shared memory each step.
// The four gates of a hidden unit live in ONE warp: lanes 0-15 of warp w own (i, g) of units 16w .. 16w+15, lanes
// 16-31 own (f, o) of the same units.  The cell update then needs three shuffles instead of a round trip through
// shared memory, every lane runs exactly one tanh(c) per row pair (upper half: even rows, lower half: odd rows),
// h is double-buffered and a step costs ONE __syncthreads.  sigmoid(x) is evaluated as 0.5 + 0.5 tanh(x / 2) so
// that the mixed-gate warps do not diverge.  (64 x 4 s DCCRN batch: 1.40 -> 1.06 -> 0.90 ms per launch.)
template <int H, int KR, int R>
__global__ void __launch_bounds__(2 * H, 1)
lstm_rec2_kernel(const float* __restrict__ pre, const float* __restrict__ w_hh0, const float* __restrict__ w_hh1,
                 float* __restrict__ h_out0, float* __restrict__ h_out1, int64_t N, int64_t T, int64_t pre_ld) {
  constexpr int NT = 2 * H, KS = H - KR, RP = (R + 1) / 2;
  const float* __restrict__ w_hh = blockIdx.y ? w_hh1 : w_hh0;
  float* __restrict__ h_out = blockIdx.y ? h_out1 : h_out0;
  pre += blockIdx.y * 4 * H;
  extern __shared__ __align__(16) float smem[];
  // weights that do not fit the register file: [KS/2][NT] float4 = (wa[k], wa[k+1], wb[k], wb[k+1]) of thread j
  float4* w_s = reinterpret_cast<float4*>(smem);
  float* h_s = smem + 2 * KS * NT;                   // [2][R][H]
  const int j = threadIdx.x, lane = j & 31;
  const bool upper = lane >= 16;
  const int u = 16 * (j >> 5) + (lane & 15);
  const int ca = (upper ? H : 0) + u;                // i | f
  const int cb = ca + 2 * H;                         // g | o
  const float bs = upper ? 0.5f : 1.f, bo = upper ? 0.5f : 0.f;   // act(b) = bo + bs * tanh(bs * b)

  // Packed FP32: a 3-register FFMA issues at half rate on sm_100 (scripts/ffma2_rate.cu: 73 FMA lanes / clk / SM against
  // 113 for FFMA2), and this loop is nothing but such FMAs.  The two accumulator chains of a column (even k, odd k) are
  // the two halves of ONE fma.rn.f32x2: weights (w[k], w[k+1]) and h (h[k], h[k+1]) are natural register pairs, the
  // per-lane arithmetic and its order are those of the scalar loop (bit-identical results), the FMA instruction count halves.
  float2 wa[KR / 2], wb[KR / 2];
#pragma unroll
  for (int k = 0; k < KR; k += 2) {
    wa[k / 2] = make_float2(__ldg(w_hh + (int64_t)ca * H + k), __ldg(w_hh + (int64_t)ca * H + k + 1));
    wb[k / 2] = make_float2(__ldg(w_hh + (int64_t)cb * H + k), __ldg(w_hh + (int64_t)cb * H + k + 1));
  }
  for (int k = 0; k < KS; k += 2)
    w_s[(k / 2) * NT + j] = make_float4(__ldg(w_hh + (int64_t)ca * H + KR + k), __ldg(w_hh + (int64_t)ca * H + KR + k + 1),
                                        __ldg(w_hh + (int64_t)cb * H + KR + k), __ldg(w_hh + (int64_t)cb * H + KR + k + 1));

  for (int64_t row0 = (int64_t)blockIdx.x * R; row0 < N; row0 += (int64_t)gridDim.x * R) {
    float c[RP];
#pragma unroll
    for (int q = 0; q < RP; ++q) c[q] = 0.f;
    const float* prow[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int64_t row = row0 + r < N ? row0 + r : N - 1;       // tail rows recompute the last row (never stored)
      prow[r] = pre + row * T * pre_ld;
    }
    __syncthreads();                                             // previous row block is done with h_s
    for (int i = j; i < R * H; i += NT) h_s[i] = 0.f;
    float nxa[R], nxb[R];
#pragma unroll
    for (int r = 0; r < R; ++r) { nxa[r] = __ldg(prow[r] + ca); nxb[r] = __ldg(prow[r] + cb); }
    __syncthreads();
    int cur = 0;
    for (int64_t t = 0; t < T; ++t) {
      const float* hc = h_s + cur * (R * H);
      float* hn = h_s + (cur ^ 1) * (R * H);
      float2 aa[R], ab[R];                                       // (even-k chain, odd-k chain)
#pragma unroll
      for (int r = 0; r < R; ++r) {
        aa[r] = make_float2(nxa[r], 0.f);
        ab[r] = make_float2(nxb[r], 0.f);
        if (t + 1 < T) {
          nxa[r] = __ldg(prow[r] + (t + 1) * pre_ld + ca);
          nxb[r] = __ldg(prow[r] + (t + 1) * pre_ld + cb);
        }
      }
#pragma unroll
      for (int k = 0; k < KR; k += 4) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const float4 hv = *reinterpret_cast<const float4*>(hc + r * H + k);
          const float2 h01 = make_float2(hv.x, hv.y), h23 = make_float2(hv.z, hv.w);
          aa[r] = __ffma2_rn(wa[k / 2], h01, aa[r]);
          aa[r] = __ffma2_rn(wa[k / 2 + 1], h23, aa[r]);
          ab[r] = __ffma2_rn(wb[k / 2], h01, ab[r]);
          ab[r] = __ffma2_rn(wb[k / 2 + 1], h23, ab[r]);
        }
      }
#pragma unroll 2
      for (int k = 0; k < KS; k += 4) {
        const float4 w01 = w_s[(k / 2) * NT + j], w23 = w_s[(k / 2 + 1) * NT + j];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const float4 hv = *reinterpret_cast<const float4*>(hc + r * H + KR + k);
          const float2 h01 = make_float2(hv.x, hv.y), h23 = make_float2(hv.z, hv.w);
          aa[r] = __ffma2_rn(make_float2(w01.x, w01.y), h01, aa[r]);
          aa[r] = __ffma2_rn(make_float2(w23.x, w23.y), h23, aa[r]);
          ab[r] = __ffma2_rn(make_float2(w01.z, w01.w), h01, ab[r]);
          ab[r] = __ffma2_rn(make_float2(w23.z, w23.w), h23, ab[r]);
        }
      }
      float ga[R], gb[R];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        ga[r] = 0.5f + 0.5f * tanhf_(0.5f * (aa[r].x + aa[r].y));            // i | f
        gb[r] = bo + bs * tanhf_(bs * (ab[r].x + ab[r].y));                 // g | o
      }
#pragma unroll
      for (int q = 0; q < RP; ++q) {
        constexpr unsigned FULL = 0xffffffffu;
        const int r0 = 2 * q, r1 = 2 * q + 1 < R ? 2 * q + 1 : 2 * q;
        // even row r0: the upper half (f, o) receives i*g; odd row r1: the lower half (i, g) receives f and o
        const float p0 = __shfl_xor_sync(FULL, ga[r0] * gb[r0], 16);
        const float f1 = __shfl_xor_sync(FULL, ga[r1], 16);
        const float o1 = __shfl_xor_sync(FULL, gb[r1], 16);
        const float cn = upper ? fmaf(ga[r0], c[q], p0) : fmaf(f1, c[q], ga[r1] * gb[r1]);
        const float og = upper ? gb[r0] : o1;
        const int r = upper ? r0 : 2 * q + 1;
        if (r < R) {
          c[q] = cn;
          const float h = og * tanhf_(cn);
          hn[r * H + u] = h;
          if (row0 + r < N) h_out[((row0 + r) * T + t) * H + u] = h;
        }
      }
      __syncthreads();
      cur ^= 1;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Cluster version (H = 128): the 512 gate columns of a row are split over a 2-CTA cluster -- CTA r owns
// hidden units [64r, 64r+64) x 4 gates = 256 columns x 128 k = 32 K weights, ALL of them in registers
// (512 threads x 64), so a step reads no weights from shared memory at all (the single-CTA kernel above
// is bound by exactly that: 128 KB of shared-memory weight reads + the h broadcast per step).
// Mapping: lane = k-slice of 4, warp = 16 columns; each thread does 16 x 4 FMAs on its own h slice (one
// conflict-free LDS.128), then a 16-shuffle reduce-scatter leaves one column sum per lane pair.  The
// new h halves are exchanged through distributed shared memory (st.shared::cluster into the peer's
// double-buffered h) and ONE cluster barrier per step.
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void st_peer_f32(const float* local_ptr, uint32_t peer, float v) {
  uint32_t la = static_cast<uint32_t>(__cvta_generic_to_shared(local_ptr)), ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(peer));
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(ra), "f"(v) : "memory");
}

template <int R>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(512, 1)
lstm_rec_cluster_kernel(const float* __restrict__ pre, const float* __restrict__ w_hh0, const float* __restrict__ w_hh1,
                        float* __restrict__ h_out0, float* __restrict__ h_out1, int64_t N, int64_t T, int64_t pre_ld) {
  constexpr int H = 128, G = 512;
  __shared__ __align__(16) float h_s[2][R][H];
  __shared__ float g_s[R][256];
  const float* __restrict__ w_hh = blockIdx.y ? w_hh1 : w_hh0;
  float* __restrict__ h_out = blockIdx.y ? h_out1 : h_out0;
  pre += blockIdx.y * G;
  const uint32_t rank = cluster_rank(), peer = rank ^ 1u;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int gate = warp >> 2;                                    // warp-uniform
  const int ubase = 64 * (int)rank + (warp & 3) * 16;            // first hidden unit of this warp's 16 columns
  // weights of columns j(c) = gate*128 + ubase + c, k-slice [4*lane, 4*lane+4)
  float w[16][4];
#pragma unroll
  for (int c = 0; c < 16; ++c) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(w_hh + (int64_t)(gate * H + ubase + c) * H + 4 * lane));
    w[c][0] = v.x; w[c][1] = v.y; w[c][2] = v.z; w[c][3] = v.w;
  }
  // after the reduce-scatter lane l holds column c_l (lanes 2m, 2m+1 hold the same one)
  const int c_l = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
  const int col_l = gate * H + ubase + c_l;                      // global gate column
  const int slot_l = gate * 64 + (warp & 3) * 16 + c_l;          // g_s slot: [gate][unit within this CTA]
  const bool loader = (lane & 1) == 0;
  const int64_t nclusters = gridDim.x / 2, cid = blockIdx.x / 2;

  for (int64_t row0 = cid * R; row0 < N; row0 += nclusters * R) {
    const float* prow[R];
    float nxt[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int64_t row = row0 + r < N ? row0 + r : N - 1;       // tail rows recompute the last row (never stored)
      prow[r] = pre + row * T * pre_ld + col_l;
      nxt[r] = loader ? __ldg(prow[r]) : 0.f;
    }
    float cst = 0.f;                                             // cell state of (row = tid/64, unit = tid%64)
    for (int i = tid; i < R * H; i += 512) h_s[0][i / H][i % H] = 0.f;
    cluster_barrier();
    int cur = 0;
    for (int64_t t = 0; t < T; ++t) {
      float pg[R];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        pg[r] = nxt[r];
        if (loader && t + 1 < T) nxt[r] = __ldg(prow[r] + (t + 1) * pre_ld);
      }
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const float4 hv = *reinterpret_cast<const float4*>(&h_s[cur][r][4 * lane]);
        float p[16];
#pragma unroll
        for (int c = 0; c < 16; ++c)
          p[c] = fmaf(w[c][3], hv.w, fmaf(w[c][2], hv.z, fmaf(w[c][1], hv.y, w[c][0] * hv.x)));
        // reduce-scatter over the 32 k-slices: 8 + 4 + 2 + 1 + 1 shuffles
        {
          const bool up = lane & 16;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const float send = up ? p[c] : p[c + 8], keep = up ? p[c + 8] : p[c];
            p[c] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
          }
        }
        {
          const bool up = lane & 8;
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const float send = up ? p[c] : p[c + 4], keep = up ? p[c + 4] : p[c];
            p[c] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
          }
        }
        {
          const bool up = lane & 4;
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            const float send = up ? p[c] : p[c + 2], keep = up ? p[c + 2] : p[c];
            p[c] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
          }
        }
        {
          const bool up = lane & 2;
          const float send = up ? p[0] : p[1], keep = up ? p[1] : p[0];
          p[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
        }
        p[0] += __shfl_xor_sync(0xffffffffu, p[0], 1);
        if (loader) {
          const float a = p[0] + pg[r];
          g_s[r][slot_l] = gate == 2 ? tanhf_(a) : sigmoidf_(a);
        }
      }
      __syncthreads();
      if (tid < 64 * R) {
        const int r = tid >> 6, u = tid & 63;
        const float* g = g_s[r];
        cst = g[64 + u] * cst + g[u] * g[128 + u];
        const float h = g[192 + u] * tanhf_(cst);
        float* dst = &h_s[cur ^ 1][r][64 * (int)rank + u];
        *dst = h;
        st_peer_f32(dst, peer, h);
        if (row0 + r < N) h_out[((row0 + r) * T + t) * H + 64 * (int)rank + u] = h;
      }
      cluster_barrier();                                         // both h halves of step t are in place everywhere
      cur ^= 1;
    }
  }
}

template <int R>
static int launch_lstm_cluster(const float* pre, int64_t pre_ld, const float* w0, const float* w1, float* h0, float* h1,
                               int nl, int64_t N, int64_t T, cudaStream_t s) {
  int64_t clusters = (N + R - 1) / R;
  const int64_t cap = num_sms() / 2 / nl;
  if (clusters > cap) clusters = cap;
  lstm_rec_cluster_kernel<R><<<dim3((unsigned)(2 * clusters), (unsigned)nl), 512, 0, s>>>(pre, w0, w1, h0, h1, N, T, pre_ld);
  return check_launch("lstm_rec_cluster_kernel");
}

template <int H, int KR, int R>
static int launch_lstm2(const float* pre, int64_t pre_ld, const float* w_hh0, const float* w_hh1, float* h_out0,
                        float* h_out1, int nl, int64_t N, int64_t T, cudaStream_t s) {
  const int smem = ((H - KR) * 4 * H + 2 * R * H) * (int)sizeof(float);
  static se::PerDevice<bool> attr_set_pd(false);
  bool& attr_set = attr_set_pd.get();
  if (!attr_set) {
    SE_CUDA(cudaFuncSetAttribute(lstm_rec2_kernel<H, KR, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  int64_t ctas = (N + R - 1) / R;
  if (ctas * nl > num_sms()) ctas = num_sms() / nl;
  lstm_rec2_kernel<H, KR, R><<<dim3((unsigned)ctas, (unsigned)nl), 2 * H, smem, s>>>(pre, w_hh0, w_hh1, h_out0, h_out1,
                                                                                      N, T, pre_ld);
  return check_launch("lstm_rec2_kernel");
}

template <int H, int KR, int R>
static int launch_lstm(const float* pre, int64_t pre_ld, const float* w_hh0, const float* w_hh1, float* h_out0,
                       float* h_out1, int nl, int64_t N, int64_t T, cudaStream_t s) {
  constexpr int G = 4 * H;
  const int smem = ((H - KR) * G + R * H + R * G) * (int)sizeof(float);
  static se::PerDevice<bool> attr_set_pd(false);
  bool& attr_set = attr_set_pd.get();
  if (!attr_set) {
    SE_CUDA(cudaFuncSetAttribute(lstm_rec_kernel<H, KR, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  int64_t ctas = (N + R - 1) / R;
  if (ctas * nl > num_sms()) ctas = num_sms() / nl;
  lstm_rec_kernel<H, KR, R><<<dim3((unsigned)ctas, (unsigned)nl), G, smem, s>>>(pre, w_hh0, w_hh1, h_out0, h_out1, N, T,
                                                                                pre_ld);
  return check_launch("lstm_rec_kernel");
}

}  // namespace se

extern "C" int se_lstm_rec_supported(int H) { return (H == 128 || H == 64) ? 1 : 0; }

template <int H, int KR>
static int dispatch_rows(const float* pre, int64_t pre_ld, const float* w0, const float* w1, float* h0, float* h1, int nl,
                         int64_t N, int64_t T, cudaStream_t s) {
  const int64_t sms = se::num_sms(), jobs = N * nl;
  if (H == 128) {
    // two columns per thread (see lstm_rec2_kernel); SE_LSTM_COLS=1 keeps the one-column kernel (tuning aid)
    static const char* cols = getenv("SE_LSTM_COLS");
    if (!(cols && atoi(cols) == 1)) {
      constexpr int H2 = 128, KR2 = 96;                          // only instantiated for H = 128
      if (jobs <= sms) return se::launch_lstm2<H2, KR2, 1>(pre, pre_ld, w0, w1, h0, h1, nl, N, T, s);
      if (jobs <= 2 * sms) return se::launch_lstm2<H2, KR2, 2>(pre, pre_ld, w0, w1, h0, h1, nl, N, T, s);
      return se::launch_lstm2<H2, KR2, 4>(pre, pre_ld, w0, w1, h0, h1, nl, N, T, s);
    }
  }
  if (jobs <= sms) return se::launch_lstm<H, KR, 1>(pre, pre_ld, w0, w1, h0, h1, nl, N, T, s);
  if (jobs <= 2 * sms) return se::launch_lstm<H, KR, 2>(pre, pre_ld, w0, w1, h0, h1, nl, N, T, s);
  return se::launch_lstm<H, KR, 4>(pre, pre_ld, w0, w1, h0, h1, nl, N, T, s);
}

static int lstm_rec_any(const float* pre, int64_t pre_ld, const float* w0, const float* w1, float* h0, float* h1, int nl,
                        int64_t N, int64_t T, int H, cudaStream_t s) {
  if (H == 128) {
    static const char* single = getenv("SE_LSTM_SINGLE");      // tuning aid: keep the single-CTA kernel
    if (single && atoi(single) != 0) return dispatch_rows<128, 64>(pre, pre_ld, w0, w1, h0, h1, nl, N, T, s);
    // The 2-CTA cluster kernel (all of W_hh in registers, h halves exchanged through DSMEM) was the small-batch path
    // while the single-CTA kernel issued scalar FFMAs; with the packed-FFMA2 time loop the single-CTA kernel has the
    // shorter step at every batch (DCCRN batch 1 / 8 / 16, 4 s: LSTM block 1.49 -> 1.12 ms, the cluster barrier +
    // DSMEM exchange cost ~0.9 us per step), so the cluster kernel now only runs on request (SE_LSTM_CLUSTER=R).
    static const char* force = getenv("SE_LSTM_CLUSTER");      // tuning aid: R of the cluster kernel to force
    const int fr = force ? atoi(force) : 0;
    if (fr == 4) return se::launch_lstm_cluster<4>(pre, pre_ld, w0, w1, h0, h1, nl, N, T, s);
    if (fr == 2) return se::launch_lstm_cluster<2>(pre, pre_ld, w0, w1, h0, h1, nl, N, T, s);
    if (fr == 1) return se::launch_lstm_cluster<1>(pre, pre_ld, w0, w1, h0, h1, nl, N, T, s);
    return dispatch_rows<128, 64>(pre, pre_ld, w0, w1, h0, h1, nl, N, T, s);
  }
  if (H == 64) return dispatch_rows<64, 64>(pre, pre_ld, w0, w1, h0, h1, nl, N, T, s);
  return se::fail(SE_ERR_UNSUPPORTED, "se_lstm_rec_fwd: hidden size %d not supported (64 or 128)", H);
}

extern "C" int se_lstm_rec_fwd(const float* pre_gates, int64_t pre_ld, const float* w_hh, int64_t N, int64_t T, int H,
                               float* h_out, void* stream) {
  SE_REQUIRE(pre_gates && w_hh && h_out, "se_lstm_rec_fwd: null pointer");
  SE_REQUIRE(N >= 0 && T >= 0 && H > 0 && pre_ld >= 4 * (int64_t)H, "se_lstm_rec_fwd: bad dims");
  if (N == 0 || T == 0) return SE_OK;
  return lstm_rec_any(pre_gates, pre_ld, w_hh, w_hh, h_out, h_out, 1, N, T, H, (cudaStream_t)stream);
}

extern "C" int se_lstm_rec_pair_fwd(const float* pre_gates, int64_t pre_ld, const float* w_hh0, const float* w_hh1,
                                    int64_t N, int64_t T, int H, float* h_out0, float* h_out1, void* stream) {
  SE_REQUIRE(pre_gates && w_hh0 && w_hh1 && h_out0 && h_out1, "se_lstm_rec_pair_fwd: null pointer");
  SE_REQUIRE(N >= 0 && T >= 0 && H > 0 && pre_ld >= 8 * (int64_t)H, "se_lstm_rec_pair_fwd: bad dims");
  if (N == 0 || T == 0) return SE_OK;
  return lstm_rec_any(pre_gates, pre_ld, w_hh0, w_hh1, h_out0, h_out1, 2, N, T, H, (cudaStream_t)stream);
}

// ---- LSTM-block output rows -> CL2 decoder input ---------------------------------------------------------------
// The reference reshapes the [B, T, C*F] output of the LSTM block back to [B, C, F, T] (_2008_00264_dccrn.py:164-165,
// column j = part * (C/2) * F + c * F + f for channel part * C/2 + c, complex halves first).  This kernel does that
// permutation while converting to the engine's CL2 format [B][2][F][T][Cp] (bf16 hi | lo), one (b, t) row per block
// staged through shared memory, 16-byte stores of 8 channels -- instead of a transposing copy + se_cl2_from_nchw.
namespace se {
__global__ void __launch_bounds__(128)
cl2_from_lstm_rows_kernel(const float* __restrict__ rows, int64_t ld, int C, int F, int T, int Cp,
                          __nv_bfloat16* __restrict__ out) {
  extern __shared__ float row_s[];                 // [C * F]
  const int64_t r = blockIdx.x;                    // row = b * T + t
  const int64_t b = r / T;
  const int t = (int)(r - b * T);
  const int K = C * F, half = (C / 2) * F;
  const float* src = rows + r * ld;
  for (int i = threadIdx.x; i < K; i += blockDim.x) row_s[i] = __ldg(src + i);
  __syncthreads();
  const int groups = Cp / 8;
  const int64_t plane = (int64_t)F * T * Cp;
  for (int i = threadIdx.x; i < F * groups; i += blockDim.x) {
    const int f = i / groups, c0 = (i - f * groups) * 8;
    __align__(16) __nv_bfloat16 hi[8], lo[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int ch = c0 + j;
      float v = 0.f;
      if (ch < C) {
        const int part = ch >= C / 2, c = ch - part * (C / 2);
        v = row_s[part * half + c * F + f];
      }
      hi[j] = __float2bfloat16_rn(v);
      lo[j] = __float2bfloat16_rn(v - __bfloat162float(hi[j]));
    }
    __nv_bfloat16* dst = out + (b * 2) * plane + ((int64_t)f * T + t) * Cp + c0;
    *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(hi);
    *reinterpret_cast<uint4*>(dst + plane) = *reinterpret_cast<const uint4*>(lo);
  }
}

// compile-time F: a thread owns 8 channels of one (b, t) row for every f -- its 8 F input floats are contiguous
// (128 bytes for F = 4), the F stores of a warp are 512-byte runs; no shared memory
template <int F>
__global__ void __launch_bounds__(256)
cl2_from_lstm_rows_f_kernel(const float* __restrict__ rows, int64_t ld, int64_t nrows, int C, int T, int Cp,
                            __nv_bfloat16* __restrict__ out) {
  const int groups = Cp / 8;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nrows * groups) return;
  const int64_t r = i / groups;
  const int c0 = (int)(i - r * groups) * 8;
  const int64_t b = r / T;
  const int t = (int)(r - b * T);
  const int64_t plane = (int64_t)F * T * Cp;
  __nv_bfloat16* dst = out + (b * 2) * plane + (int64_t)t * Cp + c0;
  float v[8 * F];
  if (c0 < C) {                                      // C / 2 is a multiple of 8: the group lies in one complex half
    const int part = c0 >= C / 2, c = c0 - part * (C / 2);
    const float4* src = reinterpret_cast<const float4*>(rows + r * ld + (int64_t)part * (C / 2) * F + (int64_t)c * F);
#pragma unroll
    for (int j = 0; j < 2 * F; ++j) {
      const float4 q = __ldg(src + j);
      v[4 * j] = q.x; v[4 * j + 1] = q.y; v[4 * j + 2] = q.z; v[4 * j + 3] = q.w;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 8 * F; ++j) v[j] = 0.f;
  }
#pragma unroll
  for (int f = 0; f < F; ++f) {
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float a0 = v[(2 * j) * F + f], a1 = v[(2 * j + 1) * F + f];
      const __nv_bfloat162 h2 = __floats2bfloat162_rn(a0, a1);
      const float2 hf = __bfloat1622float2(h2);
      const __nv_bfloat162 l2 = __floats2bfloat162_rn(a0 - hf.x, a1 - hf.y);
      hi[j] = *reinterpret_cast<const uint32_t*>(&h2);
      lo[j] = *reinterpret_cast<const uint32_t*>(&l2);
    }
    *reinterpret_cast<uint4*>(dst + (int64_t)f * T * Cp) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<uint4*>(dst + plane + (int64_t)f * T * Cp) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  }
}
}  // namespace se

extern "C" int se_cl2_from_lstm_rows(const float* rows, int64_t ld, int64_t B, int C, int F, int T, void* cl2, int Cp,
                                     void* stream) {
  SE_REQUIRE(rows && cl2, "se_cl2_from_lstm_rows: null pointer");
  SE_REQUIRE(B >= 0 && C > 0 && C % 2 == 0 && F > 0 && T > 0 && Cp >= C && Cp % 64 == 0 && ld >= (int64_t)C * F,
             "se_cl2_from_lstm_rows: bad dims");
  if ((int64_t)C * F > 8192) return se::fail(SE_ERR_UNSUPPORTED, "se_cl2_from_lstm_rows: needs C * F <= 8192");
  if (B * T > 0x7fffffffll) return se::fail(SE_ERR_UNSUPPORTED, "se_cl2_from_lstm_rows: too many rows");
  if (B == 0) return SE_OK;
  // fast path: 8-channel groups inside one complex half, 16-byte aligned 8*F-float runs
  if ((F == 2 || F == 4) && (C / 2) % 8 == 0 && ld % 4 == 0 && (reinterpret_cast<uintptr_t>(rows) & 15) == 0) {
    const int64_t total = B * T * (Cp / 8);
    const unsigned blocks = (unsigned)((total + 255) / 256);
    if (F == 4)
      se::cl2_from_lstm_rows_f_kernel<4><<<blocks, 256, 0, (cudaStream_t)stream>>>(rows, ld, B * T, C, T, Cp,
                                                                                   reinterpret_cast<__nv_bfloat16*>(cl2));
    else
      se::cl2_from_lstm_rows_f_kernel<2><<<blocks, 256, 0, (cudaStream_t)stream>>>(rows, ld, B * T, C, T, Cp,
                                                                                   reinterpret_cast<__nv_bfloat16*>(cl2));
    return se::check_launch("se_cl2_from_lstm_rows");
  }
  se::cl2_from_lstm_rows_kernel<<<(unsigned)(B * T), 128, (size_t)C * F * sizeof(float), (cudaStream_t)stream>>>(
      rows, ld, C, F, T, Cp, reinterpret_cast<__nv_bfloat16*>(cl2));
  return se::check_launch("se_cl2_from_lstm_rows");
}

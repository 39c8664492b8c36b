// ConvSTFT / ConviSTFT as shared-memory FFTs: the HBM-bound engine of the STFT front / back end.
//
// The reference projects every frame onto a dense windowed real-DFT basis with F.conv1d
// (/root/reference/models/conv_stft.py:56) and inverts with F.conv_transpose1d against pinv(basis) * window
// (:17-20, :101).  When the module's `weight` buffer IS that basis (the host checks it against the closed form
// before choosing this engine) the projection is  window -> 512-point real FFT  and the inverse is
//   y_t[m] = window[m]/N * ( z_t[m] - S_parity(m) / (N/2 + n_parity(m)) ),   z_t[m] = 2 Re sum'_f X_t[f] e^{+2 pi i f m / N}
// (sum' counts DC and Nyquist once without their imaginary parts; S_even / S_odd are the sums of z_t over the even /
// odd samples below `win`, n_even / n_odd how many there are) -- the rank-2 term is what pinv of the TRUNCATED DFT
// adds to a plain inverse FFT:  K^T K = (N/2) I + e e^T + o o^T  for the first `win` rows of rfft(eye(N)).
// ~25x fewer FLOPs than the dense projection, all in exact fp32 on the CUDA cores, so both directions move
// their algorithmic bytes once: wave -> spectrum (4(L + C2 T) B/utt) and spectrum -> wave.
//
// Work decomposition (both kernels): one CTA per (utterance, tile of <= 32 frames / hop blocks), 8 warps.
// A warp transforms TWO frames at a time as one 512-point complex FFT (frame A real part, frame B imaginary
// part): 16 points per lane, radix-16 in registers, one transpose through shared memory, radix-16 in registers
// again and a final radix-2 across lane pairs (shfl.xor 16).  Spectra / frames of the whole tile are staged in
// shared memory so that every global access is a contiguous run along the frame axis.
#include <type_traits>
#include <utility>
#include "common.cuh"
#include "umma.cuh"

namespace se {
namespace {

constexpr int NFFT = 512;       // points of the complex FFT (= fft_size)
constexpr int TILE = 32;        // frames (forward) / hop blocks (inverse) per CTA
constexpr int THREADS = 512;     // inverse kernel: 16 warps = the 16 frame pairs of a tile, one round
constexpr int WARPS = THREADS / 32;
constexpr int SROWS = 16 * 33;   // rows of the inverse kernel's tile: 2F = 514 spectrum rows, >= 16 * TPITCH so that a
                                 // pair's own column doubles as its transpose scratch
constexpr int ZPITCH = 529;     // float2 per pair row: >= 16*33 transpose scratch, odd -> conflict-free columns
constexpr int TPITCH = 33;      // transpose scratch row pitch (float2)
constexpr int MAXJ = 8;         // ceil(win / hop) covered by the inverse kernel
constexpr int FTHREADS = 512;   // forward kernel: 16 warps = the 16 frame pairs of a tile
constexpr int FWARPS = FTHREADS / 32;

template <int... I, class F>
__device__ __forceinline__ void static_for_impl(std::integer_sequence<int, I...>, F&& f) {
  (f(std::integral_constant<int, I>{}), ...);
}
template <int N, class F>
__device__ __forceinline__ void static_for(F&& f) { static_for_impl(std::make_integer_sequence<int, N>{}, f); }

__host__ __device__ constexpr int brev4(int k) { return ((k & 1) << 3) | ((k & 2) << 1) | ((k & 4) >> 1) | ((k & 8) >> 3); }

// cos / sin of 2 pi j / 32, j = 0..15
__host__ __device__ constexpr float cos32(int j) {
  constexpr float t[16] = {1.0f, 0.98078528040323043f, 0.92387953251128674f, 0.83146961230254524f,
                           0.70710678118654752f, 0.55557023301960218f, 0.38268343236508978f, 0.19509032201612825f,
                           0.0f, -0.19509032201612825f, -0.38268343236508978f, -0.55557023301960218f,
                           -0.70710678118654752f, -0.83146961230254524f, -0.92387953251128674f, -0.98078528040323043f};
  return t[j];
}
__host__ __device__ constexpr float sin32(int j) {
  constexpr float t[16] = {0.0f, 0.19509032201612825f, 0.38268343236508978f, 0.55557023301960218f,
                           0.70710678118654752f, 0.83146961230254524f, 0.92387953251128674f, 0.98078528040323043f,
                           1.0f, 0.98078528040323043f, 0.92387953251128674f, 0.83146961230254524f,
                           0.70710678118654752f, 0.55557023301960218f, 0.38268343236508978f, 0.19509032201612825f};
  return t[j];
}

// complex add / subtract as ONE packed add.f32x2 each (sm_100 FADD2): the butterflies are half of the FFT's FP instructions
#ifdef SE_FFT_SCALAR_ADDS
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
#else
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return __fadd2_rn(a, make_float2(-b.x, -b.y)); }
#endif
// a * (c - i s) forward, a * (c + i s) inverse
template <bool INV>
__device__ __forceinline__ float2 cmul_w(float2 a, float c, float s) {
  if (INV) return make_float2(fmaf(a.x, c, -a.y * s), fmaf(a.y, c, a.x * s));
  return make_float2(fmaf(a.x, c, a.y * s), fmaf(a.y, c, -a.x * s));
}
// a * W16^J (forward: exp(-2 pi i J / 16), inverse: conjugate), J in [0, 8)
template <int J, bool INV>
__device__ __forceinline__ float2 twid16(float2 a) {
  if constexpr (J == 0) return a;
  else if constexpr (J == 4) return INV ? make_float2(-a.y, a.x) : make_float2(a.y, -a.x);
  else return cmul_w<INV>(a, cos32(2 * J), sin32(2 * J));
}

// 16-point DFT in registers, radix-2 decimation in frequency; X[k] ends up in v[brev4(k)]
template <bool INV>
__device__ __forceinline__ void fft16(float2 (&v)[16]) {
  static_for<8>([&](auto j_) {
    constexpr int j = decltype(j_)::value;
    float2 u = v[j], w = v[j + 8];
    v[j] = cadd(u, w);
    v[j + 8] = twid16<j, INV>(csub(u, w));
  });
  static_for<8>([&](auto i_) {
    constexpr int j = decltype(i_)::value & 3, i = (decltype(i_)::value >> 2) * 8 + j;
    float2 u = v[i], w = v[i + 4];
    v[i] = cadd(u, w);
    v[i + 4] = twid16<2 * j, INV>(csub(u, w));
  });
  static_for<8>([&](auto i_) {
    constexpr int j = decltype(i_)::value & 1, i = (decltype(i_)::value >> 1) * 4 + j;
    float2 u = v[i], w = v[i + 2];
    v[i] = cadd(u, w);
    v[i + 2] = twid16<4 * j, INV>(csub(u, w));
  });
  static_for<8>([&](auto i_) {
    constexpr int i = decltype(i_)::value * 2;
    float2 u = v[i], w = v[i + 1];
    v[i] = cadd(u, w);
    v[i + 1] = csub(u, w);
  });
}

// 512-point complex DFT of one warp: on entry lane n1 holds z[n1 + 32 n2] in v[n2]; on exit lane (h = lane >> 4,
// k2 = lane & 15) holds Z[16 k + 256 h + k2] in v[k].  tw[k2] = exp(-+2 pi i lane k2 / 512) (cos, sin magnitudes).
// `row` is this pair's scratch row (>= 16 * TPITCH float2); the caller owns it exclusively.
template <bool INV>
__device__ __forceinline__ void warp_fft512(float2 (&v)[16], const float2 (&tw)[16], float2* row, int lane) {
  fft16<INV>(v);
  static_for<16>([&](auto k2_) {
    constexpr int k2 = decltype(k2_)::value;
    float2 y = v[brev4(k2)];
    if constexpr (k2 != 0) y = cmul_w<INV>(y, tw[k2].x, tw[k2].y);
    row[k2 * TPITCH + lane] = y;
  });
  __syncwarp();
  const int h = lane >> 4, k2l = lane & 15;
  static_for<16>([&](auto m_) {
    constexpr int m = decltype(m_)::value; v[m] = row[k2l * TPITCH + 2 * m + h]; });
  __syncwarp();
  fft16<INV>(v);
  float2 r[16];
  const float sgn = h ? -1.0f : 1.0f;
  static_for<16>([&](auto k_) {
    constexpr int k = decltype(k_)::value;
    float2 mine = v[brev4(k)];
    // odd half: times W32^k before the last butterfly
    if constexpr (k != 0) {
      float c = h ? cos32(k) : 1.0f, s = h ? sin32(k) : 0.0f;
      mine = cmul_w<INV>(mine, c, s);
    }
    float2 other = make_float2(__shfl_xor_sync(0xffffffffu, mine.x, 16), __shfl_xor_sync(0xffffffffu, mine.y, 16));
    r[k] = make_float2(fmaf(sgn, mine.x, other.x), fmaf(sgn, mine.y, other.y));   // E + O' (h = 0) / E - O' (h = 1)
  });
  static_for<16>([&](auto k_) {
    constexpr int k = decltype(k_)::value; v[k] = r[k]; });
}

// Same transform with the stage twiddles read from a shared-memory table tw_s[k2 * 32 + lane] (built once per CTA)
// and the results handed to `emit(k, Z[16 k + 256 h + k2])` instead of staying in registers: 64 registers per thread
// instead of 128, i.e. 32 resident warps per SM instead of 16.
template <bool INV, class Emit>
__device__ __forceinline__ void warp_fft512_emit(float2 (&v)[16], const float2* __restrict__ tw_s, float2* row, int lane,
                                                 Emit&& emit) {
  fft16<INV>(v);
  static_for<16>([&](auto k2_) {
    constexpr int k2 = decltype(k2_)::value;
    float2 y = v[brev4(k2)];
    if constexpr (k2 != 0) {
      const float2 t = tw_s[k2 * 32 + lane];
      y = cmul_w<INV>(y, t.x, t.y);
    }
    row[k2 * TPITCH + lane] = y;
  });
  __syncwarp();
  const int h = lane >> 4, k2l = lane & 15;
  static_for<16>([&](auto m_) {
    constexpr int m = decltype(m_)::value;
    v[m] = row[k2l * TPITCH + 2 * m + h];
  });
  __syncwarp();
  fft16<INV>(v);
  const float sgn = h ? -1.0f : 1.0f;
  static_for<16>([&](auto k_) {
    constexpr int k = decltype(k_)::value;
    float2 mine = v[brev4(k)];
    if constexpr (k != 0) {                  // odd half: times W32^k before the last butterfly
      float c = h ? cos32(k) : 1.0f, sn = h ? sin32(k) : 0.0f;
      mine = cmul_w<INV>(mine, c, sn);
    }
    float2 other = make_float2(__shfl_xor_sync(0xffffffffu, mine.x, 16), __shfl_xor_sync(0xffffffffu, mine.y, 16));
    emit(k_, make_float2(fmaf(sgn, mine.x, other.x), fmaf(sgn, mine.y, other.y)));   // E + O' (h = 0) / E - O' (h = 1)
  });
}

// warp_fft512 with the stage twiddles read from a shared-memory table tw_s[k2 * 32 + lane] and the transpose scratch
// addressed with a stride: scratch element i lives at col[i * hp].  The inverse kernel passes the pair's own column of
// the spectrum tile (already consumed into registers), so it needs no scratch buffer of its own.
template <bool INV>
__device__ __forceinline__ void warp_fft512_col(float2 (&v)[16], const float2* __restrict__ tw_s, float2* col, int hp,
                                                int lane) {
  fft16<INV>(v);
  static_for<16>([&](auto k2_) {
    constexpr int k2 = decltype(k2_)::value;
    float2 y = v[brev4(k2)];
    if constexpr (k2 != 0) {
      const float2 t = tw_s[k2 * 32 + lane];
      y = cmul_w<INV>(y, t.x, t.y);
    }
    col[(k2 * TPITCH + lane) * hp] = y;
  });
  __syncwarp();
  const int h = lane >> 4, k2l = lane & 15;
  static_for<16>([&](auto m_) {
    constexpr int m = decltype(m_)::value;
    v[m] = col[(k2l * TPITCH + 2 * m + h) * hp];
  });
  __syncwarp();
  fft16<INV>(v);
  float2 r[16];
  const float sgn = h ? -1.0f : 1.0f;
  static_for<16>([&](auto k_) {
    constexpr int k = decltype(k_)::value;
    float2 mine = v[brev4(k)];
    if constexpr (k != 0) {                  // odd half: times W32^k before the last butterfly
      float c = h ? cos32(k) : 1.0f, sn = h ? sin32(k) : 0.0f;
      mine = cmul_w<INV>(mine, c, sn);
    }
    float2 other = make_float2(__shfl_xor_sync(0xffffffffu, mine.x, 16), __shfl_xor_sync(0xffffffffu, mine.y, 16));
    r[k] = make_float2(fmaf(sgn, mine.x, other.x), fmaf(sgn, mine.y, other.y));   // E + O' (h = 0) / E - O' (h = 1)
  });
  static_for<16>([&](auto k_) {
    constexpr int k = decltype(k_)::value; v[k] = r[k]; });
}

__device__ __forceinline__ void lane_twiddles(float2 (&tw)[16], int lane) {
  static_for<16>([&](auto k2_) {
    constexpr int k2 = decltype(k2_)::value;
    float s, c;
    sincospif((float)(lane * k2) * (1.0f / 256.0f), &s, &c);     // angle 2 pi lane k2 / 512
    tw[k2] = make_float2(c, s);
  });
}

// 4-byte asynchronous global -> shared copy (LDGSTS): the staging loops issue all of a tile's copies back to back
// and wait once, so the memory system sees the whole tile in flight.  `valid == false` writes 0.
__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc, bool valid = true) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  int sz = valid ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(d), "l"(gsrc), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async8(float* smem_dst, const float* gsrc, bool valid = true) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  int sz = valid ? 8 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;\n" ::"r"(d), "l"(gsrc), "r"(sz) : "memory");
}
// warp index as a warp-uniform value: lets the compiler keep the per-pair loops (and the shuffles inside) convergent
__device__ __forceinline__ int uniform_warp_idx() { return __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }

// ------------------------------------------------------------------------------------------------ forward
// out[b, f, t] = Re X_t[f], out[b, F + f, t] = Im X_t[f],  X_t = rfft_N( window * reflect_pad(x)[t*hop : t*hop + win] )
// Persistent CTAs walk the (utterance, tile of 32 frames) list; the next tile's samples stream into the other half of
// a double buffer while this tile is transformed.
// TIn = float, or int16_t for PCM16 clips (the wav files of audio_dataset.py:12-21 are PCM_S 16;
// torchaudio.load scales by 1 / 32768): the int16 -> fp32 conversion happens in this stage load, so a PCM clip costs
// 2 bytes per sample of HBM read instead of a separate conversion pass (2 read + 4 written + 4 read).
template <typename TIn>
__global__ void __launch_bounds__(FTHREADS, 2)
stft_fft_kernel(const TIn* __restrict__ x, int64_t L, const float* __restrict__ window, int win, int hop, int pad,
                float* __restrict__ out, int T, int tiles, int total, int xs_cap) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* zbuf = reinterpret_cast<float2*>(smem_raw);                   // [TILE/2][ZPITCH]
  float2* tw_s = zbuf + (TILE / 2) * ZPITCH;                            // [16][32] stage twiddles exp(-2 pi i lane k2 / 512)
  float* wl_s = reinterpret_cast<float*>(tw_s + 16 * 32);               // [NFFT] window / 2 (0 beyond win)
  float* xs = wl_s + NFFT;                                              // [2][xs_cap] staged samples
  const int tid = threadIdx.x, lane = tid & 31, warp = uniform_warp_idx();
  constexpr int F = NFFT / 2 + 1;

  // stage a tile's samples once (each is used by ceil(win / hop) frames).  Interior tiles whose first sample is
  // 16-byte aligned are ONE TMA bulk copy (cp.async.bulk -> mbarrier) issued by a single thread; tiles that touch the
  // clip ends mirror sample by sample (F.pad(mode='reflect'), conv_stft.py:55) with 4-byte cp.async.  The NFFT - win
  // floats behind the last frame are zeroed so that the frame loads below need no bounds predicate (they are
  // multiplied by a zero window tap).
  __shared__ __align__(8) uint64_t bars[2];
  if (tid == 0) {
    ptx::mbar_init(&bars[0], 1);
    ptx::mbar_init(&bars[1], 1);
    ptx::fence_mbar_init();
  }
  __syncthreads();
  auto stage = [&](int idx, float* dst, uint64_t* bar) {
    const int b = idx / tiles, t0 = (idx - b * tiles) * TILE;
    const int nfr = min(TILE, T - t0);
    const int nsamp = (nfr - 1) * hop + win;
    const int64_t s0 = (int64_t)t0 * hop - pad;
    const TIn* xb = x + (int64_t)b * L;
    if constexpr (!std::is_same<TIn, float>::value) {
      // PCM16: plain loads with the reflect mirror, converted on the way into shared memory
      for (int i = tid; i < nsamp; i += FTHREADS) {
        int64_t sidx = s0 + i;
        if (sidx < 0) sidx = -sidx;
        else if (sidx >= L) sidx = 2 * (L - 1) - sidx;
        dst[i] = (float)__ldg(xb + sidx) * (1.0f / 32768.0f);
      }
      if (tid == 0) ptx::mbar_arrive(bar);
    } else if (s0 >= 0 && s0 + nsamp <= L && ((reinterpret_cast<uintptr_t>(xb + s0) & 15) == 0)) {
      const float* src = reinterpret_cast<const float*>(xb) + s0;
      const int n4 = nsamp & ~3;
      if (tid == 0) {
        ptx::fence_proxy_async_smem();         // the buffer was read through the generic proxy one tile ago
        ptx::mbar_arrive_expect_tx(bar, (uint32_t)n4 * 4u);
        ptx::tma_bulk_load_1d(dst, src, (uint32_t)n4 * 4u, bar);
      }
      if (tid < nsamp - n4) cp_async4(dst + n4 + tid, src + n4 + tid);
    } else {
      for (int i = tid; i < nsamp; i += FTHREADS) {
        int64_t sidx = s0 + i;
        if (sidx < 0) sidx = -sidx;
        else if (sidx >= L) sidx = 2 * (L - 1) - sidx;
        cp_async4(dst + i, reinterpret_cast<const float*>(xb) + sidx);
      }
      if (tid == 0) ptx::mbar_arrive(bar);
    }
    for (int i = tid; i < NFFT - win; i += FTHREADS) dst[nsamp + i] = 0.0f;
    cp_async_commit();
  };

  int idx = blockIdx.x;
  if (idx < total) stage(idx, xs, &bars[0]);
  for (int i = tid; i < 16 * 32; i += FTHREADS) {
    float sn, cs;
    sincospif((float)((i >> 5) * (i & 31)) * (1.0f / 256.0f), &sn, &cs);      // angle 2 pi lane k2 / 512
    tw_s[i] = make_float2(cs, sn);
  }
  // window taps times the 1/2 of the pair separation
  for (int i = tid; i < NFFT; i += FTHREADS) wl_s[i] = i < win ? 0.5f * __ldg(window + i) : 0.0f;
  const bool t_even = (T & 1) == 0 && (reinterpret_cast<uintptr_t>(out) & 7) == 0;

  for (int buf = 0, it = 0; idx < total; idx += gridDim.x, buf ^= 1, ++it) {
    const int b = idx / tiles, t0 = (idx - b * tiles) * TILE;
    const int nfr = min(TILE, T - t0);
    const float* xc = xs + buf * xs_cap;
    ptx::mbar_wait(&bars[buf], (uint32_t)(it >> 1) & 1u);
    cp_async_wait_all();
    __syncthreads();                         // this tile's samples landed; everyone left the previous write-out
    if (idx + (int)gridDim.x < total) stage(idx + gridDim.x, xs + (buf ^ 1) * xs_cap, &bars[buf ^ 1]);

    const int npairs = (nfr + 1) >> 1;
    if (warp < npairs) {                     // one pair of frames per warp
      const int p = warp;
      const float* fa = xc + (2 * p) * hop + lane;
      const bool has_b = 2 * p + 1 < nfr;
      const float* fb = has_b ? fa + hop : fa;
      float2 v[16];
      static_for<16>([&](auto n2_) {
        constexpr int n2 = decltype(n2_)::value;
        const float w = wl_s[lane + 32 * n2];
        v[n2] = make_float2(fa[32 * n2] * w, has_b ? fb[32 * n2] * w : 0.0f);
      });
      float2* row = zbuf + p * ZPITCH;
      const int base = 256 * (lane >> 4) + (lane & 15);
      warp_fft512_emit<false>(v, tw_s, row, lane, [&](auto k_, float2 z) {
        constexpr int k = decltype(k_)::value;
        row[16 * k + base] = z;
      });
    }
    __syncthreads();

    // separate the two real spectra of each pair and write runs along t:  X_A[f] = (Z[f] + conj Z[N-f]) / 2,
    // X_B[f] = (Z[f] - conj Z[N-f]) / 2i  (the 1/2 rode in on the window).  Half a warp covers one row f
    // (16 pairs = up to 32 frames = 128 contiguous bytes); t0 is a multiple of 32, so with an even T every pair is an
    // aligned float2.
    const int p = lane & 15, fo = lane >> 4;
    if (p < npairs) {
      const bool has_b = 2 * p + 1 < nfr;
      const float2* zr = zbuf + p * ZPITCH;
      float* pre = out + (int64_t)b * (2 * F) * T + (t0 + 2 * p) + (int64_t)(2 * warp + fo) * T;
      const int64_t im_off = (int64_t)F * T, step = (int64_t)(2 * FWARPS) * T;
      if (t_even && has_b) {
#pragma unroll 4
        for (int f = 2 * warp + fo; f < F; f += 2 * FWARPS, pre += step) {
          float2 z1 = zr[f], z2 = zr[(NFFT - f) & (NFFT - 1)];
          *reinterpret_cast<float2*>(pre) = make_float2(z1.x + z2.x, z1.y + z2.y);
          *reinterpret_cast<float2*>(pre + im_off) = make_float2(z1.y - z2.y, z2.x - z1.x);
        }
      } else {
#pragma unroll 2
        for (int f = 2 * warp + fo; f < F; f += 2 * FWARPS, pre += step) {
          float2 z1 = zr[f], z2 = zr[(NFFT - f) & (NFFT - 1)];
          pre[0] = z1.x + z2.x;
          pre[im_off] = z1.y - z2.y;
          if (has_b) { pre[1] = z1.y + z2.y; pre[im_off + 1] = z2.x - z1.x; }
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ inverse
// out[b, n - crop] = scale[n] * sum_t y_t[n - t*hop]   (conv_stft.py:101-114), y_t as in the file header.
// One tile owns `tq` = 32 - halo hop blocks of the output and transforms the 32 frames (16 pairs: two full rounds
// of the 8 warps) that reach them; the halo frames are recomputed by the neighbouring tile: deterministic, no
// atomics, no second pass.  halo = J - 1 rounded up to even, so that a tile's first frame is even and a pair of
// frames is one aligned 8-byte copy when T is even.
constexpr int IPAIRS = TILE / 2;            // frame pairs per tile
constexpr int IPITCH = 2 * (IPAIRS | 1);    // floats per spectrum row in shared memory; IPITCH / 2 odd

// interior tiles: every frame exists, every tap is below `win`, every sample is inside the crop
// fp32 sample -> stored sample.  int16_t = PCM_S 16 as the reference's savers write it (mix_audio.py:144-146,
// trainer.py:255-262: torchaudio.save(encoding="PCM_S", bits_per_sample=16), which in torchaudio >= 2.9 is FFmpeg's
// flt -> s16 conversion av_clip_int16(lrintf(x * 2^15))), with the models' final clamp(-1, 1)
// (_2008_00264_dccrn.py:184) fused in front of it.
template <typename TOut> __device__ __forceinline__ TOut to_sample(float v);
template <> __device__ __forceinline__ float to_sample<float>(float v) { return v; }
template <> __device__ __forceinline__ int16_t to_sample<int16_t>(float v) {
  v = fminf(fmaxf(v, -1.0f), 1.0f);
  return (int16_t)min(__float2int_rn(v * 32768.0f), 32767);
}

template <int JJ, typename TOut>
__device__ __forceinline__ void ola_interior(const float* sbuf, int total_s, int hop, int halo, unsigned hop_magic,
                                             const float* __restrict__ scale, TOut* __restrict__ dst, int tid) {
  const int jstep = hop * IPITCH - 1;
  for (int i = tid; i < total_s; i += THREADS) {
    int ql = __umulhi((unsigned)i, hop_magic);           // i / hop
    const float* col = sbuf + (i - ql * hop) * IPITCH + ql + halo;
    float acc = col[0];
#pragma unroll
    for (int j = 1; j < JJ; ++j) acc += col[j * jstep];
    dst[i] = to_sample<TOut>(acc * __ldg(scale + i));
  }
}

template <typename TOut>
__global__ void __launch_bounds__(THREADS, 2)
istft_fft_kernel(const float* __restrict__ spec, int T, const float* __restrict__ window, int win, int hop, int J,
                 float c_even, float c_odd, const float* __restrict__ scale, int64_t crop, int64_t out_len,
                 TOut* __restrict__ out, int Q, int tq, int tiles, int total, unsigned hop_magic) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int F = NFFT / 2 + 1, C2 = 2 * F;
  float* sbuf = reinterpret_cast<float*>(smem_raw);                         // [SROWS][IPITCH]: spectra, then frames
  float2* tw_s = reinterpret_cast<float2*>(sbuf + (size_t)SROWS * IPITCH);  // [16][32] stage twiddles (cos, sin magnitudes)
  float* wl_s = reinterpret_cast<float*>(tw_s + 16 * 32);                   // [NFFT] window / N (0 beyond win)
  const int tid = threadIdx.x, lane = tid & 31, warp = uniform_warp_idx();

  // 64 registers per thread (32 resident warps per SM instead of 16): the stage twiddles and the window taps live in
  // shared-memory tables, and a pair's transpose scratch is its own column of the tile
  for (int i = tid; i < 16 * 32; i += THREADS) {
    float sn, cs;
    sincospif((float)((i >> 5) * (i & 31)) * (1.0f / 256.0f), &sn, &cs);      // angle 2 pi lane k2 / 512
    tw_s[i] = make_float2(cs, sn);
  }
  for (int i = tid; i < NFFT; i += THREADS) wl_s[i] = i < win ? __ldg(window + i) * (1.0f / NFFT) : 0.0f;
  const int h = lane >> 4, m2 = lane & 15;
  const float cpar = (m2 & 1) ? c_odd : c_even;
  constexpr int hp = IPITCH >> 1;                              // row pitch in float2 (odd)
  float2* s2 = reinterpret_cast<float2*>(sbuf);
  const int halo = (J - 1 + 1) & ~1;
  const bool pair_copies = (T & 1) == 0 && (reinterpret_cast<uintptr_t>(spec) & 7) == 0;

  for (int idx = blockIdx.x; idx < total; idx += gridDim.x) {
    const int b = idx / tiles, q0 = (idx - b * tiles) * tq;
    const int nq = min(tq, Q - q0);
    const int tfirst = q0 - halo, nfr = nq + halo;
    const int npairs = (nfr + 1) >> 1;
    const float* sb = spec + (int64_t)b * C2 * T;

    // stage spec[b, :, tfirst : tfirst + 2 npairs] (zero outside [0, T)): half a warp per spectrum row (16 pairs =
    // 128 contiguous bytes), lanes along t
    if (pair_copies) {
      const int pc = lane & 15, c0 = 2 * warp + (lane >> 4);
      const int t = tfirst + 2 * pc;
      const bool ok = t >= 0 && t < T;                         // T and t even: both frames of the pair exist or neither
      if (pc < npairs) {
        const float* src = sb + (int64_t)c0 * T + (ok ? t : 0);
        float* dst = sbuf + c0 * IPITCH + 2 * pc;
#pragma unroll 4
        for (int c = c0; c < C2; c += 2 * WARPS, src += (int64_t)(2 * WARPS) * T, dst += 2 * WARPS * IPITCH)
          cp_async8(dst, src, ok);
      }
    } else {
      const int t = tfirst + lane;
      const bool ok = lane < nfr && t >= 0 && t < T;
      if (lane < 2 * npairs) {
        const float* src = sb + (int64_t)warp * T + (ok ? t : 0);
        float* dst = sbuf + warp * IPITCH + lane;
#pragma unroll 4
        for (int c = warp; c < C2; c += WARPS, src += (int64_t)WARPS * T, dst += WARPS * IPITCH) cp_async4(dst, src, ok);
      }
    }
    cp_async_commit();
    cp_async_wait_all();
    __syncthreads();

    for (int p = warp; p < npairs; p += WARPS) {
      // skip pairs whose frames both lie outside [0, T): their columns already hold zeros
      int tA = tfirst + 2 * p;
      if (tA + 1 < 0 || tA >= T) continue;
      // G[f] = X_A[f] + i X_B[f] on the Hermitian-extended axis (times 2; DC and Nyquist real, counted once)
      float2 v[16];
      const float2* colp = s2 + p;
      static_for<16>([&](auto f2_) {
        constexpr int f2 = decltype(f2_)::value;
        int g = f2 < 8 ? lane + 32 * f2 : NFFT - (lane + 32 * f2);     // f2 == 8, lane 0 -> Nyquist (g = 256)
        float2 re = colp[g * hp];                                      // (Re X_A, Re X_B)
        float2 im = colp[(F + g) * hp];                                // (Im X_A, Im X_B)
        constexpr float sg = f2 < 8 ? 1.0f : -1.0f;
        float2 gg = make_float2(re.x - sg * im.y, fmaf(sg, im.x, re.y));
        if constexpr (f2 == 0 || f2 == 8) {
          if (lane == 0) gg = make_float2(2.0f * re.x, 2.0f * re.y);
        }
        v[f2] = gg;
      });
      __syncwarp();                                      // every lane holds its share of the column: it is scratch now
      warp_fft512_col<true>(v, tw_s, s2 + p, hp, lane);
      // rank-2 term of the truncated-DFT pseudo-inverse: subtract c_parity * (sum of this parity's samples below win)
      float2 acc = make_float2(0.f, 0.f);
      static_for<16>([&](auto k_) {
        constexpr int k = decltype(k_)::value;
        if (16 * k + 256 * h + m2 < win) acc = cadd(acc, v[k]);
      });
#pragma unroll
      for (int off = 2; off <= 16; off <<= 1) {
        acc.x += __shfl_xor_sync(0xffffffffu, acc.x, off);
        acc.y += __shfl_xor_sync(0xffffffffu, acc.y, off);
      }
      acc.x *= cpar; acc.y *= cpar;
      // rows >= win get zeros (their window tap is 0); the overlap-add never reads them
      __syncwarp();                                      // the scratch reads of the other lanes are done
      float2* ycol = s2 + p + (256 * h + m2) * hp;
      const float* wlp = wl_s + 256 * h + m2;
      static_for<16>([&](auto k_) {
        constexpr int k = decltype(k_)::value;
        const float w = wlp[16 * k];
        ycol[16 * k * hp] = make_float2((v[k].x - acc.x) * w, (v[k].y - acc.y) * w);
      });
    }
    __syncthreads();

    // overlap-add in a fixed order, normalise, crop:  sample n = q*hop + r sums frame q - j at offset j*hop + r.
    // Frames outside [0, T) are skipped, not added as zeros: the partner of a real frame in a pair carries that
    // frame's rounding noise, which the 1 / (window^2 + 1e-8) normaliser would amplify at un-centred clip edges.
    TOut* ob = out + (int64_t)b * out_len;
    const int total_s = nq * hop;
    const int64_t n0 = (int64_t)q0 * hop;
    const bool interior = q0 - (J - 1) >= 0 && q0 + nq <= T && J * hop == win && n0 >= crop &&
                          n0 + total_s <= crop + out_len && scale != nullptr;
    if (interior && J == 4) ola_interior<4, TOut>(sbuf, total_s, hop, halo, hop_magic, scale + n0, ob + (n0 - crop), tid);
    else if (interior && J == 2) ola_interior<2, TOut>(sbuf, total_s, hop, halo, hop_magic, scale + n0, ob + (n0 - crop), tid);
    else {
      for (int i = tid; i < total_s; i += THREADS) {
        int ql = __umulhi((unsigned)i, hop_magic), r = i - ql * hop;
        float acc = 0.0f;
        for (int j = 0; j < J; ++j) {
          int m = j * hop + r, t = q0 + ql - j;
          if (m < win && t >= 0 && t < T) acc += sbuf[m * IPITCH + (ql - j + halo)];
        }
        int64_t n = n0 + i, pos = n - crop;
        if (pos >= 0 && pos < out_len) ob[pos] = to_sample<TOut>(scale ? acc * __ldg(scale + n) : acc);
      }
    }
    __syncthreads();                         // the next tile restages sbuf
  }
}

bool fft_framing_ok(int win, int hop, int fft) {
  return fft == NFFT && win >= 2 && win <= fft && hop >= 2 && hop <= 256 && (win + hop - 1) / hop <= MAXJ;
}

}  // namespace
}  // namespace se

extern "C" int se_stft_fft_supported(int win, int hop, int fft) { return se::fft_framing_ok(win, hop, fft) ? 1 : 0; }

namespace se {
namespace {

template <typename TIn>
int stft_fft_launch(const char* nm, const TIn* x, int64_t B, int64_t L, const float* window, int win, int hop, int fft,
                    int pad, float* out, int64_t T, void* stream) {
  SE_REQUIRE(x && window && out, "%s: null pointer", nm);
  SE_REQUIRE(B >= 0 && L > 0 && T >= 0 && pad >= 0, "%s: bad dims", nm);
  if (!fft_framing_ok(win, hop, fft))
    return fail(SE_ERR_UNSUPPORTED, "%s: win %d / hop %d / fft %d outside the FFT engine (fft 512, hop <= 256, "
                                    "ceil(win/hop) <= %d)", nm, win, hop, fft, MAXJ);
  SE_REQUIRE(pad < L && (T - 1) * hop + win <= L + 2 * (int64_t)pad, "%s: frames exceed the padded clip", nm);
  if (B == 0 || T == 0) return SE_OK;
  int64_t tiles = (T + TILE - 1) / TILE;
  SE_REQUIRE(T < (1 << 30) && tiles * B < (1ll << 31), "%s: batch x frame count too large", nm);
  int xs_cap = ((TILE - 1) * hop + NFFT + 3) & ~3;
  size_t smem = (size_t)(TILE / 2) * ZPITCH * sizeof(float2) + 16 * 32 * sizeof(float2) + NFFT * sizeof(float) +
                2 * (size_t)xs_cap * sizeof(float);
  static se::PerDevice<bool> attr_set_pd(false);
  bool& attr_set = attr_set_pd.get();
  if (!attr_set) {
    SE_CUDA(cudaFuncSetAttribute(stft_fft_kernel<TIn>, cudaFuncAttributeMaxDynamicSharedMemorySize, 144 * 1024));
    attr_set = true;
  }
  int total = (int)(tiles * B);
  int grid = total < 2 * num_sms() ? total : 2 * num_sms();
  stft_fft_kernel<TIn><<<grid, FTHREADS, smem, (cudaStream_t)stream>>>(x, L, window, win, hop, pad, out, (int)T,
                                                                      (int)tiles, total, xs_cap);
  return check_launch(nm);
}

template <typename TOut>
int istft_fft_launch(const char* nm, const float* spec, int64_t B, int64_t T, const float* window, int win, int hop,
                     int fft, const float* scale, int64_t crop_front, int64_t out_len, TOut* out, void* stream) {
  SE_REQUIRE(spec && window && out, "%s: null pointer", nm);
  SE_REQUIRE(B >= 0 && T >= 0 && crop_front >= 0 && out_len >= 0, "%s: bad dims", nm);
  if (!fft_framing_ok(win, hop, fft))
    return fail(SE_ERR_UNSUPPORTED, "%s: win %d / hop %d / fft %d outside the FFT engine (fft 512, hop <= 256, "
                                    "ceil(win/hop) <= %d)", nm, win, hop, fft, MAXJ);
  if (B == 0 || T == 0 || out_len == 0) return SE_OK;
  SE_REQUIRE(crop_front + out_len <= (T - 1) * (int64_t)hop + win, "%s: crop window outside the overlap-add length", nm);
  int J = (win + hop - 1) / hop;
  SE_REQUIRE(T < (1 << 30), "%s: frame count too large", nm);
  int Q = (int)T + J - 1;
  int halo = (J - 1 + 1) & ~1;
  int tq = TILE - halo;                                       // hop blocks of output per tile (32 frames with the halo)
  int64_t tiles = (Q + tq - 1) / tq;
  SE_REQUIRE(tiles * B < (1ll << 31), "%s: batch x frame count too large", nm);
  size_t smem = (size_t)SROWS * IPITCH * sizeof(float) + 16 * 32 * sizeof(float2) + NFFT * sizeof(float);
  static se::PerDevice<bool> attr_set_pd(false);
  bool& attr_set = attr_set_pd.get();
  if (!attr_set) {
    SE_CUDA(cudaFuncSetAttribute(istft_fft_kernel<TOut>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    attr_set = true;
  }
  // closed form of pinv for the first `win` rows of the real DFT:
  // (K^T K)^-1 = (2/N) (I - e e^T / (N/2 + n_e) - o o^T / (N/2 + n_o))
  float c_even = 1.0f / (float)(fft / 2 + (win + 1) / 2), c_odd = 1.0f / (float)(fft / 2 + win / 2);
  int total = (int)(tiles * B);
  int grid = total < 2 * num_sms() ? total : 2 * num_sms();
  // i / hop for 0 <= i < 32 * 256 as __umulhi(i, magic): exact for these small dividends
  unsigned hop_magic = (unsigned)((0x100000000ull + (unsigned)hop - 1) / (unsigned)hop);
  istft_fft_kernel<TOut><<<grid, THREADS, smem, (cudaStream_t)stream>>>(spec, (int)T, window, win, hop, J, c_even, c_odd,
                                                                       scale, crop_front, out_len, out, Q, tq, (int)tiles,
                                                                       total, hop_magic);
  return check_launch(nm);
}

}  // namespace
}  // namespace se

extern "C" int se_stft_fft_fwd(const float* x, int64_t B, int64_t L, const float* window, int win, int hop, int fft,
                               int pad, float* out, int64_t T, void* stream) {
  return se::stft_fft_launch<float>("se_stft_fft_fwd", x, B, L, window, win, hop, fft, pad, out, T, stream);
}
extern "C" int se_stft_fft_fwd_pcm16(const int16_t* x, int64_t B, int64_t L, const float* window, int win, int hop, int fft,
                                     int pad, float* out, int64_t T, void* stream) {
  return se::stft_fft_launch<int16_t>("se_stft_fft_fwd_pcm16", x, B, L, window, win, hop, fft, pad, out, T, stream);
}
extern "C" int se_istft_fft_fwd(const float* spec, int64_t B, int64_t T, const float* window, int win, int hop, int fft,
                                const float* scale, int64_t crop_front, int64_t out_len, float* out, void* stream) {
  return se::istft_fft_launch<float>("se_istft_fft_fwd", spec, B, T, window, win, hop, fft, scale, crop_front, out_len, out,
                                     stream);
}
extern "C" int se_istft_fft_fwd_pcm16(const float* spec, int64_t B, int64_t T, const float* window, int win, int hop, int fft,
                                      const float* scale, int64_t crop_front, int64_t out_len, int16_t* out, void* stream) {
  return se::istft_fft_launch<int16_t>("se_istft_fft_fwd_pcm16", spec, B, T, window, win, hop, fft, scale, crop_front,
                                       out_len, out, stream);
}

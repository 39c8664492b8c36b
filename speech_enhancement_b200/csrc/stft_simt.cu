// ConvSTFT / ConviSTFT, exact-fp32 CUDA-core engine (the tcgen05 engine lives in stft_umma.cu).
// Reference semantics: /root/reference/models/conv_stft.py:48-66 (forward), :89-116 (inverse).
#include "simt_gemm.cuh"
#include "engines.cuh"

namespace se {

// ---- forward: out[b, c, t] = sum_k W[c, k] * xpad[b, t*hop + k] ---------------------------
struct StftProb {
  int M, K; int64_t N;          // M = C2, K = win, N = B*T
  const float* x; float* out;
  int64_t L, T; int hop, pad;   // pad = 0 when center == False
  struct Col { const float* xb; int64_t base; int64_t out_off; };
  __device__ Col col(int64_t n) const {
    int64_t b = n / T, t = n - b * T;
    return {x + b * L, t * hop - pad, b * (int64_t)M * T + t};
  }
  __device__ float b(int k, const Col& c) const {
    int64_t i = c.base + k;
    if (i < 0) i = -i;                         // F.pad(mode='reflect'), conv_stft.py:55
    else if (i >= L) i = 2 * (L - 1) - i;
    return __ldg(c.xb + i);
  }
  __device__ void store(int m, int64_t n, float acc) const {
    int64_t b = n / T, t = n - b * T;
    out[b * (int64_t)M * T + (int64_t)m * T + t] = acc;
  }
};

// ---- inverse, output-stationary form ---------------------------------------------------------
// y[b, q*hop + r] = sum_{j<J} sum_c Winv[c, j*hop + r] * spec[b, c, q - j]      (conv_stft.py:101)
// GEMM: M = r (hop), K = (j, c), N = (b, q) with q in [0, T + J - 1).  The epilogue applies
// 1 / (coff + 1e-8) with coff[n] = sum_j window[j*hop + r]^2 over the frames that exist
// (conv_stft.py:104-106, computed analytically instead of through the eye(win) conv) and the
// centre crop (:109-114).
struct IstftProb {
  int M, K; int64_t N;          // M = hop, K = J*C2, N = B*Q
  const float* spec; const float* phase; const float* window; float* out;
  int C2, F, J, hop, win; int64_t T, Q, crop, out_len;
  struct Col { int64_t b, q; };
  __device__ Col col(int64_t n) const { int64_t b = n / Q; return {b, n - b * Q}; }
  __device__ float b(int k, const Col& c) const {
    int j = k / C2, ch = k - j * C2;
    int64_t t = c.q - j;
    if (t < 0 || t >= T) return 0.f;
    if (phase == nullptr) return __ldg(spec + (c.b * C2 + ch) * T + t);
    // polar input (conv_stft.py:96-100): mags, phase are [B, F, T]
    int f = ch < F ? ch : ch - F;
    float mg = __ldg(spec + (c.b * F + f) * T + t), ph = __ldg(phase + (c.b * F + f) * T + t);
    return ch < F ? mg * cosf(ph) : mg * sinf(ph);
  }
  __device__ void store(int r, int64_t n, float acc) const {
    int64_t b = n / Q, q = n - b * Q;
    int64_t pos = q * hop + r - crop;
    if (pos < 0 || pos >= out_len) return;
    float coff = 0.f;
    for (int j = 0; j < J; ++j) {
      int64_t t = q - j; int k = j * hop + r;
      if (t >= 0 && t < T && k < win) { float w = __ldg(window + k); coff = fmaf(w, w, coff); }
    }
    out[b * out_len + pos] = acc / (coff + 1e-8f);
  }
};

// A_packed[r, j*C2 + c] = Winv[c, j*hop + r]  (0 where j*hop + r >= win)
__global__ void istft_pack_kernel(const float* __restrict__ Winv, int C2, int win, int hop, int J,
                                  float* __restrict__ A) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t K = (int64_t)J * C2;
  if (i >= (int64_t)hop * K) return;
  int r = (int)(i / K); int k = (int)(i - (int64_t)r * K);
  int j = k / C2, c = k - j * C2;
  int col = j * hop + r;
  A[i] = col < win ? Winv[(int64_t)c * win + col] : 0.f;
}

__global__ void magphase_kernel(const float* __restrict__ spec, int64_t B, int F, int64_t T,
                                float* __restrict__ mags, float* __restrict__ phase) {
  int64_t plane = (int64_t)F * T, total = B * plane;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    int64_t b = i / plane, r = i - b * plane;
    float re = __ldg(spec + b * 2 * plane + r), im = __ldg(spec + b * 2 * plane + plane + r);
    mags[i] = sqrtf(re * re + im * im);            // conv_stft.py:62 (no epsilon)
    phase[i] = atan2f(im, re);                     // conv_stft.py:63
  }
}

__global__ void polar_kernel(const float* __restrict__ mags, const float* __restrict__ phase,
                             int64_t B, int F, int64_t T, float* __restrict__ spec) {
  int64_t plane = (int64_t)F * T, total = B * plane;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    int64_t b = i / plane, r = i - b * plane;
    float m = __ldg(mags + i), s, c;
    sincosf(__ldg(phase + i), &s, &c);
    spec[b * 2 * plane + r] = m * c;               // conv_stft.py:98
    spec[b * 2 * plane + plane + r] = m * s;       // conv_stft.py:99
  }
}

int stft_fwd_simt(const float* x, int64_t B, int64_t L, const float* W, int C2, int win, int hop,
                  int pad, float* out, int64_t T, cudaStream_t s) {
  StftProb p{C2, win, B * T, x, out, L, T, hop, pad};
  return launch_simt_gemm(W, p, s, "se_stft_fwd(simt)");
}

int istft_fwd_simt(const float* spec, const float* phase, const float* A_packed, const float* window,
                   int64_t B, int C2, int64_t T, int win, int hop, int64_t crop, int64_t out_len,
                   float* out, cudaStream_t s) {
  int J = (win + hop - 1) / hop;
  int64_t Q = T + J - 1;
  IstftProb p{hop, J * C2, B * Q, spec, phase, window, out, C2, C2 / 2, J, hop, win, T, Q, crop, out_len};
  return launch_simt_gemm(A_packed, p, s, "se_istft_fwd(simt)");
}

}  // namespace se

extern "C" int se_istft_pack_basis(const float* Winv, int C2, int win, int hop, float* A_packed,
                                   void* stream) {
  SE_REQUIRE(Winv && A_packed, "se_istft_pack_basis: null pointer");
  SE_REQUIRE(C2 > 0 && win > 0 && hop > 0, "se_istft_pack_basis: bad dims");
  int J = (win + hop - 1) / hop;
  int64_t total = (int64_t)hop * J * C2;
  se::istft_pack_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      Winv, C2, win, hop, J, A_packed);
  return se::check_launch("se_istft_pack_basis");
}

extern "C" int se_magphase_fwd(const float* spec, int64_t B, int F, int64_t T, float* mags,
                               float* phase, void* stream) {
  SE_REQUIRE(spec && mags && phase, "se_magphase_fwd: null pointer");
  SE_REQUIRE(B >= 0 && F > 0 && T >= 0, "se_magphase_fwd: bad dims");
  if (B * F * T == 0) return SE_OK;
  int64_t total = B * F * T;
  unsigned grid = se::grid_1d(total);
  se::magphase_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(spec, B, F, T, mags, phase);
  return se::check_launch("se_magphase_fwd");
}

extern "C" int se_polar_to_rect(const float* mags, const float* phase, int64_t B, int F, int64_t T,
                                float* spec, void* stream) {
  SE_REQUIRE(spec && mags && phase, "se_polar_to_rect: null pointer");
  SE_REQUIRE(B >= 0 && F > 0 && T >= 0, "se_polar_to_rect: bad dims");
  if (B * F * T == 0) return SE_OK;
  int64_t total = B * F * T;
  unsigned grid = se::grid_1d(total);
  se::polar_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(mags, phase, B, F, T, spec);
  return se::check_launch("se_polar_to_rect");
}

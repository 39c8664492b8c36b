// ComplexConv2d / ComplexConvTranspose2d, exact-fp32 CUDA-core engine + the weight packer.
// Reference semantics: /root/reference/models/modules/complex_nn.py:52-91 -- four real convs
// on the channel halves, out = [rr - ii, ri + ir]; here ONE real implicit GEMM against the
// packed operand [[Wr,-Wi],[Wi,Wr]].
#include "simt_gemm.cuh"
#include "engines.cuh"

namespace se {

// packed[m, (ci2, dh, dw)], m in [0, 2Co), ci2 in [0, 2Ci)
__global__ void cconv_pack_kernel(const float* __restrict__ Wr, const float* __restrict__ Wi,
                                  const float* __restrict__ br, const float* __restrict__ bi,
                                  int Co, int Ci, int taps, int transposed,
                                  float* __restrict__ packed, float* __restrict__ bias_packed) {
  int64_t K = (int64_t)2 * Ci * taps, total = (int64_t)2 * Co * K;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < total) {
    int m = (int)(i / K); int64_t k = i - (int64_t)m * K;
    int ci2 = (int)(k / taps), tap = (int)(k - (int64_t)ci2 * taps);
    bool out_imag = m >= Co, in_imag = ci2 >= Ci;
    int co = out_imag ? m - Co : m, ci = in_imag ? ci2 - Ci : ci2;
    // real out: Wr*xr - Wi*xi ; imag out: Wi*xr + Wr*xi     (complex_nn.py:55-62)
    const float* src = (out_imag == in_imag) ? Wr : Wi;
    float sign = (!out_imag && in_imag) ? -1.f : 1.f;
    int64_t off = transposed ? ((int64_t)ci * Co + co) * taps + tap      // [Ci, Co, kh, kw]
                             : ((int64_t)co * Ci + ci) * taps + tap;     // [Co, Ci, kh, kw]
    packed[i] = sign * src[off];
  }
  if (bias_packed != nullptr && br != nullptr && bi != nullptr && i < 2 * Co) {
    // each sub-conv adds its own bias: real = (rr+br) - (ii+bi), imag = (ri+bi) + (ir+br)
    int m = (int)i;
    bias_packed[m] = m < Co ? br[m] - bi[m] : br[m - Co] + bi[m - Co];
  }
}

template <bool TRANSPOSED>
struct CConvProb {
  int M, K; int64_t N;          // M = Cout2, K = Cin2*kh*kw, N = B*Ho*Wo
  const float* x; const float* bias; float* y;
  ConvGeom g;
  struct Col { const float* xb; int h0, w0; };
  __device__ Col col(int64_t n) const {
    int64_t hw = (int64_t)g.Ho * g.Wo;
    int64_t b = n / hw; int r = (int)(n - b * hw);
    int ho = r / g.Wo, wo = r - ho * g.Wo;
    const float* xb = x + b * (int64_t)g.Cin2 * g.H * g.W;
    if (TRANSPOSED) return {xb, ho + g.ph, wo + g.pw};
    return {xb, ho * g.sh - g.ph, wo * g.sw - g.pw};
  }
  __device__ float b(int k, const Col& c) const {
    int taps = g.kh * g.kw;
    int ci = k / taps, tap = k - ci * taps;
    int dh = tap / g.kw, dw = tap - dh * g.kw;
    int hi, wi;
    if (TRANSPOSED) {
      // ho = hi*sh - ph + dh  =>  hi = (ho + ph - dh) / sh when divisible
      int hn = c.h0 - dh, wn = c.w0 - dw;
      if (hn < 0 || wn < 0) return 0.f;
      hi = hn / g.sh; wi = wn / g.sw;
      if (hi * g.sh != hn || wi * g.sw != wn) return 0.f;
    } else {
      hi = c.h0 + dh; wi = c.w0 + dw;
      if (hi < 0 || wi < 0) return 0.f;
    }
    if (hi >= g.H || wi >= g.W) return 0.f;
    return __ldg(c.xb + ((int64_t)ci * g.H + hi) * g.W + wi);
  }
  __device__ void store(int m, int64_t n, float acc) const {
    int64_t hw = (int64_t)g.Ho * g.Wo;
    int64_t b = n / hw, r = n - b * hw;
    if (bias != nullptr) acc += __ldg(bias + m);
    y[(b * g.Cout2 + m) * hw + r] = acc;
  }
};

int cconv_fwd_simt(bool transposed, const float* x, const float* packed, const float* bias,
                   int64_t B, const ConvGeom& g, float* y, cudaStream_t s) {
  int K = g.Cin2 * g.kh * g.kw;
  int64_t N = B * g.Ho * g.Wo;
  if (transposed) {
    CConvProb<true> p{g.Cout2, K, N, x, bias, y, g};
    return launch_simt_gemm(packed, p, s, "se_cconvT2d_fwd(simt)");
  }
  CConvProb<false> p{g.Cout2, K, N, x, bias, y, g};
  return launch_simt_gemm(packed, p, s, "se_cconv2d_fwd(simt)");
}

}  // namespace se

extern "C" int se_cconv_pack_weight(const float* Wr, const float* Wi, const float* br, const float* bi,
                                    int Co, int Ci, int kh, int kw, int transposed,
                                    float* packed, float* bias_packed, void* stream) {
  SE_REQUIRE(Wr && Wi && packed, "se_cconv_pack_weight: null pointer");
  SE_REQUIRE(Co > 0 && Ci > 0 && kh > 0 && kw > 0, "se_cconv_pack_weight: bad dims");
  SE_REQUIRE((br == nullptr) == (bi == nullptr), "se_cconv_pack_weight: need both biases or none");
  int64_t total = (int64_t)4 * Co * Ci * kh * kw;
  se::cconv_pack_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      Wr, Wi, br, bi, Co, Ci, kh * kw, transposed, packed, bias_packed);
  return se::check_launch("se_cconv_pack_weight");
}

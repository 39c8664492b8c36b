// extern "C" entry points of libse_b200.so that dispatch between engines (include/se_b200.h).
#include "common.cuh"
#include "engines.cuh"

namespace se { std::atomic<int64_t> g_launch_count{0}; }

extern "C" int64_t se_launch_count(void) { return se::g_launch_count.load(); }
extern "C" int se_version(void) { return 100; }   // 0.1.0
extern "C" const char* se_last_error(void) { return se::last_error_buf(); }

extern "C" int se_device_ok(void) {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return se::fail(SE_ERR_CUDA, "cudaGetDevice failed");
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess)
    return se::fail(SE_ERR_CUDA, "cudaDeviceGetAttribute failed");
  return major == 10 ? 1 : 0;
}

extern "C" int se_stft_fwd(const float* x, int64_t B, int64_t L, const float* W, int C2, int win,
                           int hop, int pad, int center, float* out, int64_t T, void* stream) {
  SE_REQUIRE(x && W && out, "se_stft_fwd: null pointer");
  SE_REQUIRE(B >= 0 && L > 0 && C2 > 0 && win > 0 && hop > 0 && pad >= 0, "se_stft_fwd: bad dims");
  int p = center ? pad : 0;
  SE_REQUIRE(p < L, "se_stft_fwd: reflect padding (%d) must be smaller than the signal (%lld)", p,
             (long long)L);
  SE_REQUIRE(L + 2 * p >= win, "se_stft_fwd: signal shorter than one window");
  SE_REQUIRE(T == (L + 2 * p - win) / hop + 1, "se_stft_fwd: T does not match L/win/hop");
  if (B == 0) return SE_OK;
  return se::stft_fwd_simt(x, B, L, W, C2, win, hop, p, out, T, (cudaStream_t)stream);
}

extern "C" int se_istft_fwd(const float* spec, const float* phase, const float* A_packed,
                            const float* window, int64_t B, int C2, int64_t T, int win, int hop,
                            int64_t crop_front, int64_t out_len, float* out, void* stream) {
  SE_REQUIRE(spec && A_packed && window && out, "se_istft_fwd: null pointer");
  SE_REQUIRE(B >= 0 && C2 > 0 && (C2 % 2) == 0 && T > 0 && win > 0 && hop > 0, "se_istft_fwd: bad dims");
  SE_REQUIRE(crop_front >= 0 && out_len >= 0 && crop_front + out_len <= (T - 1) * hop + win,
             "se_istft_fwd: crop window outside the overlap-add length");
  if (B == 0 || out_len == 0) return SE_OK;
  return se::istft_fwd_simt(spec, phase, A_packed, window, B, C2, T, win, hop, crop_front, out_len,
                            out, (cudaStream_t)stream);
}

static int cconv_dispatch(bool transposed, const float* x, const float* packed, const float* bias,
                          int64_t B, int Cin2, int H, int W, int Cout2, int kh, int kw, int sh, int sw,
                          int ph, int pw, float* y, int Ho, int Wo, int engine, void* ws,
                          int64_t ws_bytes, void* stream) {
  const char* nm = transposed ? "se_cconvT2d_fwd" : "se_cconv2d_fwd";
  SE_REQUIRE(x && packed && y, "%s: null pointer", nm);
  SE_REQUIRE(B >= 0 && Cin2 > 0 && Cout2 > 0 && (Cin2 % 2) == 0 && (Cout2 % 2) == 0,
             "%s: channel counts must be positive and even", nm);
  SE_REQUIRE(H > 0 && W > 0 && kh > 0 && kw > 0 && sh > 0 && sw > 0 && ph >= 0 && pw >= 0 &&
             Ho > 0 && Wo > 0, "%s: bad geometry", nm);
  if (!transposed)
    SE_REQUIRE(Ho == (H + 2 * ph - kh) / sh + 1 && Wo == (W + 2 * pw - kw) / sw + 1,
               "%s: output size does not match geometry", nm);
  else
    SE_REQUIRE(Ho >= (H - 1) * sh - 2 * ph + kh && Ho < (H - 1) * sh - 2 * ph + kh + sh &&
               Wo >= (W - 1) * sw - 2 * pw + kw && Wo < (W - 1) * sw - 2 * pw + kw + sw,
               "%s: output size does not match geometry (output_padding must be < stride)", nm);
  if (B == 0) return SE_OK;
  se::ConvGeom g{Cin2, H, W, Cout2, kh, kw, sh, sw, ph, pw, Ho, Wo};
  if (engine == SE_ENGINE_AUTO) {
    // degenerate layers are HBM-bound: direct fp32 kernels
    if (!transposed && Cin2 <= 8 && Cin2 * kh * kw <= 512 && Ho <= 65535 && B <= 65535)
      return se_cconv_small_cin(x, (int64_t)Cin2 * H * W, (int64_t)H * W, W, B, Cin2, H, W, 0, 0, packed, bias, 1.f,
                                Cout2, kh, kw, sh, sw, ph, pw, y, nullptr, 0, Ho, Wo, stream);
    if (Cout2 <= 4 && kh * kw * Cin2 <= 12800 && Ho <= 65535 && B <= 65535)
      return se::cconv_fwd_small_cout(transposed, x, packed, bias, B, g, y, (cudaStream_t)stream);
  }
  if (engine == SE_ENGINE_UMMA || (engine == SE_ENGINE_AUTO && se::cconv_umma_supported(transposed, B, g))) {
    return se::cconv_fwd_umma(transposed, x, packed, bias, B, g, y, ws, ws_bytes, (cudaStream_t)stream);
  }
  if (engine != SE_ENGINE_AUTO && engine != SE_ENGINE_SIMT)
    return se::fail(SE_ERR_BAD_ARG, "%s: unknown engine %d", nm, engine);
  return se::cconv_fwd_simt(transposed, x, packed, bias, B, g, y, (cudaStream_t)stream);
}

extern "C" int se_cconv2d_fwd(const float* x, const float* packed, const float* bias, int64_t B,
                              int Cin2, int H, int W, int Cout2, int kh, int kw, int sh, int sw, int ph,
                              int pw, float* y, int Ho, int Wo, int engine, void* ws, int64_t ws_bytes,
                              void* stream) {
  return cconv_dispatch(false, x, packed, bias, B, Cin2, H, W, Cout2, kh, kw, sh, sw, ph, pw, y, Ho, Wo,
                        engine, ws, ws_bytes, stream);
}

extern "C" int se_cconvT2d_fwd(const float* x, const float* packed, const float* bias, int64_t B,
                               int Cin2, int H, int W, int Cout2, int kh, int kw, int sh, int sw, int ph,
                               int pw, float* y, int Ho, int Wo, int engine, void* ws, int64_t ws_bytes,
                               void* stream) {
  return cconv_dispatch(true, x, packed, bias, B, Cin2, H, W, Cout2, kh, kw, sh, sw, ph, pw, y, Ho, Wo,
                        engine, ws, ws_bytes, stream);
}

extern "C" int64_t se_cconv_workspace_bytes(int64_t B, int Cin2, int H, int W, int Cout2, int kh, int kw,
                                            int sh, int sw, int transposed) {
  return se::cconv_umma_workspace_bytes(transposed != 0, B, Cin2, H, W, Cout2, kh, kw, sh, sw);
}

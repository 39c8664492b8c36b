// Fused complex-mask kernels: 4 strided input planes -> the [B, 2F, T] real|imag stack the
// iSTFT consumes, one pass (the reference issues ~24 pointwise launches + pad + cat for
// DCCRN-E).  file:line of each variant is in include/se_b200.h.
#include "common.cuh"

namespace se {

struct Plane { const float* p; int64_t bs, fs; };

__device__ __forceinline__ void mag_phase(float re, float im, float& mag, float& ph) {
  mag = sqrtf(re * re + im * im + 1e-8f);      // _2008_00264_dccrn.py:210, _1903_03107_dcunet.py:187
  ph = atan2f(im, re);
}

// (cos, sin) of atan2(im, re) without the transcendental round trip: the unit vector (re, im) / |(re, im)|.
// Exactly-zero, denormal-scale and huge inputs take the reference's own atan2 -> sincos route (signed zeros and
// the atan2(0, 0) = 0 convention included); the branch is warp-divergent only on such bins.
__device__ __forceinline__ void unit_dir(float re, float im, float& c, float& s) {
  const float d2 = re * re + im * im;
  if (d2 > 1e-30f && d2 < 1e30f) {
    const float r = rsqrtf(d2);
    c = re * r; s = im * r;
  } else {
    sincosf(atan2f(im, re), &s, &c);
  }
}

// The polar masks are evaluated in rectangular form:  mag * (cos, sin)(a + b) is the complex product of the unit
// vectors of a and b, so DCCRN-E needs no atan2 / sincos at all (the reference's eps sits only inside the two
// magnitudes, _2008_00264_dccrn.py:210, and is kept there) and DCUNet's bounded-tanh needs one atan2 + one sincos
// for its angle / magnitude quirk (_1903_03107_dcunet.py:177).  ~60 instead of ~250 instructions per bin: the
// kernel is HBM-bound instead of SFU/issue-bound.
template <int KIND>
__device__ __forceinline__ void mask_eval(float nr, float ni, float mr, float mi, float& re, float& im) {
  if (KIND == SE_MASK_DCCRN_R || KIND == SE_MASK_DCUNET_UNB) {
    re = nr * mr; im = ni * mi;
  } else if (KIND == SE_MASK_DCCRN_C) {
    re = nr * mr - ni * mi; im = nr * mi + ni * mr;
  } else if (KIND == SE_MASK_CARN) {
    re = mr * nr - mi * ni; im = mr * ni - mi * nr;       // _2104_05267_carn.py:166 sign kept
  } else if (KIND == SE_MASK_DCUNET_SIGMOID) {
    re = nr * (1.f / (1.f + expf(-mr))); im = ni * (1.f / (1.f + expf(-mi)));
  } else if (KIND == SE_MASK_FRCRN) {
    re = tanhf(mr) * nr; im = tanhf(mi) * ni;
  } else {
    const float n_mag = sqrtf(nr * nr + ni * ni + 1e-8f), m_mag = sqrtf(mr * mr + mi * mi + 1e-8f);
    float cn, sn, cm, sm;
    unit_dir(nr, ni, cn, sn);                                              // noisy phase
    if (KIND == SE_MASK_DCCRN_E) unit_dir(mr / m_mag, mi / m_mag, cm, sm); // dccrn.py:198-202 (scale invariant)
    else sincosf(atan2f(mi, mr) / m_mag, &sm, &cm);                        // dcunet.py:177: angle / magnitude
    const float g = n_mag * tanhf(m_mag);
    re = g * (cn * cm - sn * sm); im = g * (cn * sm + sn * cm);
  }
}

// one warp per (b, f) row: lanes walk t in 128-byte runs, four runs (16 loads) in flight per lane
template <int KIND>
__global__ void __launch_bounds__(256)
mask_kernel(Plane nr, Plane ni, Plane mr, Plane mi, int64_t rows, int F, int64_t T, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int64_t b = row / F; const int f = (int)(row - b * F);
  const float* pnr = nr.p + b * nr.bs + f * nr.fs;
  const float* pni = ni.p + b * ni.bs + f * ni.fs;
  float* ore = out + (b * 2 * F + f) * T;
  float* oim = ore + (int64_t)F * T;
  if (KIND == SE_MASK_FRCRN && f < 2) {
    for (int64_t t = lane; t < T; t += 32) { ore[t] = 0.f; oim[t] = 0.f; }
    return;
  }
  const int fm = KIND == SE_MASK_FRCRN ? f - 2 : f;
  const float* pmr = mr.p + b * mr.bs + fm * mr.fs;
  const float* pmi = mi.p + b * mi.bs + fm * mi.fs;
  for (int64_t t0 = lane; t0 < T; t0 += 128) {
    float a[4], c[4], d[4], e[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t t = t0 + 32 * j;
      const bool ok = t < T;
      a[j] = ok ? __ldg(pnr + t) : 0.f; c[j] = ok ? __ldg(pni + t) : 0.f;
      d[j] = ok ? __ldg(pmr + t) : 1.f; e[j] = ok ? __ldg(pmi + t) : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t t = t0 + 32 * j;
      float re, im;
      mask_eval<KIND>(a[j], c[j], d[j], e[j], re, im);
      if (t < T) { ore[t] = re; oim[t] = im; }
    }
  }
}

template <int KIND>
int launch_mask(Plane nr, Plane ni, Plane mr, Plane mi, int64_t B, int F, int64_t T, float* out,
                cudaStream_t s) {
  const int64_t rows = B * F, blocks = (rows + 7) / 8;
  if (blocks > 0x7fffffffll) return fail(SE_ERR_UNSUPPORTED, "se_mask_fwd: B*F too large");
  mask_kernel<KIND><<<(unsigned)blocks, 256, 0, s>>>(nr, ni, mr, mi, rows, F, T, out);
  return check_launch("se_mask_fwd");
}

// ---- backward: (d re, d im) -> gradients of the four input planes, same strided inputs --------
// Closed forms of what autograd derives from the reference's pointwise graph (file:line as in forward).
template <int KIND>
__device__ __forceinline__ void mask_grad(float nr, float ni, float mr, float mi, float gr, float gi,
                                          float& dnr, float& dni, float& dmr, float& dmi) {
  if (KIND == SE_MASK_DCCRN_R || KIND == SE_MASK_DCUNET_UNB) {
    dnr = gr * mr; dmr = gr * nr; dni = gi * mi; dmi = gi * ni;
  } else if (KIND == SE_MASK_DCCRN_C) {
    dnr = gr * mr + gi * mi; dni = -gr * mi + gi * mr; dmr = gr * nr + gi * ni; dmi = -gr * ni + gi * nr;
  } else if (KIND == SE_MASK_CARN) {
    dnr = gr * mr - gi * mi; dni = -gr * mi + gi * mr; dmr = gr * nr + gi * ni; dmi = -gr * ni - gi * nr;
  } else if (KIND == SE_MASK_DCUNET_SIGMOID) {
    const float sr = 1.f / (1.f + expf(-mr)), si = 1.f / (1.f + expf(-mi));
    dnr = gr * sr; dmr = gr * nr * sr * (1.f - sr); dni = gi * si; dmi = gi * ni * si * (1.f - si);
  } else if (KIND == SE_MASK_FRCRN) {
    const float tr = tanhf(mr), ti = tanhf(mi);
    dnr = gr * tr; dmr = gr * nr * (1.f - tr * tr); dni = gi * ti; dmi = gi * ni * (1.f - ti * ti);
  } else {
    float n_mag, n_ph, m_mag, m_at;
    mag_phase(nr, ni, n_mag, n_ph);
    mag_phase(mr, mi, m_mag, m_at);
    const float m_ph = KIND == SE_MASK_DCCRN_E ? atan2f(mi / m_mag, mr / m_mag) : m_at / m_mag;
    const float th = tanhf(m_mag), A = n_mag * th;
    float sn, cs;
    sincosf(n_ph + m_ph, &sn, &cs);
    const float dA = gr * cs + gi * sn, dth = A * (-gr * sn + gi * cs);
    const float dn_mag = dA * th, dm_mag = dA * n_mag * (1.f - th * th);
    const float n2 = nr * nr + ni * ni, m2 = mr * mr + mi * mi;
    dnr = dn_mag * nr / n_mag - dth * ni / n2;
    dni = dn_mag * ni / n_mag + dth * nr / n2;
    if (KIND == SE_MASK_DCCRN_E) {
      // atan2 of the normalised mask is scale invariant: d/dm = (-mi, mr) / (mr^2 + mi^2)
      dmr = dm_mag * mr / m_mag - dth * mi / m2;
      dmi = dm_mag * mi / m_mag + dth * mr / m2;
    } else {
      // phase = atan2(mi, mr) / |m|   (_1903_03107_dcunet.py:177)
      const float k = m_at / (m_mag * m_mag * m_mag);
      dmr = dm_mag * mr / m_mag + dth * (-mi / (m2 * m_mag) - k * mr);
      dmi = dm_mag * mi / m_mag + dth * (mr / (m2 * m_mag) - k * mi);
    }
  }
}

// grid: (ceil(T / blockDim), min(rows = B*F, 65535)), rows strided over grid.y; outputs contiguous: dn_* [B,F,T], dm_* [B,Fm,T]
template <int KIND>
__global__ void __launch_bounds__(128)
mask_bwd_kernel(Plane nr, Plane ni, Plane mr, Plane mi, int64_t rows, int F, int64_t T, const float* __restrict__ g,
                float* __restrict__ dnr, float* __restrict__ dni, float* __restrict__ dmr, float* __restrict__ dmi) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const int Fm = KIND == SE_MASK_FRCRN ? F - 2 : F;
  for (int64_t row = blockIdx.y; row < rows; row += gridDim.y) {
  const int64_t b = row / F; const int f = (int)(row - b * F);
  const int64_t on = (b * F + f) * T + t;
  if (KIND == SE_MASK_FRCRN && f < 2) { dnr[on] = 0.f; dni[on] = 0.f; continue; }
  const int fm = KIND == SE_MASK_FRCRN ? f - 2 : f;
  const float a = __ldg(nr.p + b * nr.bs + f * nr.fs + t), c = __ldg(ni.p + b * ni.bs + f * ni.fs + t);
  const float d = __ldg(mr.p + b * mr.bs + fm * mr.fs + t), e = __ldg(mi.p + b * mi.bs + fm * mi.fs + t);
  const float gr = __ldg(g + (b * 2 * F + f) * T + t), gi = __ldg(g + (b * 2 * F + F + f) * T + t);
  float o0, o1, o2, o3;
  mask_grad<KIND>(a, c, d, e, gr, gi, o0, o1, o2, o3);
  dnr[on] = o0; dni[on] = o1;
  const int64_t om = (b * Fm + fm) * T + t;
  dmr[om] = o2; dmi[om] = o3;
  }
}

template <int KIND>
int launch_mask_bwd(Plane nr, Plane ni, Plane mr, Plane mi, int64_t B, int F, int64_t T, const float* g, float* dnr,
                    float* dni, float* dmr, float* dmi, cudaStream_t s) {
  const int64_t rows = B * F;
  dim3 grid((unsigned)((T + 127) / 128), (unsigned)(rows < 65535 ? rows : 65535));    // rows beyond grid.y are looped
  mask_bwd_kernel<KIND><<<grid, 128, 0, s>>>(nr, ni, mr, mi, rows, F, T, g, dnr, dni, dmr, dmi);
  return check_launch("se_mask_bwd");
}

}  // namespace se

extern "C" int se_mask_fwd(int kind,
                           const float* nr, int64_t nr_bs, int64_t nr_fs,
                           const float* ni, int64_t ni_bs, int64_t ni_fs,
                           const float* mr, int64_t mr_bs, int64_t mr_fs,
                           const float* mi, int64_t mi_bs, int64_t mi_fs,
                           int64_t B, int F, int64_t T, float* out, void* stream) {
  SE_REQUIRE(nr && ni && mr && mi && out, "se_mask_fwd: null pointer");
  SE_REQUIRE(B >= 0 && F > 0 && T >= 0, "se_mask_fwd: bad dims");
  if (B == 0 || T == 0) return SE_OK;
  se::Plane a{nr, nr_bs, nr_fs}, c{ni, ni_bs, ni_fs}, d{mr, mr_bs, mr_fs}, e{mi, mi_bs, mi_fs};
  cudaStream_t s = (cudaStream_t)stream;
  switch (kind) {
    case SE_MASK_DCCRN_R:        return se::launch_mask<SE_MASK_DCCRN_R>(a, c, d, e, B, F, T, out, s);
    case SE_MASK_DCCRN_C:        return se::launch_mask<SE_MASK_DCCRN_C>(a, c, d, e, B, F, T, out, s);
    case SE_MASK_DCCRN_E:        return se::launch_mask<SE_MASK_DCCRN_E>(a, c, d, e, B, F, T, out, s);
    case SE_MASK_DCUNET_UNB:     return se::launch_mask<SE_MASK_DCUNET_UNB>(a, c, d, e, B, F, T, out, s);
    case SE_MASK_DCUNET_SIGMOID: return se::launch_mask<SE_MASK_DCUNET_SIGMOID>(a, c, d, e, B, F, T, out, s);
    case SE_MASK_DCUNET_TANH:    return se::launch_mask<SE_MASK_DCUNET_TANH>(a, c, d, e, B, F, T, out, s);
    case SE_MASK_CARN:           return se::launch_mask<SE_MASK_CARN>(a, c, d, e, B, F, T, out, s);
    case SE_MASK_FRCRN:
      SE_REQUIRE(F >= 3, "se_mask_fwd: FRCRN mask needs F >= 3");
      return se::launch_mask<SE_MASK_FRCRN>(a, c, d, e, B, F, T, out, s);
  }
  return se::fail(SE_ERR_BAD_ARG, "se_mask_fwd: unknown kind %d", kind);
}

extern "C" int se_mask_bwd(int kind,
                           const float* nr, int64_t nr_bs, int64_t nr_fs,
                           const float* ni, int64_t ni_bs, int64_t ni_fs,
                           const float* mr, int64_t mr_bs, int64_t mr_fs,
                           const float* mi, int64_t mi_bs, int64_t mi_fs,
                           int64_t B, int F, int64_t T, const float* grad_out, float* d_nr, float* d_ni,
                           float* d_mr, float* d_mi, void* stream) {
  SE_REQUIRE(nr && ni && mr && mi && grad_out && d_nr && d_ni && d_mr && d_mi, "se_mask_bwd: null pointer");
  SE_REQUIRE(B >= 0 && F > 0 && T >= 0, "se_mask_bwd: bad dims");
  if (B == 0 || T == 0) return SE_OK;
  se::Plane a{nr, nr_bs, nr_fs}, c{ni, ni_bs, ni_fs}, d{mr, mr_bs, mr_fs}, e{mi, mi_bs, mi_fs};
  cudaStream_t s = (cudaStream_t)stream;
#define SE_BWD(K) return se::launch_mask_bwd<K>(a, c, d, e, B, F, T, grad_out, d_nr, d_ni, d_mr, d_mi, s)
  switch (kind) {
    case SE_MASK_DCCRN_R:        SE_BWD(SE_MASK_DCCRN_R);
    case SE_MASK_DCCRN_C:        SE_BWD(SE_MASK_DCCRN_C);
    case SE_MASK_DCCRN_E:        SE_BWD(SE_MASK_DCCRN_E);
    case SE_MASK_DCUNET_UNB:     SE_BWD(SE_MASK_DCUNET_UNB);
    case SE_MASK_DCUNET_SIGMOID: SE_BWD(SE_MASK_DCUNET_SIGMOID);
    case SE_MASK_DCUNET_TANH:    SE_BWD(SE_MASK_DCUNET_TANH);
    case SE_MASK_CARN:           SE_BWD(SE_MASK_CARN);
    case SE_MASK_FRCRN:
      SE_REQUIRE(F >= 3, "se_mask_bwd: FRCRN mask needs F >= 3");
      SE_BWD(SE_MASK_FRCRN);
  }
#undef SE_BWD
  return se::fail(SE_ERR_BAD_ARG, "se_mask_bwd: unknown kind %d", kind);
}

// ---- PCM_S 16 <-> fp32 (the wire format of audio_dataset.py:12-21 / mix_audio.py:144-146) ----------
// For the framings the FFT engine does not cover: the fused forms live in stft_fft.cu (se_stft_fft_fwd_pcm16 /
// se_istft_fft_fwd_pcm16).  from_f32 applies the models' clamp(-1, 1) and FFmpeg's flt -> s16 rule
// av_clip_int16(lrintf(x * 2^15)) (what torchaudio.save(encoding="PCM_S", bits_per_sample=16) runs); to_f32 is
// torchaudio.load's x / 32768.
namespace se {
__global__ void __launch_bounds__(256) pcm16_to_f32_kernel(const int16_t* __restrict__ x, float* __restrict__ y, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = (float)__ldg(x + i) * (1.0f / 32768.0f);
}
__global__ void __launch_bounds__(256) pcm16_from_f32_kernel(const float* __restrict__ x, int16_t* __restrict__ y, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = fminf(fmaxf(__ldg(x + i), -1.0f), 1.0f);
    y[i] = (int16_t)min(__float2int_rn(v * 32768.0f), 32767);
  }
}
}  // namespace se

extern "C" int se_pcm16_to_f32(const int16_t* x, int64_t n, float* y, void* stream) {
  SE_REQUIRE(n >= 0 && (n == 0 || (x && y)), "se_pcm16_to_f32: bad arguments");
  if (n == 0) return SE_OK;
  se::pcm16_to_f32_kernel<<<se::grid_1d(n), 256, 0, (cudaStream_t)stream>>>(x, y, n);
  return se::check_launch("se_pcm16_to_f32");
}
extern "C" int se_pcm16_from_f32(const float* x, int64_t n, int16_t* y, void* stream) {
  SE_REQUIRE(n >= 0 && (n == 0 || (x && y)), "se_pcm16_from_f32: bad arguments");
  if (n == 0) return SE_OK;
  se::pcm16_from_f32_kernel<<<se::grid_1d(n), 256, 0, (cudaStream_t)stream>>>(x, y, n);
  return se::check_launch("se_pcm16_from_f32");
}

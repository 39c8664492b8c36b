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
    float n_mag, n_ph, m_mag, m_ph;
    mag_phase(nr, ni, n_mag, n_ph);
    mag_phase(mr, mi, m_mag, m_ph);
    if (KIND == SE_MASK_DCCRN_E) m_ph = atan2f(mi / m_mag, mr / m_mag);   // dccrn.py:198-202
    else                          m_ph = m_ph / m_mag;                     // dcunet.py:177
    float g = n_mag * tanhf(m_mag), s, c;
    sincosf(n_ph + m_ph, &s, &c);
    re = g * c; im = g * s;
  }
}

// grid: (ceil(T / (blockDim*4)), rows = B*F); each thread handles 4 t's strided by blockDim
template <int KIND>
__global__ void __launch_bounds__(128)
mask_kernel(Plane nr, Plane ni, Plane mr, Plane mi, int F, int64_t T, float* __restrict__ out) {
  const int64_t row = blockIdx.y;
  const int64_t b = row / F; const int f = (int)(row - b * F);
  const float* pnr = nr.p + b * nr.bs + f * nr.fs;
  const float* pni = ni.p + b * ni.bs + f * ni.fs;
  float* ore = out + (b * 2 * F + f) * T;
  float* oim = ore + (int64_t)F * T;
  const int64_t t0 = (int64_t)blockIdx.x * (blockDim.x * 4) + threadIdx.x;
  if (KIND == SE_MASK_FRCRN && f < 2) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int64_t t = t0 + j * blockDim.x;
      if (t < T) { ore[t] = 0.f; oim[t] = 0.f; }
    }
    return;
  }
  const int fm = KIND == SE_MASK_FRCRN ? f - 2 : f;
  const float* pmr = mr.p + b * mr.bs + fm * mr.fs;
  const float* pmi = mi.p + b * mi.bs + fm * mi.fs;
  float a[4], c[4], d[4], e[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    int64_t t = t0 + j * blockDim.x;
    bool ok = t < T;
    a[j] = ok ? __ldg(pnr + t) : 0.f; c[j] = ok ? __ldg(pni + t) : 0.f;
    d[j] = ok ? __ldg(pmr + t) : 1.f; e[j] = ok ? __ldg(pmi + t) : 0.f;
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    int64_t t = t0 + j * blockDim.x;
    float re, im;
    mask_eval<KIND>(a[j], c[j], d[j], e[j], re, im);
    if (t < T) { ore[t] = re; oim[t] = im; }
  }
}

template <int KIND>
int launch_mask(Plane nr, Plane ni, Plane mr, Plane mi, int64_t B, int F, int64_t T, float* out,
                cudaStream_t s) {
  int64_t rows = B * F;
  // rows go on grid.y (max 65535); larger problems are split along the batch below
  if (rows <= 65535) {
    dim3 grid((unsigned)((T + 511) / 512), (unsigned)rows);
    mask_kernel<KIND><<<grid, 128, 0, s>>>(nr, ni, mr, mi, F, T, out);
    return check_launch("se_mask_fwd");
  }
  // split along the batch so each launch has <= 65535 rows
  int64_t bper = 65535 / F;
  if (bper < 1) return fail(SE_ERR_UNSUPPORTED, "se_mask_fwd: F too large");
  for (int64_t b0 = 0; b0 < B; b0 += bper) {
    int64_t nb = B - b0 < bper ? B - b0 : bper;
    Plane a = {nr.p + b0 * nr.bs, nr.bs, nr.fs}, c = {ni.p + b0 * ni.bs, ni.bs, ni.fs};
    Plane d = {mr.p + b0 * mr.bs, mr.bs, mr.fs}, e = {mi.p + b0 * mi.bs, mi.bs, mi.fs};
    dim3 grid((unsigned)((T + 511) / 512), (unsigned)(nb * F));
    mask_kernel<KIND><<<grid, 128, 0, s>>>(a, c, d, e, F, T, out + b0 * 2 * F * T);
    int rc = check_launch("se_mask_fwd");
    if (rc != SE_OK) return rc;
  }
  return SE_OK;
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

// FP32 issue / pipe rates on sm_100a: scalar FFMA vs packed FFMA2 (fma.rn.f32x2), with and without LDS in the mix.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ffma2_rate scripts/ffma2_rate.cu && /tmp/ffma2_rate
// Prints FMA lanes per clock per SM for: 8 independent FFMA chains, 8 independent FFMA2 chains (16 FMAs),
// at 8 / 16 / 32 warps per SM.  If FFMA2 doubles the FMA rate per issue slot, the LSTM matvec and the FFT butterflies
// (both issue-bound) gain from packing; if the pipe is the limit it only frees issue slots.
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(float* out, int iters, float a0, float b0) {
  float2 a = make_float2(a0, a0 * 1.0001f), b = make_float2(b0, b0 * 0.9999f);
  float2 acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = make_float2(i * 0.1f + threadIdx.x, i * 0.2f);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) {           // 16 scalar FFMA
        acc[i].x = fmaf(a.x, acc[i].x, b.x);
        acc[i].y = fmaf(a.y, acc[i].y, b.y);
      } else {                   // 8 FFMA2
        acc[i] = __ffma2_rn(a, acc[i], b);
      }
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += acc[i].x + acc[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  int dev = 0, sms = 0, khz = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  float* out;
  cudaMalloc(&out, sizeof(float) * sms * 1024 * 4);
  const int iters = 20000;
  for (int warps : {4, 8, 16, 32}) {
    for (int mode = 0; mode < 2; ++mode) {
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      auto launch = [&]() {
        if (mode == 0) k<0><<<sms, warps * 32>>>(out, iters, 0.999f, 0.001f);
        else k<1><<<sms, warps * 32>>>(out, iters, 0.999f, 0.001f);
      };
      launch(); cudaDeviceSynchronize();
      cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
      const double fma = (double)iters * 16 * warps * 32 * sms;
      printf("%s warps/SM=%2d: %.3f ms  %.1f GFMA/s  = %.1f FMA lanes / ns / SM\n", mode ? "FFMA2" : "FFMA ", warps, ms,
             fma / ms / 1e6, fma / ms / 1e6 / sms);
    }
  }
  printf("max SM clock %.0f MHz (divide lanes/ns by the clock under load for lanes per cycle)\n", khz / 1000.0);
  return 0;
}

// How much does the store pattern of the GEMM epilogue's fp32 row-major output cost?  Each warp writes a 32-row x 32-column
// fp32 chunk (rows `ld` floats apart).  Pattern A (the epilogue today): lane = row, 4 x st.global.v8 (32 B) along its row.
// Pattern B: 4 lanes per row, each one st.v8 -> every instruction writes 8 rows x 128 contiguous bytes.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/store_pattern scripts/store_pattern.cu && /tmp/store_pattern
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void st8(float* p, float v) {
  asm volatile("st.global.v8.f32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1};" ::"l"(p), "f"(v) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(256, 1) k(float* out, int64_t M, int ld, int ncol) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int64_t tile = blockIdx.x; tile * 128 < M; tile += gridDim.x) {
    const int64_t row0 = tile * 128 + (warp & 3) * 32;
    for (int c0 = (warp >> 2) * 32; c0 < ncol; c0 += 64) {
      if (MODE == 0) {
        float* p = out + (row0 + lane) * ld + c0;
#pragma unroll
        for (int j = 0; j < 4; ++j) st8(p + 8 * j, (float)lane);
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float* p = out + (row0 + 8 * i + (lane >> 2)) * ld + c0 + 8 * (lane & 3);
          st8(p, (float)lane);
        }
      }
    }
  }
}

int main() {
  const int64_t M = 41088 * 2;
  const int ld = 1024, ncol = 256;
  float* out;
  cudaMalloc(&out, M * ld * sizeof(float));
  int sms = 148;
  for (int mode = 0; mode < 2; ++mode) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) k<0><<<sms, 256>>>(out, M, ld, ncol); else k<1><<<sms, 256>>>(out, M, ld, ncol);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
    }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%s: %.3f ms for %.1f MB = %.0f GB/s\n", mode ? "B: 4 lanes per row (128 B runs)" : "A: lane = row (32 B per lane)", ms,
           M * ncol * 4 / 1e6, M * ncol * 4 / ms / 1e6);
  }
  return 0;
}

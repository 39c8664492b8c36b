// Micro-benchmark: raw tcgen05.mma issue rate on B200 with operands resident in shared memory
// (no TMA traffic).  Answers: what does an SS-mode 128xNx16 bf16 MMA cost with cta_group::1 vs
// cta_group::2, 128B vs 64B swizzle, one vs several accumulators?  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I speech_enhancement_b200/csrc \
//        scripts/mma_rate.cu -o speech_enhancement_b200/csrc/build/mma_rate
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "umma.cuh"

using namespace se;

__device__ __forceinline__ uint64_t mk_desc(uint32_t addr, bool sw128) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)((sw128 ? 1024 : 512) >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(sw128 ? 2 : 4) << 61;
  return d;
}

__device__ __forceinline__ void umma_bf16_cg2(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit_cg2_mc(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(ptx::smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r;
}

// mode bits: sw128 (1), nacc (1,2,4), cg2
template <int cg2>
__global__ void __launch_bounds__(256, 1)
rate_kernel(int N, int sw128, int nacc, int iters, int a_bytes_stride, long long* out, int bg, const uint8_t* gsrc) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ __align__(8) uint64_t bgbar[4];
  __shared__ volatile int stop_flag;
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x3c003c00u + (i & 0xff);
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); for (int i = 0; i < 4; ++i) ptx::mbar_init(&bgbar[i], 1); stop_flag = 0; ptx::fence_mbar_init(); }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if constexpr (cg2 != 0) cluster_sync_all(); else __syncthreads();
  if (warp == 0) {
    if constexpr (cg2 != 0) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(ptx::smem_u32(&tmem_slot)), "r"(512) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      ptx::tmem_alloc(&tmem_slot, 512);
      ptx::tmem_relinquish();
    }
  }
  ptx::tc_fence_before();
  if constexpr (cg2 != 0) cluster_sync_all(); else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tm = tmem_slot;
  uint32_t rank = 0;
  if constexpr (cg2 != 0) rank = cluster_ctarank();
  long long t0 = 0, t1 = 0;
  if (warp == 1 && rank == 0) {
    if (ptx::elect_one()) {
      const uint32_t idesc = ptx::make_idesc_bf16_f32(cg2 ? 256 : 128, N);
      const uint32_t base = ptx::smem_u32(smem);
      t0 = clock64();
      for (int it = 0; it < iters; ++it) {
        // 4 rotating A tiles (16 KB apart by default) and B slabs behind them; 4 k-steps of 32 B each
        const uint32_t a_addr = base + (uint32_t)((it & 3) * a_bytes_stride);
        const uint32_t b_addr = base + 96 * 1024 + (uint32_t)((it & 1) * 32 * 1024);
        const uint64_t da = mk_desc(a_addr, sw128), db = mk_desc(b_addr, sw128);
        const uint32_t d = tm + (uint32_t)((it % nacc) * N);
        const int ks = sw128 ? 4 : 2;
        for (int k = 0; k < ks; ++k) {
          if constexpr (cg2 != 0) umma_bf16_cg2(d, da + 2 * k, db + 2 * k, idesc, 1);
          else ptx::umma_bf16(d, da + 2 * k, db + 2 * k, idesc, 1);
        }
      }
      if constexpr (cg2 != 0) commit_cg2_mc(&bar, 3); else ptx::umma_commit(&bar);
      ptx::mbar_wait(&bar, 0);
      t1 = clock64();
      if (blockIdx.x == 0) { out[0] = t1 - t0; }
      stop_flag = 1;
    }
  } else if (cg2 && warp == 1 && rank == 1) {
    if (ptx::elect_one()) ptx::mbar_wait(&bar, 0);
  }
  if (warp == 2 && (bg & 1)) {
    // background TMA-style traffic: 1-D bulk copies global -> a spare 64 KB smem region, 4 x 16 KB in flight
    if (ptx::elect_one()) {
      uint8_t* dst = smem + 160 * 1024;
      uint32_t n = 0;
      long long bytes = 0;
      while (!stop_flag) {
        const uint32_t sl = n & 3, ph = (n >> 2) & 1;
        if (n >= 4) ptx::mbar_wait(&bgbar[sl], ph ^ 1u);
        ptx::mbar_arrive_expect_tx(&bgbar[sl], 16384);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(ptx::smem_u32(dst + sl * 16384)), "l"(gsrc + ((size_t)(blockIdx.x * 64 + (n & 63)) * 16384)),
                       "r"(16384), "r"(ptx::smem_u32(&bgbar[sl])) : "memory");
        ++n; bytes += 16384;
      }
      for (uint32_t k = (n >= 4 ? n - 4 : 0); k < n; ++k) ptx::mbar_wait(&bgbar[k & 3], (k >> 2) & 1);
      if (blockIdx.x == 0) out[1] = bytes;
    }
  }
  if (warp >= 4 && (bg & 2)) {
    // background epilogue-style TMEM reads: warps 4..7 own lane quarters 0..3
    uint32_t r[32];
    uint32_t acc = 0;
    long long cnt = 0;
    while (!stop_flag) {
      ptx::tmem_ld_32x32(tm + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((cnt * 32) % N), r);
      ptx::tmem_ld_wait();
      acc += r[0] ^ r[31];
      ++cnt;
    }
    if (acc == 0x12345678u) out[3] = acc;
    if (blockIdx.x == 0 && threadIdx.x == 128) out[2] = cnt;
  }
  __syncwarp();
  ptx::tc_fence_before();
  if constexpr (cg2 != 0) cluster_sync_all(); else __syncthreads();
  if (warp == 0) {
    if constexpr (cg2 != 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(512) : "memory");
    else ptx::tmem_dealloc(tm, 512);
  }
}

int main() {
  long long* d_out; cudaMalloc(&d_out, 64);
  const int smem = 226 * 1024;
  uint8_t* gsrc; cudaMalloc(&gsrc, (size_t)148 * 64 * 16384); cudaMemset(gsrc, 0x3c, (size_t)148 * 64 * 16384);
  cudaFuncSetAttribute(rate_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(rate_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  int nsm = 148;
  const int iters = 2000;
  for (int bg = 0; bg < 4; ++bg)
  for (int cg2 = 0; cg2 <= 1; ++cg2)
    for (int sw128 = 1; sw128 >= (bg ? 1 : 0); --sw128)
      for (int N : {256, 128, 64, 32})
        for (int nacc : {1, 2}) {
          if (bg && (cg2 || nacc != 1)) continue;
          cudaMemset(d_out, 0, 64);
          if (nacc * N > 512) continue;
          cudaLaunchConfig_t cfg{};
          cfg.gridDim = dim3(cg2 ? (nsm / 2) * 2 : nsm); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = smem;
          cudaLaunchAttribute at[1];
          at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cg2 ? 2 : 1;
          at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
          cfg.attrs = at; cfg.numAttrs = cg2 ? 1 : 0;
          int a_stride = 16 * 1024;
          cudaError_t e = cg2 ? cudaLaunchKernelEx(&cfg, rate_kernel<1>, N, sw128, nacc, iters, a_stride, d_out, bg, (const uint8_t*)gsrc)
                              : cudaLaunchKernelEx(&cfg, rate_kernel<0>, N, sw128, nacc, iters, a_stride, d_out, bg, (const uint8_t*)gsrc);
          cudaError_t e2 = cudaDeviceSynchronize();
          long long o4[4] = {0, 0, 0, 0}; cudaMemcpy(o4, d_out, 32, cudaMemcpyDeviceToHost);
          long long cyc = o4[0];
          const int ks = sw128 ? 4 : 2;
          double per = (double)cyc / (iters * ks);
          double floor_c = 128.0 * N / 256.0;
          printf("bg%d cg%d sw%-3d N=%3d nacc=%d : %8.1f cyc/MMA (floor %5.1f, %5.1f%%)  bgTMA %.1f B/clk  bgLD %.0f cyc/ld [%s %s]\n", bg, cg2 + 1, sw128 ? 128 : 64, N, nacc,
                 per, floor_c, 100.0 * floor_c / per, (double)o4[1] / (double)cyc, o4[2] ? (double)cyc / (double)o4[2] : 0.0,
                 cudaGetErrorString(e), cudaGetErrorString(e2));
          if (e2 != cudaSuccess) return 1;
        }
  return 0;
}

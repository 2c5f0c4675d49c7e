// Micro-benchmark: tcgen05.mma issue-to-completion rate for the shapes the prefill kernel uses.
//   SS: A (128 x 16) and B (N x 16) from 128B-swizzled shared memory (QK^T)
//   TS: A from TMEM, B (16 x N, MN-major) from shared memory (PV)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t umma_desc(uint32_t a, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0; d |= (uint64_t)((a & 0x3FFFFu) >> 4); d |= (uint64_t)((lbo >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFFu) << 32; d |= (uint64_t)1 << 46; d |= (uint64_t)2 << 61; return d;
}
__host__ __device__ constexpr uint32_t idesc(bool b_mn, int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((b_mn ? 1u : 0u) << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t id, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(id), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t id, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(b), "r"(id), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void wait(uint64_t* bar, uint32_t ph) {
  uint32_t ok = 0;
  while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(ph) : "memory");
}
// MODE 0: SS N=128; 1: TS N=128; 2: alternate 8 TS + 8 SS (the kernel's PV,QK order); 3: SS N=64; 4: SS N=256; 5: SS with A and B the same buffer
template <int MODE>
__global__ void __launch_bounds__(128, 1) k(long long* cyc, int iters) {
  extern __shared__ __align__(1024) uint8_t raw[];
  uint8_t* sm = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar; __shared__ uint32_t tptr;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) ((uint32_t*)sm)[i] = 0;
  if (threadIdx.x == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;"); }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tptr)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tm = tptr;
  if (threadIdx.x == 0) {
    const uint32_t q = smem_u32(sm), kb = smem_u32(sm) + 32768, vb = smem_u32(sm) + 65536 + 32768;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (MODE == 0 || MODE == 3 || MODE == 4 || MODE == 5) {
        const int N = MODE == 3 ? 64 : (MODE == 4 ? 256 : 128);
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
          mma_ss(tm + (it & 1) * 256, umma_desc(q + off, 16, 1024), umma_desc((MODE == 5 ? q : kb) + off * (N == 256 ? 2 : 1), 16, 1024), idesc(false, 128, N), kk > 0);
        }
      } else if (MODE == 1) {
        for (int kk = 0; kk < 8; ++kk) mma_ts(tm + 256 + (it & 1) * 128, tm + kk * 8, umma_desc(vb + kk * 2048, 16384, 1024), idesc(true, 128, 128), 1);
      } else {
        for (int kk = 0; kk < 8; ++kk) mma_ts(tm + 256 + (it & 1) * 128, tm + (it & 1) * 128 + kk * 8, umma_desc(vb + kk * 2048, 16384, 1024), idesc(true, 128, 128), 1);
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
          mma_ss(tm + (it & 1) * 128, umma_desc(q + (it & 1) * 32768 * 0 + off, 16, 1024), umma_desc(kb + off, 16, 1024), idesc(false, 128, 128), kk > 0);
        }
      }
    }
    commit(&bar);
    wait(&bar, 0);
    long long t1 = clock64();
    cyc[blockIdx.x] = t1 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}
template <int MODE> void run(const char* name, int mmas_per_iter, double flop_per_mma) {
  long long* cyc; cudaMalloc(&cyc, 148 * 8);
  const int iters = 2000;
  cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  k<MODE><<<148, 128, 200 * 1024>>>(cyc, iters); cudaDeviceSynchronize();
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0); k<MODE><<<148, 128, 200 * 1024>>>(cyc, iters); cudaEventRecord(e1);
  cudaError_t e = cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-34s %s  %.1f cyc per MMA   (%.0f TFLOP/s chip, kernel %.3f ms)\n", name, cudaGetErrorString(e),
         (double)h[0] / ((double)iters * mmas_per_iter), 148.0 * iters * mmas_per_iter * flop_per_mma / (ms * 1e-3) / 1e12, ms);
}
int main() {
  const double f128 = 2.0 * 128 * 128 * 16;
  run<0>("SS M128 N128 K16 (QK^T)", 8, f128);
  run<5>("SS M128 N128 K16, A==B buffer", 8, f128);
  run<1>("TS M128 N128 K16 (PV)", 8, f128);
  run<2>("8 TS + 8 SS alternating", 16, f128);
  run<3>("SS M128 N64 K16", 8, f128 / 2);
  run<4>("SS M128 N256 K16", 8, f128 * 2);
  return 0;
}

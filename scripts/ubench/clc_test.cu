#include <cstdint>
#include <cstdio>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void k(int* out, int* count) {
  extern __shared__ char big[];
  __shared__ __align__(16) uint4 resp;
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  __syncthreads();
  int bx = blockIdx.x, by = blockIdx.y;
  uint32_t phase = 0;
  int n = 0;
  while (true) {
    if (threadIdx.x == 0) {
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], 16;" ::"r"(smem_u32(&bar)) : "memory");
      asm volatile("clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes.b128 [%0], [%1];" ::"r"(smem_u32(&resp)), "r"(smem_u32(&bar)) : "memory");
    }
    // work
    if (threadIdx.x == 0) { atomicAdd(&out[by * gridDim.x + bx], 1); big[bx] = 1; }
    { long long t0 = clock64(); while (clock64() - t0 < 20000) {} }
    n++;
    uint32_t ok = 0;
    while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(&bar)), "r"(phase) : "memory");
    phase ^= 1;
    uint32_t valid, nx, ny;
    asm volatile("{\n\t.reg .b128 r;\n\t.reg .pred p;\n\tld.shared.b128 r, [%3];\n\tclusterlaunchcontrol.query_cancel.is_canceled.pred.b128 p, r;\n\tselp.u32 %0, 1, 0, p;\n\t@p clusterlaunchcontrol.query_cancel.get_first_ctaid.v4.b32.b128 {%1, %2, _, _}, r;\n\t}" : "=r"(valid), "=r"(nx), "=r"(ny) : "r"(smem_u32(&resp)) : "memory");
    __syncthreads();
    if (!valid) break;
    bx = nx, by = ny;
  }
  if (threadIdx.x == 0) atomicAdd(count, n > 1 ? 1 : 0);
}
int main() {
  int *out, *cnt; cudaMalloc(&out, 129 * 8 * 4); cudaMalloc(&cnt, 4); cudaMemset(out, 0, 129 * 8 * 4); cudaMemset(cnt, 0, 4);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  k<<<dim3(129, 8), 128, 200 * 1024>>>(out, cnt);
  cudaError_t e = cudaDeviceSynchronize();
  int h[129 * 8], c; cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost); cudaMemcpy(&c, cnt, 4, cudaMemcpyDeviceToHost);
  int bad = 0; for (int i = 0; i < 129 * 8; ++i) bad += h[i] != 1;
  printf("%s: items not processed exactly once: %d; CTAs that stole work: %d\n", cudaGetErrorString(e), bad, c);
}

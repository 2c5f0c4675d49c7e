// Micro-benchmark: issue throughput of the instructions the softmax loop is made of
// (one CTA per SM, W warps per SMSP).  Prints cycles per warp-instruction per SMSP.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define ITERS 4096
template <int OP>
__global__ void k(float* out, long long* cyc, float seed) {
  float a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  uint64_t p0, p1, p2, p3, p4, p5, p6, p7, c;
  asm("mov.b64 %0, {%1,%2};" : "=l"(p0) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1,%2};" : "=l"(p1) : "f"(a1), "f"(a2));
  asm("mov.b64 %0, {%1,%2};" : "=l"(p2) : "f"(a2), "f"(a3));
  asm("mov.b64 %0, {%1,%2};" : "=l"(p3) : "f"(a3), "f"(a4));
  asm("mov.b64 %0, {%1,%2};" : "=l"(p4) : "f"(a4), "f"(a5));
  asm("mov.b64 %0, {%1,%2};" : "=l"(p5) : "f"(a5), "f"(a6));
  asm("mov.b64 %0, {%1,%2};" : "=l"(p6) : "f"(a6), "f"(a7));
  asm("mov.b64 %0, {%1,%2};" : "=l"(p7) : "f"(a7), "f"(a0));
  asm("mov.b64 %0, {%1,%2};" : "=l"(c) : "f"(1.0001f), "f"(0.9999f));
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < ITERS; ++i) {
    if (OP == 0) {  // FFMA (3 reg)
#define F(x) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x) : "f"(a7 * 0 + 1.0001f), "f"(seed));
      F(a0) F(a1) F(a2) F(a3) F(a4) F(a5) F(a6) F(a0) F(a1) F(a2) F(a3) F(a4) F(a5) F(a6) F(a0) F(a1)
#undef F
    } else if (OP == 1) {  // FFMA2
#define F(x) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(x) : "l"(c));
      F(p0) F(p1) F(p2) F(p3) F(p4) F(p5) F(p6) F(p7) F(p0) F(p1) F(p2) F(p3) F(p4) F(p5) F(p6) F(p7)
#undef F
    } else if (OP == 2) {  // FADD2
#define F(x) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(x) : "l"(c));
      F(p0) F(p1) F(p2) F(p3) F(p4) F(p5) F(p6) F(p7) F(p0) F(p1) F(p2) F(p3) F(p4) F(p5) F(p6) F(p7)
#undef F
    } else if (OP == 3) {  // MUFU.EX2
#define F(x) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x));
      F(a0) F(a1) F(a2) F(a3) F(a4) F(a5) F(a6) F(a7) F(a0) F(a1) F(a2) F(a3) F(a4) F(a5) F(a6) F(a7)
#undef F
    } else if (OP == 4) {  // FMNMX3
#define F(x, y, z) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x) : "f"(y), "f"(z));
      F(a0, a1, a2) F(a1, a2, a3) F(a2, a3, a4) F(a3, a4, a5) F(a4, a5, a6) F(a5, a6, a7) F(a6, a7, a0) F(a7, a0, a1)
      F(a0, a1, a2) F(a1, a2, a3) F(a2, a3, a4) F(a3, a4, a5) F(a4, a5, a6) F(a5, a6, a7) F(a6, a7, a0) F(a7, a0, a1)
#undef F
    } else if (OP == 5) {  // F2FP bf16x2
      uint32_t r;
#define F(x, y) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(x), "f"(y)); x = __uint_as_float(r);
      F(a0, a1) F(a1, a2) F(a2, a3) F(a3, a4) F(a4, a5) F(a5, a6) F(a6, a7) F(a7, a0)
      F(a0, a1) F(a1, a2) F(a2, a3) F(a3, a4) F(a4, a5) F(a5, a6) F(a6, a7) F(a7, a0)
#undef F
    } else if (OP == 6) {  // FADD (2 reg + 1)
#define F(x) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(x) : "f"(seed));
      F(a0) F(a1) F(a2) F(a3) F(a4) F(a5) F(a6) F(a7) F(a0) F(a1) F(a2) F(a3) F(a4) F(a5) F(a6) F(a7)
#undef F
    } else if (OP == 7) {  // mixed: 1 MUFU + 3 FFMA2
#define M(x) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x));
#define F(x) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(x) : "l"(c));
      M(a0) F(p0) F(p1) F(p2) M(a1) F(p3) F(p4) F(p5) M(a2) F(p6) F(p7) F(p0) M(a3) F(p1) F(p2) F(p3)
#undef F
#undef M
    } else if (OP == 8) {  // FMNMX (2-input)
#define F(x, y) asm volatile("max.f32 %0, %0, %1;" : "+f"(x) : "f"(y));
      F(a0, a1) F(a1, a2) F(a2, a3) F(a3, a4) F(a4, a5) F(a5, a6) F(a6, a7) F(a7, a0)
      F(a0, a1) F(a1, a2) F(a2, a3) F(a3, a4) F(a4, a5) F(a5, a6) F(a6, a7) F(a7, a0)
#undef F
    } else if (OP == 9) {  // mixed: 1 MUFU + 1 FFMA2 + 1 FMNMX3 + 1 F2FP (the MUFU softmax mix)
      uint32_t r;
#define M(x) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x));
#define F(x) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(x) : "l"(c));
#define X(x, y, z) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x) : "f"(y), "f"(z));
#define C(x, y) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(x), "f"(y)); x = __uint_as_float(r);
      M(a0) F(p0) X(a4, a5, a6) C(a7, a6) M(a1) F(p1) X(a5, a6, a4) C(a6, a7) M(a2) F(p2) X(a4, a5, a6) C(a7, a6) M(a3) F(p3) X(a5, a6, a4) C(a6, a7)
#undef M
#undef F
#undef X
#undef C
    }
  }
  long long t1 = clock64();
  float s0, s1;
  asm("mov.b64 {%0,%1}, %2;" : "=f"(s0), "=f"(s1) : "l"(p0 ^ p1 ^ p2 ^ p3 ^ p4 ^ p5 ^ p6 ^ p7));
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 + s0 + s1;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int OP> void run(const char* name, int warps_per_smsp) {
  float* out; long long* cyc;
  int threads = 128 * warps_per_smsp;
  cudaMalloc(&out, 148 * threads * 4); cudaMalloc(&cyc, 148 * 8);
  k<OP><<<148, threads>>>(out, cyc, 0.5f); cudaDeviceSynchronize();
  k<OP><<<148, threads>>>(out, cyc, 0.5f); cudaDeviceSynchronize();
  long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double per = (double)h[0] / ((double)ITERS * 16 * warps_per_smsp);
  printf("%-28s warps/SMSP=%d  %.2f cyc per warp-instr per SMSP\n", name, warps_per_smsp, per);
  cudaFree(out); cudaFree(cyc);
}
int main() {
  for (int w = 1; w <= 4; w *= 2) {
    run<0>("FFMA", w); run<1>("FFMA2", w); run<2>("FADD2", w); run<6>("FADD", w); run<3>("MUFU.EX2", w); run<4>("FMNMX3", w);
    run<8>("FMNMX", w); run<5>("F2FP.BF16x2", w); run<7>("1 MUFU + 3 FFMA2", w); run<9>("MUFU+FFMA2+FMNMX3+F2FP", w);
  }
  return 0;
}

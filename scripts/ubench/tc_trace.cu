// Debug harness: runs the tcgen05 prefill kernel (built with -DB200ATTN_TC_TRACE) on one
// 8192-token causal prompt and prints the clock64 stamps of CTA (0,0)'s pipeline events.
//   nvcc -DB200ATTN_TC_TRACE ... scripts/ubench/tc_trace.cu vllm_triton_backend_b200/csrc/{cabi,tma_maps,...}.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "../../vllm_triton_backend_b200/csrc/attn_prefill_tc.cu"
using namespace b200attn;
int main(int argc, char** argv) {
  const int T = 8192, Hq = 32, Hkv = 8, D = 128, bs = 16, NB = T / bs + 8;
  const int variant = argc > 1 ? atoi(argv[1]) : 1;
  b200attn_set_tuning(0, 0, 0, variant << 4);
  void *q, *out, *kc, *vc; int *bt, *sl, *cu;
  cudaMalloc(&q, (size_t)T * Hq * D * 2); cudaMalloc(&out, (size_t)T * Hq * D * 2);
  cudaMalloc(&kc, (size_t)NB * bs * Hkv * D * 2); cudaMalloc(&vc, (size_t)NB * bs * Hkv * D * 2);
  cudaMemset(q, 0x3c, (size_t)T * Hq * D * 2); cudaMemset(kc, 0x3c, (size_t)NB * bs * Hkv * D * 2); cudaMemset(vc, 0x3c, (size_t)NB * bs * Hkv * D * 2);
  std::vector<int> hbt(T / bs); for (int i = 0; i < T / bs; ++i) hbt[i] = (i * 37) % (T / bs);
  int hsl = T, hcu[2] = {0, T};
  cudaMalloc(&bt, hbt.size() * 4); cudaMalloc(&sl, 4); cudaMalloc(&cu, 8);
  cudaMemcpy(bt, hbt.data(), hbt.size() * 4, cudaMemcpyHostToDevice); cudaMemcpy(sl, &hsl, 4, cudaMemcpyHostToDevice); cudaMemcpy(cu, hcu, 8, cudaMemcpyHostToDevice);
  b200attn_attention_params p = {};
  p.out = out; p.q = q; p.k_cache = kc; p.v_cache = vc; p.block_tables = bt; p.seq_lens = sl; p.cu_seqlens_q = cu;
  p.softmax_scale = 0.088f; p.num_tokens = T; p.num_seqs = 1; p.num_q_heads = Hq; p.num_kv_heads = Hkv; p.head_size = D; p.block_size = bs;
  p.max_seqlen_q = T; p.max_seqlen_k = T; p.q_dtype = B200ATTN_BF16; p.kv_dtype = B200ATTN_BF16; p.path = B200ATTN_PATH_PREFILL_TC;
  p.num_blocks = NB; p.q_stride0 = Hq * D; p.q_stride1 = D; p.out_stride0 = Hq * D; p.out_stride1 = D; p.bt_stride0 = T / bs;
  p.kc_block_stride = p.vc_block_stride = (int64_t)bs * Hkv * D; p.kc_slot_stride = p.vc_slot_stride = Hkv * D; p.kc_head_stride = p.vc_head_stride = D;
  cudaEvent_t ev0, ev1; cudaEventCreate(&ev0); cudaEventCreate(&ev1);
  for (int it = 0; it < 3; ++it) {
    if (it == 2) cudaEventRecord(ev0);
    int rc = b200attn_unified_attention(&p, nullptr);
    if (rc) { printf("rc=%d %s\n", rc, b200attn_last_error()); return 1; }
  }
  cudaEventRecord(ev1);
  cudaError_t e = cudaDeviceSynchronize();
  float kms; cudaEventElapsedTime(&kms, ev0, ev1);
  {
    static long long ct[4096 * 4];
    cudaMemcpyFromSymbol(ct, g_tc_cta, sizeof(ct));
    long long tmin = 1LL << 62, tmax = 0; double busy[160] = {0}; long long tiles[160] = {0}; int ncta[160] = {0};
    double per_tile_sum = 0; int nw = 0; double ovh = 0;
    for (int c = 0; c < 4096; ++c) {
      if (ct[c * 4 + 3] <= 0 || ct[c * 4 + 1] == 0) continue;
      int sm = (int)ct[c * 4]; long long a0 = ct[c * 4 + 1], a1 = ct[c * 4 + 2];
      if (a0 < tmin) tmin = a0; if (a1 > tmax) tmax = a1;
      busy[sm] += (double)(a1 - a0); tiles[sm] += ct[c * 4 + 3]; ncta[sm]++;
      nw++;
    }
    double bsum = 0, bmax = 0, bmin = 1e30; long long tsum = 0;
    for (int i = 0; i < 160; ++i) if (ncta[i]) { bsum += busy[i]; tsum += tiles[i]; if (busy[i] > bmax) bmax = busy[i]; if (busy[i] < bmin) bmin = busy[i]; }
    printf("kernel %.1f us (event), CTA span %.1f us; %d working CTAs, %lld tile-pairs; SM busy avg %.1f us min %.1f max %.1f; ns per tile-pair (busy/tiles) %.1f\n",
           kms * 1e3, (tmax - tmin) * 1e-3, nw, tsum, bsum / 148 * 1e-3, bmin * 1e-3, bmax * 1e-3, bsum / tsum);
    // fit busy = a + b * tiles over CTAs
    double sx = 0, sy = 0, sxx = 0, sxy = 0; int n = 0;
    for (int c = 0; c < 4096; ++c) { if (ct[c * 4 + 3] <= 0 || ct[c * 4 + 1] == 0) continue; double x = (double)ct[c * 4 + 3], y = (double)(ct[c * 4 + 2] - ct[c * 4 + 1]); sx += x; sy += y; sxx += x * x; sxy += x * y; n++; }
    double bb = (n * sxy - sx * sy) / (n * sxx - sx * sx), aa = (sy - bb * sx) / n;
    printf("per-CTA time ~= %.0f ns + %.1f ns x tiles  (n=%d)\n", aa, bb, n);
  }
  printf("sync: %s\n", cudaGetErrorString(e));
  static long long tr[65 * 32];
  cudaMemcpyFromSymbol(tr, g_tc_trace, sizeof(tr));
  const char* names[21] = {"s0.wake", "s0.ld", "s0.max", "s0.Pa", "s0.Pb", "s1.wake", "s1.ld", "s1.max", "s1.Pa", "s1.Pb", "m.Pa0", "m.Pb0", "m.QK0", "m.Pa1", "m.Pb1", "m.QK1", "m.Kw", "m.Kok", "m.Vok", "t.Kfree", "t.Vfree"};
  {
    const char* mk[8] = {"start", "setup done", "S(0) ready", "S(1) ready", "S(last) ready", "softmax1 loop done", "O(last) done", "teardown"};
    printf("CTA(1,0) phases (cycles since start):");
    for (int k2 = 0; k2 < 8; ++k2) printf("  %s=%lld", mk[k2], tr[64 * 32 + k2] - tr[64 * 32]);
    printf("\n");
  }
  long long base = tr[20 * 32 + 10];
  printf("tile  "); for (int e2 = 0; e2 < 21; ++e2) printf("%8s", names[e2]); printf("\n");
  for (int j = 20; j < 28; ++j) { printf("%4d  ", j); for (int e2 = 0; e2 < 21; ++e2) printf("%8lld", tr[j * 32 + e2] - base); printf("\n"); }
  printf("period (m.Pa0 j=40 - j=20)/20 = %.0f cycles\n", (double)(tr[40 * 32 + 10] - tr[20 * 32 + 10]) / 20);
  return 0;
}

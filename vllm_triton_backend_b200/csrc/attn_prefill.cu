// K3 (general): varlen / prefix / chunked-prefill / mixed-batch paged attention.
//
// Computes what the reference's 2D kernel computes
// (reference ibm-triton-lib/ibm_triton_lib/kernels/triton_unified_attention.py:
//  kernel_unified_attention_2d :275-523): per (q-block, kv-head), rows of the
// tile are (query token, query head of the GQA group) pairs, KV tiles are
// gathered through the block table, causal + context mask, optional softcap /
// sliding window / ALiBi, online softmax in fp32, acc / L epilogue.
// The launch grid is the same host-int upper bound the reference uses
// (T // BLOCK_Q + num_seqs q-blocks, :886-889) with the same binary search from
// q-block to sequence (:32-52), so it needs no host sync and is graph safe.
//
// Implementation: 4 warps x 16 rows, 64-token KV tiles moved with cp.async into
// XOR-swizzled shared memory (double buffered), mma.sync m16n8k16 with fp32
// accumulation; FP8-e4m3 KV is converted to f16 on its way into shared memory
// and the per-tensor scales are folded into the softmax scale / epilogue.
// This kernel takes every head size in {64,128,256} and any GQA group <= 64; the
// tcgen05/TMEM kernel in attn_prefill_tc.cu is the fast path for D=128.
#include "common.cuh"

namespace b200attn {
namespace {

constexpr int BM = 64, BN = 64, NWARPS = 4;

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// byte offset of 16-byte chunk `c` of row `r` in a [rows][ROWB] tile, XOR swizzled
template <int ROWB> __device__ __forceinline__ uint32_t swz_off(int r, int c) {
  return (uint32_t)(r * ROWB + ((c ^ (r & 7)) << 4));
}

// kPlain: no softcap, no ALiBi and a positive scale -- the scale then folds into the exp2 argument
// (one FFMA per logit) and the row max is taken over the raw dot products.
// bs_shift: log2(block_size) when it is a power of two (page = pos >> bs_shift), else -1 (division).
template <int D, bool MMA_BF16, bool Q_BF16, bool KV_FP8, bool kPlain>
__global__ void __launch_bounds__(NWARPS * 32, (D <= 128) ? 2 : 1)
attn_prefill_kernel(const b200attn_attention_params p, const int block_q, const int group, const int min_q_len,
                    const int max_q_len, const int qb_per_seq, const int32_t* skip_staged, const int bs_shift) {
  constexpr int ROWB = D * 2;           // smem row bytes (operands are 2-byte in smem)
  constexpr int CPR = ROWB / 16;        // 16-byte chunks per row
  constexpr int KSTEPS = D / 16;
  constexpr int DT = D / 8;             // n8 tiles of the output
  constexpr int TILE_B = BN * ROWB;
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* sQ = smem;                   // [BM][ROWB]
  uint8_t* sK = sQ + BM * ROWB;         // [2][BN][ROWB]
  uint8_t* sV = sK + 2 * TILE_B;        // [2][BN][ROWB]
  float* sKs = reinterpret_cast<float*>(sV + 2 * TILE_B);  // [2][BN] per-(token, head) key scales (pth only)
  float* sVs = sKs + 2 * BN;                               // [2][BN] value scales

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int gid = lane >> 2, t4 = lane & 3;
  const int q_block_global = blockIdx.x, kvh = blockIdx.y;

  // ---- q-block -> sequence ------------------------------------------------------
  int b, q_block_local;
  if (qb_per_seq > 0) {
    // launch restricted to short queries: a fixed number of q-blocks per sequence, no search
    // (most CTAs of a mixed batch exit here, so the exit has to be cheap)
    b = q_block_global / qb_per_seq, q_block_local = q_block_global % qb_per_seq;
  } else {
    // binary search over cu_seqlens_q // BLOCK_Q + idx (reference :32-52)
    int left = 0, right = p.num_seqs;
    while (left < right) {
      const int mid = (left + right) >> 1;
      const int mid_val = p.cu_seqlens_q[mid] / block_q + mid;
      if (mid_val <= q_block_global) left = mid + 1; else right = mid;
    }
    b = left - 1;
    q_block_local = q_block_global - (p.cu_seqlens_q[b] / block_q + b);
  }
  const int q_start = p.cu_seqlens_q[b];
  const int q_len = p.cu_seqlens_q[b + 1] - q_start;
  if (q_block_local * block_q >= q_len) return;
  if (q_len < min_q_len || q_len > max_q_len) return;  // another launch covers this sequence
  if (skip_staged && skip_staged[b] >= 0) return;      // staged for the tcgen05 kernel (kv_stage.cu)
  const int seq_len = p.seq_lens[b];
  const int ctx = seq_len - q_len;
  const int qpos0 = q_block_local * block_q;                   // first query position of this block
  const int qpos_last = min(qpos0 + block_q, q_len) - 1;       // last valid one
  int kv_limit = min(ctx + qpos_last + 1, seq_len);            // keys [0, kv_limit) can be visible
  int kv_first = 0;
  if (p.sliding_window > 0) kv_first = max(0, ctx + qpos0 - p.sliding_window + 1);
  const int tile0 = kv_first / BN;
  const int n_tiles = (kv_limit + BN - 1) / BN - tile0;

  const int32_t* bt = p.block_tables + (int64_t)b * p.bt_stride0;
  // per-(token, head) scales (b200attn.h k_scale_cache; general-logits variant only): K and V are used as stored,
  // every key's own scale multiplies its logit and every value's own scale multiplies P before the rounding
  const bool pth = KV_FP8 && !kPlain && p.k_scale_cache != nullptr;
  const bool kv_i8 = KV_FP8 && p.kv_dtype == B200ATTN_INT8;
  const float k_scale = (KV_FP8 && !pth && p.k_scale) ? *p.k_scale : 1.0f;
  const float v_scale = (KV_FP8 && !pth && p.v_scale) ? *p.v_scale : 1.0f;
  const bool has_cap = p.softcap > 0.f;
  const float inv_cap = has_cap ? 1.f / p.softcap : 0.f;
  const float qk_scale = p.softmax_scale * k_scale;
  const float scale_log2 = qk_scale * kLog2e;

  // ---- load the Q tile: row r = (token r / group, head r % group) --------------
  const int rows_used = block_q * group;
  for (int idx = tid; idx < BM * CPR; idx += NWARPS * 32) {
    const int r = idx / CPR, c = idx % CPR;
    const int tok = qpos0 + r / group, hq = kvh * group + r % group;
    const bool ok = r < rows_used && tok < q_len;
    int4 v = make_int4(0, 0, 0, 0);
    if (ok) {
      const char* src = static_cast<const char*>(p.q) + ((int64_t)(q_start + tok) * p.q_stride0 + (int64_t)hq * p.q_stride1) * 2 + c * 16;
      v = *reinterpret_cast<const int4*>(src);
      if constexpr (Q_BF16 && !MMA_BF16) {
        uint32_t* w = reinterpret_cast<uint32_t*>(&v);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 f = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&w[i]));
          w[i] = pack2<false>(fminf(fmaxf(f.x, -65504.f), 65504.f), fminf(fmaxf(f.y, -65504.f), 65504.f));
        }
      }
    }
    *reinterpret_cast<int4*>(sQ + swz_off<ROWB>(r, c)) = v;
  }

  // ---- KV tile loader -----------------------------------------------------------
  // Thread `tid` always copies the same 16-byte column of rows ld_r0, ld_r0 + RSTEP, ...: the column, the
  // head offset and the swizzle are hoisted out of the tile loop; per row there is one block-table load and
  // one 32x32->64-bit multiply-add per cache (strides < 2^31 elements, checked on the host).
  constexpr int ESZ = KV_FP8 ? 1 : 2;
  constexpr int CPT = KV_FP8 ? D / 16 : CPR;       // 16-byte global chunks per cache row
  constexpr int RSTEP = NWARPS * 32 / CPT;         // rows between two chunks of one thread
  const int ld_c = tid % CPT, ld_r0 = tid / CPT;
  const char* k_col = static_cast<const char*>(p.k_cache) + ((int64_t)kvh * p.kc_head_stride) * ESZ + ld_c * 16;
  const char* v_col = static_cast<const char*>(p.v_cache) + ((int64_t)kvh * p.vc_head_stride) * ESZ + ld_c * 16;
  const int kc_bs = (int)p.kc_block_stride * ESZ, kc_ss = (int)p.kc_slot_stride * ESZ;
  const int vc_bs = (int)p.vc_block_stride * ESZ, vc_ss = (int)p.vc_slot_stride * ESZ;
  const int bs_mask = p.block_size - 1;
  auto load_tile = [&](int tile, int buf) {
    const int n0 = tile * BN;
    uint8_t* dK = sK + buf * TILE_B;
    uint8_t* dV = sV + buf * TILE_B;
#pragma unroll
    for (int k = 0; k < BN / RSTEP; ++k) {
      const int r = ld_r0 + k * RSTEP;
      const int pos = n0 + r;
      const bool ok = pos < kv_limit;
      int page = 0, slot = 0;
      if (ok) {
        if (bs_shift >= 0) page = bt[pos >> bs_shift], slot = pos & bs_mask;
        else page = bt[pos / p.block_size], slot = pos % p.block_size;
      }
      const char* ksrc = k_col + ((int64_t)page * kc_bs + (int64_t)slot * kc_ss);
      const char* vsrc = v_col + ((int64_t)page * vc_bs + (int64_t)slot * vc_ss);
      if constexpr (!KV_FP8) {
        const uint32_t off = swz_off<ROWB>(r, ld_c);
        cp_async_16(smem_u32(dK + off), ksrc, ok);
        cp_async_16(smem_u32(dV + off), vsrc, ok);
      } else {
        // 16 fp8 -> 16 f16 = two 16-byte smem chunks
        int4 kraw = make_int4(0, 0, 0, 0), vraw = kraw;
        if (ok) {
          kraw = *reinterpret_cast<const int4*>(ksrc);
          vraw = *reinterpret_cast<const int4*>(vsrc);
        }
        const uint32_t* kw = reinterpret_cast<const uint32_t*>(&kraw);
        const uint32_t* vw = reinterpret_cast<const uint32_t*>(&vraw);
        if (pth && ld_c == 0) {  // one thread per row also brings the row's two scales
          float kq = 1.f, vq = 1.f;
          if (ok) {
            kq = p.k_scale_cache[(int64_t)page * p.ks_block_stride + (int64_t)slot * p.ks_slot_stride + (int64_t)kvh * p.ks_head_stride];
            vq = p.v_scale_cache[(int64_t)page * p.vs_block_stride + (int64_t)slot * p.vs_slot_stride + (int64_t)kvh * p.vs_head_stride];
          }
          sKs[buf * BN + r] = kq, sVs[buf * BN + r] = vq;
        }
        auto cv = [&](uint32_t x) { return kv_i8 ? s8x2_to_f16x2((uint16_t)x) : e4m3x2_to_f16x2((uint16_t)x); };
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          int4 ko, vo;
          ko.x = cv(kw[2 * h] & 0xffff), ko.y = cv(kw[2 * h] >> 16);
          ko.z = cv(kw[2 * h + 1] & 0xffff), ko.w = cv(kw[2 * h + 1] >> 16);
          vo.x = cv(vw[2 * h] & 0xffff), vo.y = cv(vw[2 * h] >> 16);
          vo.z = cv(vw[2 * h + 1] & 0xffff), vo.w = cv(vw[2 * h + 1] >> 16);
          *reinterpret_cast<int4*>(dK + swz_off<ROWB>(r, 2 * ld_c + h)) = ko;
          *reinterpret_cast<int4*>(dV + swz_off<ROWB>(r, 2 * ld_c + h)) = vo;
        }
      }
    }
  };

  if (n_tiles > 0) load_tile(tile0, 0);
  cp_async_commit();

  // ---- per-thread row bookkeeping (rows gid and gid+8 of this warp's 16) ---------
  int row_qpos[2], row_ok[2];
  float row_slope[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int r = warp * 16 + gid + 8 * i;
    row_qpos[i] = qpos0 + r / group;
    row_ok[i] = (r < rows_used) && (row_qpos[i] < q_len);
    row_slope[i] = (p.alibi_slopes && row_ok[i]) ? p.alibi_slopes[kvh * group + r % group] : 0.f;
  }
  // rows of this warp span query positions [wq_min, wq_max]
  const int wq_min = qpos0 + (warp * 16) / group;
  const int wq_max = min(qpos0 + (warp * 16 + 15) / group, qpos_last);

  float oacc[DT][4];
#pragma unroll
  for (int i = 0; i < DT; ++i) oacc[i][0] = oacc[i][1] = oacc[i][2] = oacc[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  constexpr bool kQInRegs = (D <= 128);  // D=256: re-read Q fragments from smem to stay under 255 registers
  uint32_t qf[kQInRegs ? KSTEPS : 1][4];
  bool q_loaded = false;

  for (int it = 0; it < n_tiles; ++it) {
    const int buf = it & 1;
    if (it + 1 < n_tiles) load_tile(tile0 + it + 1, buf ^ 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    if (kQInRegs && !q_loaded) {
      // Q fragments stay in registers for the whole KV loop
#pragma unroll
      for (int ks = 0; ks < (kQInRegs ? KSTEPS : 1); ++ks) {
        const int r = warp * 16 + (lane & 7) + 8 * ((lane >> 3) & 1);
        const int c = 2 * ks + (lane >> 4);
        ldmatrix_x4(qf[ks], smem_u32(sQ + swz_off<ROWB>(r, c)));
      }
      q_loaded = true;
    }
    const int n0 = (tile0 + it) * BN;
    const uint32_t kb = smem_u32(sK + buf * TILE_B), vb = smem_u32(sV + buf * TILE_B);
    // warp-level causal skip: nothing in this tile is visible to this warp's rows
    const bool warp_live = (n0 <= ctx + wq_max) && (warp * 16 < rows_used);
    if (warp_live) {
      // ---- S = Q K^T : 16 x 64 ---------------------------------------------------
      float sacc[BN / 8][4];
#pragma unroll
      for (int i = 0; i < BN / 8; ++i) sacc[i][0] = sacc[i][1] = sacc[i][2] = sacc[i][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ++ks) {
        if constexpr (!kQInRegs) {
          const int r = warp * 16 + (lane & 7) + 8 * ((lane >> 3) & 1);
          ldmatrix_x4(qf[0], smem_u32(sQ + swz_off<ROWB>(r, 2 * ks + (lane >> 4))));
        }
        const uint32_t (&qa)[4] = qf[kQInRegs ? ks : 0];
#pragma unroll
        for (int np = 0; np < BN / 16; ++np) {  // pairs of n8 tiles
          uint32_t r4[4];
          // matrices: (tok 0-7, c) (tok 0-7, c+1) (tok 8-15, c) (tok 8-15, c+1)
          const int r = np * 16 + (lane & 7) + 8 * (lane >> 4);
          const int c = 2 * ks + ((lane >> 3) & 1);
          ldmatrix_x4(r4, kb + swz_off<ROWB>(r, c));
          mma_16816<MMA_BF16>(sacc[2 * np], qa, r4[0], r4[1]);
          mma_16816<MMA_BF16>(sacc[2 * np + 1], qa, r4[2], r4[3]);
        }
      }
      // ---- scale / softcap / bias / mask -----------------------------------------
      const bool full = (n0 + BN <= ctx + wq_min + 1) && (n0 + BN <= kv_limit) &&
                        (p.sliding_window <= 0 || (ctx + wq_max - n0) < p.sliding_window) && (warp * 16 + 16 <= rows_used) &&
                        (qpos0 + (warp * 16 + 15) / group < q_len);
      float mx[2] = {-INFINITY, -INFINITY};
      if constexpr (kPlain) {
        // max over the raw dot products (scale > 0); only boundary tiles run the masking code
        if (!full) {
          // visible keys of row ri: [lo, hi)
          int lo[2], hi[2];
#pragma unroll
          for (int ri = 0; ri < 2; ++ri) {
            hi[ri] = row_ok[ri] ? min(ctx + row_qpos[ri] + 1, seq_len) : 0;
            lo[ri] = p.sliding_window > 0 ? ctx + row_qpos[ri] - p.sliding_window + 1 : 0;
          }
#pragma unroll
          for (int nt = 0; nt < BN / 8; ++nt) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int ri = i >> 1;
              const int pos = n0 + nt * 8 + 2 * t4 + (i & 1);
              if (pos >= hi[ri] || pos < lo[ri]) sacc[nt][i] = -INFINITY;
            }
          }
        }
#pragma unroll
        for (int nt = 0; nt < BN / 8; ++nt) {
          mx[0] = fmaxf(mx[0], fmaxf(sacc[nt][0], sacc[nt][1]));
          mx[1] = fmaxf(mx[1], fmaxf(sacc[nt][2], sacc[nt][3]));
        }
        mx[0] *= scale_log2, mx[1] *= scale_log2;
      } else {
        // general path: scale, softcap (before masking, reference :467-468), ALiBi, then the mask
        const float cap_log2 = p.softcap * kLog2e, pre = qk_scale * inv_cap;
#pragma unroll
        for (int nt = 0; nt < BN / 8; ++nt) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float kq = pth ? sKs[buf * BN + nt * 8 + 2 * t4 + (i & 1)] : 1.f;  // upstream: dot * (scale * k_scale[key])
            float x = has_cap ? cap_log2 * tanh_approx(sacc[nt][i] * (pre * kq)) : sacc[nt][i] * (scale_log2 * kq);
            if (p.alibi_slopes) x = fmaf(row_slope[i >> 1] * kLog2e, (float)(n0 + nt * 8 + 2 * t4 + (i & 1) - ctx), x);
            if (p.qq_bias) {
              // bias between this step's own query tokens (upstream qq_bias, triton_unified_attention.py:301-307 of
              // the vLLM copy): the keys at positions ctx .. ctx + stride - 1 are the step's tokens 0 .. stride - 1
              const int rel = n0 + nt * 8 + 2 * t4 + (i & 1) - ctx;
              if (rel >= 0 && rel < (int)p.qq_bias_stride0 && row_ok[i >> 1])
                x += p.qq_bias[(int64_t)row_qpos[i >> 1] * p.qq_bias_stride0 + rel] * kLog2e;
            }
            sacc[nt][i] = x;
          }
        }
        if (!full) {
          int lo[2], hi[2];
#pragma unroll
          for (int ri = 0; ri < 2; ++ri) {
            hi[ri] = row_ok[ri] ? min(ctx + row_qpos[ri] + 1, seq_len) : 0;
            lo[ri] = p.sliding_window > 0 ? ctx + row_qpos[ri] - p.sliding_window + 1 : 0;
          }
#pragma unroll
          for (int nt = 0; nt < BN / 8; ++nt) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int pos = n0 + nt * 8 + 2 * t4 + (i & 1);
              if (pos >= hi[i >> 1] || pos < lo[i >> 1]) sacc[nt][i] = -INFINITY;
            }
          }
        }
#pragma unroll
        for (int nt = 0; nt < BN / 8; ++nt) {
          mx[0] = fmaxf(mx[0], fmaxf(sacc[nt][0], sacc[nt][1]));
          mx[1] = fmaxf(mx[1], fmaxf(sacc[nt][2], sacc[nt][3]));
        }
      }
      float alpha[2], mu[2];
#pragma unroll
      for (int ri = 0; ri < 2; ++ri) {
        mx[ri] = fmaxf(mx[ri], __shfl_xor_sync(0xffffffffu, mx[ri], 1));
        mx[ri] = fmaxf(mx[ri], __shfl_xor_sync(0xffffffffu, mx[ri], 2));
        const float mn = fmaxf(m_run[ri], mx[ri]);
        mu[ri] = (mn == -INFINITY) ? 0.f : mn;
        alpha[ri] = ex2f(m_run[ri] - mu[ri]);
        m_run[ri] = mn;
      }
      float ls[2] = {0.f, 0.f};
      uint32_t pa[BN / 16][4];
#pragma unroll
      for (int nt = 0; nt < BN / 8; ++nt) {
        float p0, p1, p2, p3;
        if constexpr (kPlain) {  // exp2(s * scale_log2 - m): the scale rides in the FFMA
          p0 = ex2f(fmaf(sacc[nt][0], scale_log2, -mu[0])), p1 = ex2f(fmaf(sacc[nt][1], scale_log2, -mu[0]));
          p2 = ex2f(fmaf(sacc[nt][2], scale_log2, -mu[1])), p3 = ex2f(fmaf(sacc[nt][3], scale_log2, -mu[1]));
        } else {
          p0 = ex2f(sacc[nt][0] - mu[0]), p1 = ex2f(sacc[nt][1] - mu[0]);
          p2 = ex2f(sacc[nt][2] - mu[1]), p3 = ex2f(sacc[nt][3] - mu[1]);
        }
        ls[0] += p0 + p1, ls[1] += p2 + p3;
        if constexpr (!kPlain) {
          if (pth) {  // upstream: (P * v_scale[key]).to(V.dtype); the exp-sum takes P itself
            const float v0 = sVs[buf * BN + nt * 8 + 2 * t4], v1 = sVs[buf * BN + nt * 8 + 2 * t4 + 1];
            p0 *= v0, p1 *= v1, p2 *= v0, p3 *= v1;
          }
        }
        // C fragments of n-tiles (2j, 2j+1) are the A fragment of k16 step j
        pa[nt >> 1][(nt & 1) * 2 + 0] = pack2<MMA_BF16>(p0, p1);
        pa[nt >> 1][(nt & 1) * 2 + 1] = pack2<MMA_BF16>(p2, p3);
      }
      l_run[0] = l_run[0] * alpha[0] + ls[0];
      l_run[1] = l_run[1] * alpha[1] + ls[1];
      // the running output only needs rescaling when some row of the warp saw a new maximum
      if (__any_sync(0xffffffffu, alpha[0] != 1.f || alpha[1] != 1.f)) {
#pragma unroll
        for (int dt = 0; dt < DT; ++dt) {
          oacc[dt][0] *= alpha[0], oacc[dt][1] *= alpha[0];
          oacc[dt][2] *= alpha[1], oacc[dt][3] *= alpha[1];
        }
      }
      // ---- O += P V ----------------------------------------------------------------
#pragma unroll
      for (int kt = 0; kt < BN / 16; ++kt) {
#pragma unroll
        for (int dp = 0; dp < DT / 2; ++dp) {  // pairs of d n8-tiles
          uint32_t r4[4];
          // trans matrices: (tok 0-7, dchunk c) (tok 8-15, c) (tok 0-7, c+1) (tok 8-15, c+1)
          const int r = kt * 16 + (lane & 7) + 8 * ((lane >> 3) & 1);
          const int c = 2 * dp + (lane >> 4);
          ldmatrix_x4_trans(r4, vb + swz_off<ROWB>(r, c));
          mma_16816<MMA_BF16>(oacc[2 * dp], pa[kt], r4[0], r4[1]);
          mma_16816<MMA_BF16>(oacc[2 * dp + 1], pa[kt], r4[2], r4[3]);
        }
      }
    }
    __syncthreads();
  }
  cp_async_wait<0>();

  // ---- epilogue: acc / L --------------------------------------------------------
#pragma unroll
  for (int ri = 0; ri < 2; ++ri) {
    float l = l_run[ri];
    l += __shfl_xor_sync(0xffffffffu, l, 1);
    l += __shfl_xor_sync(0xffffffffu, l, 2);
    if (!row_ok[ri]) continue;
    const int r = warp * 16 + gid + 8 * ri;
    const int hq = kvh * group + r % group;
    float f = 1.f;
    const bool seen = m_run[ri] != -INFINITY;  // some key was visible (a sink alone does not make an LSE)
    if (p.sinks) f = sink_rescale<true>(p.sinks[hq], m_run[ri], l);
    const float inv = (l > 0.f) ? f * v_scale / l : 0.f;
    if (p.lse && t4 == 0) p.lse[(int64_t)(q_start + row_qpos[ri]) * p.lse_stride0 + hq] = seen ? row_lse<true>(m_run[ri], l) : -INFINITY;
    const int64_t oel = (int64_t)(q_start + row_qpos[ri]) * p.out_stride0 + (int64_t)hq * p.out_stride1;
    if (p.out_scale) {  // fused e4m3 output quantisation (b200attn.h: out_scale)
      const float oq = inv * out_quant_factor(p.out_scale);
      char* orow = static_cast<char*>(p.out) + oel;
#pragma unroll
      for (int dt = 0; dt < DT; ++dt)
        *reinterpret_cast<uint16_t*>(orow + dt * 8 + 2 * t4) = (uint16_t)fp8x2_out(oacc[dt][2 * ri] * oq, oacc[dt][2 * ri + 1] * oq);
      continue;
    }
    char* orow = static_cast<char*>(p.out) + oel * 2;
#pragma unroll
    for (int dt = 0; dt < DT; ++dt) {
      const float a = oacc[dt][2 * ri] * inv, c = oacc[dt][2 * ri + 1] * inv;
      *reinterpret_cast<uint32_t*>(orow + (dt * 8 + 2 * t4) * 2) = pack2<Q_BF16>(a, c);
    }
  }
}

template <int D, bool MMA_BF16, bool Q_BF16, bool KV_FP8, bool kPlain>
int launch_variant(const b200attn_attention_params& a, cudaStream_t stream, QLenRange range, const int32_t* skip_staged) {
  const int group = a.num_q_heads / a.num_kv_heads;
  const int block_q = BM / group;
  const int smem = BM * D * 2 + 4 * BN * D * 2 + (KV_FP8 && !kPlain ? 4 * BN * 4 : 0);  // + the tiles' scale rows
  auto kern = attn_prefill_kernel<D, MMA_BF16, Q_BF16, KV_FP8, kPlain>;
  int bs_shift = -1;
  if ((a.block_size & (a.block_size - 1)) == 0)
    for (bs_shift = 0; (1 << bs_shift) < a.block_size; ++bs_shift) {}
  if (int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(kern), smem, "prefill")) return rc;
  // host-int upper bounds on the q-blocks: the reference's (T // BLOCK_Q + B, :886-889), or, when the launch
  // only takes queries up to max_q_len, B * ceil(max_q_len / BLOCK_Q) with a direct block -> sequence map
  long long nblocks = (long long)a.num_tokens / block_q + a.num_seqs;
  int qb_per_seq = 0;
  // (max_seqlen_q describes the batch in hand; a captured launch is replayed with others, so it only trusts the range)
  const bool captured = launch_is_captured();
  const int longest = (captured || range.max_q_len < a.max_seqlen_q) ? range.max_q_len : a.max_seqlen_q;
  const long long per_seq = ((long long)longest + block_q - 1) / block_q;
  if (per_seq >= 1 && per_seq * a.num_seqs < nblocks) qb_per_seq = (int)per_seq, nblocks = per_seq * a.num_seqs;
  const dim3 grid((unsigned)nblocks, (unsigned)a.num_kv_heads);
  kern<<<grid, NWARPS * 32, smem, stream>>>(a, block_q, group, range.min_q_len, range.max_q_len, qb_per_seq, skip_staged, bs_shift);
  note_launch("attn_prefill_mma");
  return check_launch("attn_prefill");
}

template <int D, bool kPlain>
int dispatch_types(const b200attn_attention_params& a, cudaStream_t stream, QLenRange range, const int32_t* skip) {
  const bool qbf = a.q_dtype == B200ATTN_BF16;
  if (dtype_bytes(a.kv_dtype) == 1)  // e4m3, or int8 with per-(token, head) scales
    return qbf ? launch_variant<D, false, true, true, kPlain>(a, stream, range, skip)
               : launch_variant<D, false, false, true, kPlain>(a, stream, range, skip);
  return qbf ? launch_variant<D, true, true, false, kPlain>(a, stream, range, skip)
             : launch_variant<D, false, false, false, kPlain>(a, stream, range, skip);
}

}  // namespace

bool prefill_supported(const b200attn_attention_params& p, const char** why) {
  auto no = [&](const char* w) { if (why) *why = w; return false; };
  if (p.q_dtype != B200ATTN_BF16 && p.q_dtype != B200ATTN_F16) return no("query dtype must be bf16/f16");
  const bool kv8 = dtype_bytes(p.kv_dtype) == 1;
  if (p.kv_dtype != p.q_dtype && !kv8) return no("kv dtype");
  if (p.head_size != 64 && p.head_size != 128 && p.head_size != 256) return no("head_size not in {64,128,256}");
  if (p.num_kv_heads <= 0 || p.num_q_heads % p.num_kv_heads) return no("heads");
  if (p.num_q_heads / p.num_kv_heads > BM) return no("GQA group > 64");
  const int esz = kv8 ? 1 : 2;
  auto m16 = [&](int64_t s) { return (s * esz) % 16 == 0; };
  if (!m16(p.kc_block_stride) || !m16(p.kc_slot_stride) || !m16(p.kc_head_stride) || !m16(p.vc_block_stride) ||
      !m16(p.vc_slot_stride) || !m16(p.vc_head_stride))
    return no("cache strides not 16-byte multiples");
  auto fits = [&](int64_t s) { return s >= 0 && s * esz < (int64_t(1) << 31); };  // the loader multiplies in 32 x 32 -> 64 bits
  if (!fits(p.kc_block_stride) || !fits(p.kc_slot_stride) || !fits(p.vc_block_stride) || !fits(p.vc_slot_stride))
    return no("cache block/slot stride >= 2 GiB");
  if ((reinterpret_cast<uintptr_t>(p.k_cache) & 15) || (reinterpret_cast<uintptr_t>(p.v_cache) & 15)) return no("cache base alignment");
  if ((reinterpret_cast<uintptr_t>(p.q) & 15) || (p.q_stride0 % 8) || (p.q_stride1 % 8)) return no("query alignment");
  if ((reinterpret_cast<uintptr_t>(p.out) & 3) || (p.out_stride0 % 2) || (p.out_stride1 % 2)) return no("output alignment");
  return true;
}

int launch_attn_prefill(const b200attn_attention_params& a, cudaStream_t stream, QLenRange range,
                        const int32_t* skip_staged) {
  const char* why = nullptr;
  if (!prefill_supported(a, &why)) {
    set_error("prefill: unsupported configuration (%s)", why);
    return B200ATTN_E_UNSUPPORTED;
  }
  // the scale of an FP8 cache is a device scalar: it is taken as positive (a quantisation scale always is)
  const bool plain = a.softcap == 0.f && !a.alibi_slopes && !a.qq_bias && a.softmax_scale > 0.f && !per_token_head(a);
  switch (a.head_size) {
    case 64: return plain ? dispatch_types<64, true>(a, stream, range, skip_staged) : dispatch_types<64, false>(a, stream, range, skip_staged);
    case 128: return plain ? dispatch_types<128, true>(a, stream, range, skip_staged) : dispatch_types<128, false>(a, stream, range, skip_staged);
    case 256: return plain ? dispatch_types<256, true>(a, stream, range, skip_staged) : dispatch_types<256, false>(a, stream, range, skip_staged);
  }
  return B200ATTN_E_UNSUPPORTED;
}

}  // namespace b200attn

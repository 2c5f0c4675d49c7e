// K0: generic paged attention (SIMT).  One warp per (query token, query head);
// any head size, any GQA group, fp32/bf16/f16 queries, same-dtype or e4m3 KV.
//
// This is the correctness net for shapes the tensor-core kernels do not take
// (head sizes that are not 64/128/256, fp32 queries, very large GQA groups).
// Its arithmetic follows the reference kernel step by step
// (reference ibm-triton-lib/ibm_triton_lib/kernels/triton_unified_attention.py:
//  400-511): scale after the dot (:465), softcap before masking (:467), causal +
// context mask (:460), sliding window (:474-479), ALiBi (:481-482), online
// softmax with the all-masked guard (:486-492), P rounded to V's dtype before
// PV (:508), fp8 K/V dequantised as round_to_q_dtype(float(x) * scale) (:438).
// Per-(token, head) scales (b200attn.h k_scale_cache; upstream KVQuantMode 2 / 3): K and V are used as stored, the
// key's scale multiplies the softmax scale and the value's multiplies P before its rounding.
#include "common.cuh"

namespace b200attn {
namespace {

constexpr int kMaxDPerLane = 8;  // head_size <= 256

template <typename QT, typename KVT>
__global__ void __launch_bounds__(128) attn_generic_kernel(b200attn_attention_params p) {
  extern __shared__ float q_smem[];  // [4][head_size]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t wid = (int64_t)blockIdx.x * 4 + warp;
  const int t = (int)(wid / p.num_q_heads), h = (int)(wid % p.num_q_heads);
  if (t >= p.num_tokens) return;
  const int D = p.head_size;
  const int g = p.num_q_heads / p.num_kv_heads;
  const int kvh = h / g;

  // sequence of this token: largest b with cu[b] <= t
  int lo_b = 0, hi_b = p.num_seqs;
  while (lo_b < hi_b) {
    int mid = (lo_b + hi_b) >> 1;
    if (p.cu_seqlens_q[mid] <= t) lo_b = mid + 1; else hi_b = mid;
  }
  const int b = lo_b - 1;
  if (b < 0 || t >= p.cu_seqlens_q[p.num_seqs]) return;
  const int q_start = p.cu_seqlens_q[b];
  const int q_len = p.cu_seqlens_q[b + 1] - q_start;
  const int seq_len = p.seq_lens[b];
  const int ctx = seq_len - q_len;
  const int qpos = t - q_start;
  int limit = ctx + qpos + 1;  // keys [0, limit) are causally visible
  if (limit > seq_len) limit = seq_len;
  int first = 0;
  if (p.sliding_window > 0) first = max(0, ctx + qpos - p.sliding_window + 1);

  const QT* qrow = static_cast<const QT*>(p.q) + (int64_t)t * p.q_stride0 + (int64_t)h * p.q_stride1;
  float* qs = q_smem + warp * D;
  for (int d = lane; d < D; d += 32) qs[d] = to_float<QT>(qrow[d]);
  __syncwarp();

  const float k_scale = p.k_scale ? *p.k_scale : 1.0f;
  const float v_scale = p.v_scale ? *p.v_scale : 1.0f;
  const float slope = p.alibi_slopes ? p.alibi_slopes[h] : 0.0f;
  const float inv_cap = p.softcap > 0.f ? 1.0f / p.softcap : 0.f;
  const KVT* kc = static_cast<const KVT*>(p.k_cache);
  const KVT* vc = static_cast<const KVT*>(p.v_cache);
  const int32_t* bt = p.block_tables + (int64_t)b * p.bt_stride0;
  constexpr bool kFp8 = sizeof(KVT) == 1;
  const bool pth = kFp8 && p.k_scale_cache != nullptr;

  float acc[kMaxDPerLane];
#pragma unroll
  for (int i = 0; i < kMaxDPerLane; ++i) acc[i] = 0.f;
  float m = -INFINITY, l = 0.f;

  for (int base = (first / 32) * 32; base < limit; base += 32) {
    const int pos = base + lane;
    const bool valid = pos < limit && pos >= first;
    float s = -INFINITY;
    int64_t voff = 0;
    float vs_tok = 1.f;
    if (valid) {
      const int64_t page = bt[pos / p.block_size];
      const int slot = pos % p.block_size;
      float ks_tok = 1.f;
      if (pth) {
        ks_tok = p.k_scale_cache[page * p.ks_block_stride + (int64_t)slot * p.ks_slot_stride + (int64_t)kvh * p.ks_head_stride];
        vs_tok = p.v_scale_cache[page * p.vs_block_stride + (int64_t)slot * p.vs_slot_stride + (int64_t)kvh * p.vs_head_stride];
      }
      const KVT* krow = kc + page * p.kc_block_stride + (int64_t)slot * p.kc_slot_stride + (int64_t)kvh * p.kc_head_stride;
      voff = page * p.vc_block_stride + (int64_t)slot * p.vc_slot_stride + (int64_t)kvh * p.vc_head_stride;
      float dot = 0.f;
      for (int d = 0; d < D; ++d) {
        float kx = to_float<KVT>(krow[d]);
        if constexpr (kFp8) {
          if (!pth) kx = to_float<QT>(from_float<QT>(kx * k_scale));
        }
        dot = fmaf(qs[d], kx, dot);
      }
      s = pth ? dot * (p.softmax_scale * ks_tok) : dot * p.softmax_scale;
      if (p.softcap > 0.f) s = softcap_fn(s, p.softcap, inv_cap);
      if (p.alibi_slopes) s += slope * (float)(pos - ctx);
      if (p.qq_bias) {  // bias between this step's query tokens (upstream load_qq_bias_tile)
        const int rel = pos - ctx;
        if (rel >= 0 && rel < p.qq_bias_stride0) s += p.qq_bias[(int64_t)qpos * p.qq_bias_stride0 + rel];
      }
    }
    float m_new = fmaxf(m, warp_max(s));
    const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
    float pr = valid ? __expf(s - m_use) : 0.f;
    const float alpha = __expf(m - m_use);  // m == -inf -> 0
    l = l * alpha + warp_sum(pr);
    m = m_new;
    pr = to_float<QT>(from_float<QT>(pr * vs_tok));  // P (times the value's own scale) is cast to V's (= Q's) dtype before PV
#pragma unroll
    for (int i = 0; i < kMaxDPerLane; ++i) acc[i] *= alpha;
    const unsigned live = __ballot_sync(0xffffffffu, valid);
    for (int j = 0; j < 32; ++j) {
      if (!((live >> j) & 1u)) continue;
      const float pj = __shfl_sync(0xffffffffu, pr, j);
      const int64_t vo = __shfl_sync(0xffffffffu, voff, j);
      const KVT* vrow = vc + vo;
#pragma unroll
      for (int i = 0; i < kMaxDPerLane; ++i) {
        const int d = lane + 32 * i;
        if (d < D) {
          float vx = to_float<KVT>(vrow[d]);
          if constexpr (kFp8) {
            if (!pth) vx = to_float<QT>(from_float<QT>(vx * v_scale));
          }
          acc[i] = fmaf(pj, vx, acc[i]);
        }
      }
    }
  }
  const int64_t oel = (int64_t)t * p.out_stride0 + (int64_t)h * p.out_stride1;
  // L == 0 happens only when no key is visible; the reference's reduce returns 0 there (:828)
  float f = 1.f;
  const bool seen = m != -INFINITY;  // some key was visible (a sink alone does not make an LSE)
  if (p.sinks) f = sink_rescale<false>(p.sinks[h], m, l);
  const float inv_l = l > 0.f ? f / l : 0.f;
  if (p.lse && lane == 0) p.lse[(int64_t)t * p.lse_stride0 + h] = seen ? row_lse<false>(m, l) : -INFINITY;
  if (p.out_scale) {  // fused e4m3 output quantisation (b200attn.h: out_scale)
    const float oq = out_quant_factor(p.out_scale);
    uint8_t* orow = static_cast<uint8_t*>(p.out) + oel;
#pragma unroll
    for (int i = 0; i < kMaxDPerLane; ++i) {
      const int d = lane + 32 * i;
      if (d < D) orow[d] = fp8_out(acc[i] * inv_l * oq);
    }
    return;
  }
  QT* orow = static_cast<QT*>(p.out) + oel;
#pragma unroll
  for (int i = 0; i < kMaxDPerLane; ++i) {
    const int d = lane + 32 * i;
    if (d < D) orow[d] = from_float<QT>(acc[i] * inv_l);
  }
}

template <typename QT>
int dispatch_kv(const b200attn_attention_params& p, cudaStream_t stream) {
  const int64_t warps = (int64_t)p.num_tokens * p.num_q_heads;
  const unsigned grid = (unsigned)((warps + 3) / 4);
  const size_t smem = 4 * p.head_size * sizeof(float);
  if (p.kv_dtype == B200ATTN_FP8_E4M3) {
    attn_generic_kernel<QT, __nv_fp8_e4m3><<<grid, 128, smem, stream>>>(p);
  } else if (p.kv_dtype == B200ATTN_INT8) {
    attn_generic_kernel<QT, int8_t><<<grid, 128, smem, stream>>>(p);
  } else if (p.kv_dtype == p.q_dtype) {
    attn_generic_kernel<QT, QT><<<grid, 128, smem, stream>>>(p);
  } else {
    set_error("attention(generic): kv_dtype %d must equal q_dtype %d or be fp8_e4m3", p.kv_dtype, p.q_dtype);
    return B200ATTN_E_BAD_DTYPE;
  }
  note_launch("attn_generic");
  return check_launch("attn_generic");
}

}  // namespace

int launch_attn_generic(const b200attn_attention_params& p, cudaStream_t stream) {
  if (p.head_size > 32 * kMaxDPerLane) {
    set_error("attention(generic): head_size %d > 256", p.head_size);
    return B200ATTN_E_UNSUPPORTED;
  }
  switch (p.q_dtype) {
    case B200ATTN_BF16: return dispatch_kv<__nv_bfloat16>(p, stream);
    case B200ATTN_F16: return dispatch_kv<__half>(p, stream);
    case B200ATTN_F32: return dispatch_kv<float>(p, stream);
    default:
      set_error("attention(generic): unsupported q_dtype %d", p.q_dtype);
      return B200ATTN_E_BAD_DTYPE;
  }
}

}  // namespace b200attn

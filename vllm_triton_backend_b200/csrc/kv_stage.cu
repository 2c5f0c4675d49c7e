// K6: FP8 (or int8) KV pages -> 16-bit staging pages for the tcgen05 prefill kernel.
//
// tcgen05.mma kind::f16 takes 16-bit operands from shared memory, so an FP8-e4m3 cache cannot
// feed it directly.  For the sequences of a batch that the tcgen05 kernel takes (query length in
// a range), their pages are dequantised ONCE per call into a caller-owned staging pool laid out
// as a private paged cache ([pages][block_size][Hkv][D], K plane then V plane) and the prefill
// kernel reads that pool through `staged_base[b] + page` instead of the block table.  The value
// written is what the reference's kernel forms on the fly, round_to_q_dtype(float(fp8) * scale)
// (reference triton_unified_attention.py:434-455), so the staged run is an ordinary 16-bit
// attention with scales of 1.  Cost: (1 + 2) bytes per cached element moved once, against
// attention FLOPs that grow with query_len x context.
//
// Two launches, grids from host ints only (graph safe):
//   stage_plan    one warp: greedy first-fit in sequence order, staged_base[b] = first staging
//                 page of sequence b, or -1 (not this kernel's, or the pool is full - such
//                 sequences stay on the mma.sync kernel).
//   stage_dequant (page, sequence) CTAs: 16 fp8 bytes in, 32 bytes out per thread and step.
#include "common.cuh"

namespace b200attn {
namespace {

struct StageArgs {
  const uint8_t* kc;
  const uint8_t* vc;
  int64_t kc_block_stride, kc_slot_stride, kc_head_stride, vc_block_stride, vc_slot_stride, vc_head_stride;
  const int32_t* block_tables;
  int64_t bt_stride0;
  const int32_t* seq_lens;
  const int32_t* cu_seqlens_q;
  const float* k_scale;
  const float* v_scale;
  const float* k_scale_cache;  // per-(token, head) scales (b200attn.h), or nullptr
  const float* v_scale_cache;
  int64_t ks_block_stride, ks_slot_stride, ks_head_stride, vs_block_stride, vs_slot_stride, vs_head_stride;
  int kv_int8;
  int32_t* staged_base;
  void* k_stage;
  void* v_stage;
  int num_seqs, min_q_len, max_q_len, block_size, num_kv_heads, head_size, cap_pages;
};

__global__ void __launch_bounds__(32) stage_plan_kernel(const StageArgs a) {
  const int lane = threadIdx.x;
  long long run = 0;
  for (int b0 = 0; b0 < a.num_seqs; b0 += 32) {
    const int b = b0 + lane;
    int pages = 0;
    if (b < a.num_seqs) {
      const int ql = a.cu_seqlens_q[b + 1] - a.cu_seqlens_q[b];
      const int sl = a.seq_lens[b];
      if (ql >= a.min_q_len && ql <= a.max_q_len && sl > 0) pages = (sl + a.block_size - 1) / a.block_size;
    }
    int incl = pages;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (b < a.num_seqs) a.staged_base[b] = (pages > 0 && run + incl <= a.cap_pages) ? (int)(run + incl - pages) : -1;
    run += __shfl_sync(0xffffffffu, incl, 31);
  }
}

template <typename T> __device__ __forceinline__ uint32_t pack_t(float a, float b);
template <> __device__ __forceinline__ uint32_t pack_t<__nv_bfloat16>(float a, float b) { return pack2<true>(a, b); }
template <> __device__ __forceinline__ uint32_t pack_t<__half>(float a, float b) { return pack2<false>(a, b); }

template <typename T>
__global__ void __launch_bounds__(256) stage_dequant_kernel(const StageArgs a) {
  const int b = blockIdx.y;
  const int base = a.staged_base[b];
  if (base < 0) return;
  const int D = a.head_size, H = a.num_kv_heads, bs = a.block_size;
  const int d16 = D / 16, per_slot = H * d16, units = bs * per_slot;
  const float ks = a.k_scale ? *a.k_scale : 1.f, vs = a.v_scale ? *a.v_scale : 1.f;
  const int seq_len = a.seq_lens[b];
  // pages are strided over the grid's x dimension: the grid is an upper bound from host ints, the
  // length is device data (a captured graph replays with longer sequences than it was captured with)
  for (int p = blockIdx.x; p * bs < seq_len; p += gridDim.x) {
  const int64_t page = a.block_tables[(int64_t)b * a.bt_stride0 + p];
  const int64_t dst_page = (int64_t)(base + p) * bs * H * D;  // elements
  for (int u = threadIdx.x; u < 2 * units; u += blockDim.x) {
    const bool is_v = u >= units;
    const int uu = is_v ? u - units : u;
    const int s = uu / per_slot, r = uu - s * per_slot, h = r / d16, c = r - h * d16;
    const uint8_t* src = is_v ? a.vc + page * a.vc_block_stride + (int64_t)s * a.vc_slot_stride + (int64_t)h * a.vc_head_stride
                              : a.kc + page * a.kc_block_stride + (int64_t)s * a.kc_slot_stride + (int64_t)h * a.kc_head_stride;
    int4 raw = *reinterpret_cast<const int4*>(src + c * 16);
    if (p * bs + s >= seq_len) raw = make_int4(0, 0, 0, 0);  // slots past the end were never written: stage zeros, not whatever bytes (NaN patterns) lie there
    float sc = is_v ? vs : ks;
    if (a.k_scale_cache && p * bs + s < seq_len)  // the row's own scale; the staged value is round_to_q_dtype(float(x) * scale) as for a per-tensor scale
      sc = is_v ? a.v_scale_cache[page * a.vs_block_stride + (int64_t)s * a.vs_slot_stride + (int64_t)h * a.vs_head_stride]
                : a.k_scale_cache[page * a.ks_block_stride + (int64_t)s * a.ks_slot_stride + (int64_t)h * a.ks_head_stride];
    const uint32_t w[4] = {(uint32_t)raw.x, (uint32_t)raw.y, (uint32_t)raw.z, (uint32_t)raw.w};
    uint32_t o[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t lo_bits = a.kv_int8 ? s8x2_to_f16x2((uint16_t)(w[i] & 0xffffu)) : e4m3x2_to_f16x2((uint16_t)(w[i] & 0xffffu));  // exact
      const uint32_t hi_bits = a.kv_int8 ? s8x2_to_f16x2((uint16_t)(w[i] >> 16)) : e4m3x2_to_f16x2((uint16_t)(w[i] >> 16));
      const __half2 lo = *reinterpret_cast<const __half2*>(&lo_bits), hi = *reinterpret_cast<const __half2*>(&hi_bits);
      const float2 fl = __half22float2(lo), fh = __half22float2(hi);
      o[2 * i] = pack_t<T>(fl.x * sc, fl.y * sc);
      o[2 * i + 1] = pack_t<T>(fh.x * sc, fh.y * sc);
    }
    T* dst = static_cast<T*>(is_v ? a.v_stage : a.k_stage) + dst_page + ((int64_t)s * H + h) * D + c * 16;
    reinterpret_cast<int4*>(dst)[0] = make_int4(o[0], o[1], o[2], o[3]);
    reinterpret_cast<int4*>(dst)[1] = make_int4(o[4], o[5], o[6], o[7]);
  }
  }
}

}  // namespace

// Bytes of staging a pool of `pages` pages needs after the decode region of the workspace.
uint64_t stage_bytes_for_pages(int num_seqs, int64_t pages, int block_size, int num_kv_heads, int head_size) {
  const uint64_t index = (((uint64_t)num_seqs * 4 + 255) / 256) * 256;
  return index + 2ull * (uint64_t)pages * block_size * num_kv_heads * head_size * 2 + 256;
}

// Carves the staging pool out of the part of the caller's workspace that lies past the decode
// kernel's region.  Returns false when there is no room worth using.
bool plan_kv_stage(const b200attn_attention_params& p, KvStage* st) {
  if (!p.workspace) return false;
  const uint64_t dec = ((decode_workspace_bytes(p.num_seqs, p.num_q_heads, p.num_kv_heads, p.head_size) + 255) / 256) * 256;
  const uint64_t index = (((uint64_t)p.num_seqs * 4 + 255) / 256) * 256;
  if (p.workspace_bytes < dec + index + 512) return false;
  const uint64_t page_bytes = (uint64_t)p.block_size * p.num_kv_heads * p.head_size * 2;
  uint64_t cap = (p.workspace_bytes - dec - index - 256) / (2 * page_bytes);
  const uint64_t useful = (uint64_t)p.num_seqs * ((p.max_seqlen_k + p.block_size - 1) / p.block_size);
  if (cap > useful && !launch_is_captured()) cap = useful;  // never needs more than every sequence at full length
  if (cap < 8 || cap > 0x7fffffffull) return false;
  char* w = static_cast<char*>(p.workspace) + dec;
  st->staged_base = reinterpret_cast<int32_t*>(w);
  st->k_stage = w + index;
  st->v_stage = w + index + cap * page_bytes;
  st->cap_pages = (int)cap;
  return true;
}

// The staged cache as attention parameters: a dense NHD 16-bit cache of cap_pages pages.
b200attn_attention_params staged_params(const b200attn_attention_params& p, const KvStage& st) {
  b200attn_attention_params s = p;
  s.k_cache = st.k_stage, s.v_cache = st.v_stage;
  s.kv_dtype = p.q_dtype;
  s.k_scale = nullptr, s.v_scale = nullptr;
  s.k_scale_cache = nullptr, s.v_scale_cache = nullptr;
  s.num_blocks = st.cap_pages;
  s.kc_head_stride = s.vc_head_stride = p.head_size;
  s.kc_slot_stride = s.vc_slot_stride = (int64_t)p.num_kv_heads * p.head_size;
  s.kc_block_stride = s.vc_block_stride = (int64_t)p.block_size * p.num_kv_heads * p.head_size;
  return s;
}

int launch_kv_stage(const b200attn_attention_params& p, const KvStage& st, QLenRange range, cudaStream_t stream) {
  if (dtype_bytes(p.kv_dtype) != 1 || (p.q_dtype != B200ATTN_BF16 && p.q_dtype != B200ATTN_F16) || p.head_size % 16) {
    set_error("kv_stage: needs an fp8 cache, bf16/f16 queries and head_size %% 16 == 0");
    return B200ATTN_E_UNSUPPORTED;
  }
  StageArgs a;
  a.kc = static_cast<const uint8_t*>(p.k_cache), a.vc = static_cast<const uint8_t*>(p.v_cache);
  a.kc_block_stride = p.kc_block_stride, a.kc_slot_stride = p.kc_slot_stride, a.kc_head_stride = p.kc_head_stride;
  a.vc_block_stride = p.vc_block_stride, a.vc_slot_stride = p.vc_slot_stride, a.vc_head_stride = p.vc_head_stride;
  a.block_tables = p.block_tables, a.bt_stride0 = p.bt_stride0, a.seq_lens = p.seq_lens, a.cu_seqlens_q = p.cu_seqlens_q;
  a.k_scale = p.k_scale, a.v_scale = p.v_scale;
  a.k_scale_cache = p.k_scale_cache, a.v_scale_cache = p.v_scale_cache;
  a.ks_block_stride = p.ks_block_stride, a.ks_slot_stride = p.ks_slot_stride, a.ks_head_stride = p.ks_head_stride;
  a.vs_block_stride = p.vs_block_stride, a.vs_slot_stride = p.vs_slot_stride, a.vs_head_stride = p.vs_head_stride;
  a.kv_int8 = p.kv_dtype == B200ATTN_INT8 ? 1 : 0;
  a.staged_base = st.staged_base, a.k_stage = st.k_stage, a.v_stage = st.v_stage;
  a.num_seqs = p.num_seqs, a.min_q_len = range.min_q_len, a.max_q_len = range.max_q_len;
  a.block_size = p.block_size, a.num_kv_heads = p.num_kv_heads, a.head_size = p.head_size, a.cap_pages = st.cap_pages;
  stage_plan_kernel<<<1, 32, 0, stream>>>(a);
  note_launch("kv_stage_plan");
  int rc = check_launch("kv_stage_plan");
  if (rc) return rc;
  int max_pages = (p.max_seqlen_k + p.block_size - 1) / p.block_size;
  if (max_pages > 1024 || launch_is_captured()) max_pages = 1024;  // the kernel strides over a sequence's pages
  const dim3 grid((unsigned)(max_pages > 0 ? max_pages : 1), (unsigned)p.num_seqs);
  if (p.q_dtype == B200ATTN_BF16) stage_dequant_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(a);
  else stage_dequant_kernel<__half><<<grid, 256, 0, stream>>>(a);
  note_launch("kv_stage_dequant");
  return check_launch("kv_stage_dequant");
}

}  // namespace b200attn

// K5: rotary embedding of Q and K in place + KV-cache scatter write, one launch.
//
// Replaces the pair of launches vLLM issues before attention,
//   rotary_embedding(positions, q, k, head_size, cos_sin_cache, is_neox)
//       (reference vllm/csrc/pos_encoding_kernels.cu: apply_token_rotary_embedding)
//   reshape_and_cache_flash(k, v, key_cache, value_cache, slot_mapping, ...)
//       (reference vllm/csrc/cache_kernels.cu:286-358)
// behind AttentionImpl.do_rope_and_kv_cache_update (vLLM 0.22
// vllm/v1/attention/backend.py:823-841).  K is read once, rotated, written back and
// written to the cache from registers; V goes straight to the cache.
//
// Arithmetic: x' = R(x*cos - y*sin), y' = R(y*cos + x*sin) evaluated in fp32 and
// rounded once to the tensor's 16-bit type.  That is what the compiled
// `_C.rotary_embedding` of the vLLM this plugs into (0.22) produces - probed on the
// GPU box with scripts/probe_rope_rounding.py: 0 mismatches for this model, 32 % for
// per-operation rounding in `scalar_t` - so q, k and the cache match the unfused
// pair bit for bit (tests/test_gpu_rope_cache.py).
#include "common.cuh"

namespace b200attn {
namespace {

struct RopeArgs {
  void* q;
  void* k;
  const void* v;
  const int64_t* positions;
  const void* cos_sin;  // [max_pos, rot_dim]: cos | sin halves, same 16-bit type as q
  void* kc;
  void* vc;
  const int64_t* slot_mapping;
  const float* k_scale;
  const float* v_scale;
  int64_t q_stride0, q_stride1, k_stride0, k_stride1, v_stride0, v_stride1;
  int64_t kc_block_stride, kc_slot_stride, kc_head_stride, vc_block_stride, vc_slot_stride, vc_head_stride;
  int64_t rope_tokens, cache_tokens;
  int num_q_heads, num_kv_heads, head_size, rot_dim, block_size;
};

template <typename T> __device__ __forceinline__ T rnd(float x);
template <> __device__ __forceinline__ __nv_bfloat16 rnd<__nv_bfloat16>(float x) { return __float2bfloat16_rn(x); }
template <> __device__ __forceinline__ __half rnd<__half>(float x) { return __float2half_rn(x); }

// one rotation in fp32, rounded once (products of two 16-bit floats are exact in fp32, so
// whether the compiler contracts them into FMAs does not change the result)
template <typename T> __device__ __forceinline__ void rotate(T& x, T& y, T c, T s) {
  const float xf = to_float<T>(x), yf = to_float<T>(y), cf = to_float<T>(c), sf = to_float<T>(s);
  x = rnd<T>(xf * cf - yf * sf);
  y = rnd<T>(yf * cf + xf * sf);
}

template <typename T> struct Vec8 {
  T e[8];
};
template <typename T> __device__ __forceinline__ Vec8<T> ld8(const T* p) {
  Vec8<T> v;
  *reinterpret_cast<int4*>(v.e) = *reinterpret_cast<const int4*>(p);
  return v;
}
template <typename T> __device__ __forceinline__ void st8(T* p, const Vec8<T>& v) {
  *reinterpret_cast<int4*>(p) = *reinterpret_cast<const int4*>(v.e);
}

// 8 elements -> the cache: 16-byte copy, or 8 e4m3 bytes of float(x) / scale (true division,
// quant_utils.cuh:479-502, as K1 does)
template <typename T, bool kQuant>
__device__ __forceinline__ void to_cache(void* cache_row, int d, const Vec8<T>& v, float scale) {
  if constexpr (!kQuant) {
    st8<T>(static_cast<T*>(cache_row) + d, v);
  } else {
    uint32_t w[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const float2 a = make_float2(__fdiv_rn(to_float<T>(v.e[4 * i]), scale), __fdiv_rn(to_float<T>(v.e[4 * i + 1]), scale));
      const float2 b = make_float2(__fdiv_rn(to_float<T>(v.e[4 * i + 2]), scale), __fdiv_rn(to_float<T>(v.e[4 * i + 3]), scale));
      w[i] = (uint32_t)__nv_cvt_float2_to_fp8x2(a, __NV_SATFINITE, __NV_E4M3) |
             ((uint32_t)__nv_cvt_float2_to_fp8x2(b, __NV_SATFINITE, __NV_E4M3) << 16);
    }
    *reinterpret_cast<uint2*>(static_cast<uint8_t*>(cache_row) + d) = make_uint2(w[0], w[1]);
  }
}

template <typename T, bool kNeox, bool kQuant>
__global__ void __launch_bounds__(256) rope_cache_kernel(const RopeArgs a) {
  pdl_launch_dependents();  // the attention launch behind us may start its prologue now (common.cuh)
  pdl_wait();               // ... and this grid may itself have been scheduled early: wait for the kernel in front
  const int64_t t = blockIdx.x;
  const bool do_rope = t < a.rope_tokens;
  int64_t slot = -1;
  if (t < a.cache_tokens) slot = a.slot_mapping[t];
  const bool do_cache = slot >= 0;
  const int D = a.head_size, rot = a.rot_dim, half = rot / 2;
  const int Hq = a.num_q_heads, Hkv = a.num_kv_heads;
  const int64_t esz_c = kQuant ? 1 : (int64_t)sizeof(T);
  char* kc_tok = nullptr;
  char* vc_tok = nullptr;
  float ks = 1.f, vs = 1.f;
  if (do_cache) {
    const int64_t blk = slot / a.block_size, off = slot % a.block_size;
    kc_tok = static_cast<char*>(a.kc) + (blk * a.kc_block_stride + off * a.kc_slot_stride) * esz_c;
    vc_tok = static_cast<char*>(a.vc) + (blk * a.vc_block_stride + off * a.vc_slot_stride) * esz_c;
    if constexpr (kQuant) ks = *a.k_scale, vs = *a.v_scale;
  }
  const T* cs = nullptr;
  if (do_rope) cs = static_cast<const T*>(a.cos_sin) + a.positions[t] * (int64_t)rot;

  // ---- A: rotate q and k heads in place; k also goes to the cache -----------------
  const int upr = (kNeox ? half : rot) / 8;  // 16-byte units per head in the rotated part
  for (int u = threadIdx.x; u < (Hq + Hkv) * upr; u += blockDim.x) {
    const int h = u / upr, g = u - h * upr;
    const bool is_k = h >= Hq;
    if (!do_rope && !(is_k && do_cache)) continue;
    T* row = is_k ? static_cast<T*>(a.k) + t * a.k_stride0 + (int64_t)(h - Hq) * a.k_stride1
                  : static_cast<T*>(a.q) + t * a.q_stride0 + (int64_t)h * a.q_stride1;
    void* crow = is_k && do_cache ? kc_tok + (int64_t)(h - Hq) * a.kc_head_stride * esz_c : nullptr;
    if constexpr (kNeox) {
      Vec8<T> x = ld8<T>(row + 8 * g), y = ld8<T>(row + half + 8 * g);
      if (do_rope) {
        const Vec8<T> c = ld8<T>(cs + 8 * g), s = ld8<T>(cs + half + 8 * g);
#pragma unroll
        for (int i = 0; i < 8; ++i) rotate<T>(x.e[i], y.e[i], c.e[i], s.e[i]);
        st8<T>(row + 8 * g, x);
        st8<T>(row + half + 8 * g, y);
      }
      if (crow) {
        to_cache<T, kQuant>(crow, 8 * g, x, ks);
        to_cache<T, kQuant>(crow, half + 8 * g, y, ks);
      }
    } else {
      Vec8<T> x = ld8<T>(row + 8 * g);  // pairs (0,1) (2,3) (4,5) (6,7) <-> cos/sin index 4g..4g+3
      if (do_rope) {
        T c[4], s[4];
        *reinterpret_cast<uint2*>(c) = *reinterpret_cast<const uint2*>(cs + 4 * g);
        *reinterpret_cast<uint2*>(s) = *reinterpret_cast<const uint2*>(cs + half + 4 * g);
#pragma unroll
        for (int i = 0; i < 4; ++i) rotate<T>(x.e[2 * i], x.e[2 * i + 1], c[i], s[i]);
        st8<T>(row + 8 * g, x);
      }
      if (crow) to_cache<T, kQuant>(crow, 8 * g, x, ks);
    }
  }
  if (!do_cache) return;
  // ---- B: dims of k past rot_dim are not rotated, only cached ----------------------
  const int upp = (D - rot) / 8;
  for (int u = threadIdx.x; u < Hkv * upp; u += blockDim.x) {
    const int h = u / upp, g = u - h * upp;
    const T* row = static_cast<const T*>(a.k) + t * a.k_stride0 + (int64_t)h * a.k_stride1;
    to_cache<T, kQuant>(kc_tok + (int64_t)h * a.kc_head_stride * esz_c, rot + 8 * g, ld8<T>(row + rot + 8 * g), ks);
  }
  // ---- C: v -> cache ----------------------------------------------------------------
  const int upv = D / 8;
  for (int u = threadIdx.x; u < Hkv * upv; u += blockDim.x) {
    const int h = u / upv, g = u - h * upv;
    const T* row = static_cast<const T*>(a.v) + t * a.v_stride0 + (int64_t)h * a.v_stride1;
    to_cache<T, kQuant>(vc_tok + (int64_t)h * a.vc_head_stride * esz_c, 8 * g, ld8<T>(row + 8 * g), vs);
  }
}

}  // namespace

int launch_rope_cache(const b200attn_rope_cache_params& p, cudaStream_t stream) {
  const int64_t tokens = p.rope_tokens > p.cache_tokens ? p.rope_tokens : p.cache_tokens;
  if (tokens <= 0) return B200ATTN_OK;
  if (!p.q || !p.k || !p.v || !p.k_cache || !p.v_cache || (p.rope_tokens > 0 && (!p.positions || !p.cos_sin_cache)) ||
      (p.cache_tokens > 0 && !p.slot_mapping)) {
    set_error("rope_and_cache_flash: null pointer");
    return B200ATTN_E_NULL;
  }
  if (p.src_dtype != B200ATTN_BF16 && p.src_dtype != B200ATTN_F16) {
    set_error("rope_and_cache_flash: q/k/v/cos_sin_cache must be bf16 or f16 (got %d)", p.src_dtype);
    return B200ATTN_E_BAD_DTYPE;
  }
  const bool quant = p.kv_dtype == B200ATTN_FP8_E4M3;
  if (!quant && p.kv_dtype != p.src_dtype) {
    set_error("rope_and_cache_flash: kv_dtype %d must equal src_dtype %d or be fp8_e4m3", p.kv_dtype, p.src_dtype);
    return B200ATTN_E_BAD_DTYPE;
  }
  if (quant && (!p.k_scale || !p.v_scale)) {
    set_error("rope_and_cache_flash: fp8 cache needs k_scale and v_scale");
    return B200ATTN_E_NULL;
  }
  const int D = p.head_size, rot = p.rot_dim;
  if (p.num_q_heads <= 0 || p.num_kv_heads <= 0 || D <= 0 || p.block_size <= 0 || rot <= 0 || rot > D || (rot & 1)) {
    set_error("rope_and_cache_flash: bad shape (Hq %d Hkv %d D %d rot %d bs %d)", p.num_q_heads, p.num_kv_heads, D, rot,
              p.block_size);
    return B200ATTN_E_BAD_SHAPE;
  }
  auto m8 = [](int64_t x) { return x % 8 == 0; };
  auto a16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  const int cm = quant ? 16 : 8;  // cache strides: 16-byte multiples in elements of the cache type
  const bool geom = m8(D) && m8(D - rot) && (p.is_neox ? m8(rot / 2) : m8(rot)) && m8(p.q_stride0) && m8(p.q_stride1) &&
                    m8(p.k_stride0) && m8(p.k_stride1) && m8(p.v_stride0) && m8(p.v_stride1) && a16(p.q) && a16(p.k) &&
                    a16(p.v) && a16(p.cos_sin_cache) && a16(p.k_cache) && a16(p.v_cache) &&
                    p.kc_block_stride % cm == 0 && p.kc_slot_stride % cm == 0 && p.kc_head_stride % cm == 0 &&
                    p.vc_block_stride % cm == 0 && p.vc_slot_stride % cm == 0 && p.vc_head_stride % cm == 0;
  if (!geom) {
    set_error("rope_and_cache_flash: needs 16-byte aligned rows (head_size, rot_dim halves and strides multiples of 8)");
    return B200ATTN_E_ALIGNMENT;
  }
  RopeArgs a;
  a.q = p.q, a.k = p.k, a.v = p.v, a.positions = p.positions, a.cos_sin = p.cos_sin_cache;
  a.kc = p.k_cache, a.vc = p.v_cache, a.slot_mapping = p.slot_mapping, a.k_scale = p.k_scale, a.v_scale = p.v_scale;
  a.q_stride0 = p.q_stride0, a.q_stride1 = p.q_stride1, a.k_stride0 = p.k_stride0, a.k_stride1 = p.k_stride1;
  a.v_stride0 = p.v_stride0, a.v_stride1 = p.v_stride1;
  a.kc_block_stride = p.kc_block_stride, a.kc_slot_stride = p.kc_slot_stride, a.kc_head_stride = p.kc_head_stride;
  a.vc_block_stride = p.vc_block_stride, a.vc_slot_stride = p.vc_slot_stride, a.vc_head_stride = p.vc_head_stride;
  a.rope_tokens = p.rope_tokens, a.cache_tokens = p.cache_tokens;
  a.num_q_heads = p.num_q_heads, a.num_kv_heads = p.num_kv_heads, a.head_size = D, a.rot_dim = rot;
  a.block_size = p.block_size;
  int units = (p.num_q_heads + p.num_kv_heads) * (p.is_neox ? rot / 2 : rot) / 8;
  int threads = units < 64 ? 64 : (units > 256 ? 256 : ((units + 31) / 32) * 32);
  const unsigned grid = (unsigned)tokens;
  const bool bf = p.src_dtype == B200ATTN_BF16;
#define B200_ROPE_LAUNCH(T, NEOX, QUANT) launch_pdl(rope_cache_kernel<T, NEOX, QUANT>, grid, threads, 0, stream, !(tuning().reserved & 512), a)
  if (bf) {
    if (p.is_neox) { if (quant) B200_ROPE_LAUNCH(__nv_bfloat16, true, true); else B200_ROPE_LAUNCH(__nv_bfloat16, true, false); }
    else { if (quant) B200_ROPE_LAUNCH(__nv_bfloat16, false, true); else B200_ROPE_LAUNCH(__nv_bfloat16, false, false); }
  } else {
    if (p.is_neox) { if (quant) B200_ROPE_LAUNCH(__half, true, true); else B200_ROPE_LAUNCH(__half, true, false); }
    else { if (quant) B200_ROPE_LAUNCH(__half, false, true); else B200_ROPE_LAUNCH(__half, false, false); }
  }
#undef B200_ROPE_LAUNCH
  note_launch("rope_cache_write");
  return check_launch("rope_and_cache_flash");
}

}  // namespace b200attn

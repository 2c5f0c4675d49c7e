// K1: KV-cache scatter write (+ optional fused FP8-e4m3 quantise), K4: static
// per-tensor FP8 quantisation of Q.
//
// Semantics follow the reference's cache write, reshape_and_cache_flash_kernel
// (reference vllm/csrc/cache_kernels.cu:286-358): one source row per token,
// slot < 0 skipped, fp8 = e4m3_rn_satfinite(float(x) / scale) with a true fp32
// division (quant_utils.cuh:479-502).  The implementation is new: every thread
// moves 16 bytes per store (ld.global.nc.v4 / st.global.v4), K and V rows of a
// token go to different CTAs (grid.y), and the index math is 64-bit for any
// block/slot/head stride (NHD, HND, K/V-interleaved allocations).
#include "common.cuh"

namespace b200attn {
namespace {

__device__ __forceinline__ int4 ld_nc_v4(const void* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_v4(void* p, const int4& v) {
  asm volatile("st.global.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void st_v2(void* p, const int2& v) {
  asm volatile("st.global.v2.s32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// x / scale, RNE, saturating; two values -> two e4m3 bytes
__device__ __forceinline__ uint16_t quant2(float a, float b, float scale) {
  float2 v = make_float2(__fdiv_rn(a, scale), __fdiv_rn(b, scale));
  return (uint16_t)__nv_cvt_float2_to_fp8x2(v, __NV_SATFINITE, __NV_E4M3);
}

template <typename SrcT> __device__ __forceinline__ void unpack8(const int4& raw, float (&f)[8]);
template <> __device__ __forceinline__ void unpack8<__nv_bfloat16>(const int4& raw, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
template <> __device__ __forceinline__ void unpack8<__half>(const int4& raw, float (&f)[8]) {
  const __half2* h = reinterpret_cast<const __half2*>(&raw);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __half22float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}

struct CacheWriteArgs {
  const void* src[2];  // key, value
  void* dst[2];        // k_cache, v_cache
  const int64_t* slot_mapping;
  int64_t src_stride0[2];
  int64_t block_stride[2], slot_stride[2], head_stride[2];
  const float* scale[2];
  int num_kv_heads, head_size, block_size;
};

// A CTA takes kTokPerCta consecutive tokens (K and V rows of each) when there are enough tokens to fill the GPU that
// way: with one (token, K|V) row per CTA - 128 threads moving 16 bytes each - a prompt-sized write was 16 384 tiny CTAs
// and ran at half the HBM rate (8192 tokens: 20.5 us, profiles/r02_kbench_baseline.jsonl).  Decode-sized writes keep
// one token per CTA (they are launch latency either way).
constexpr int kMaxTokPerCta = 8;

struct RowRef {
  const char* src;
  char* dst;
};

// Copy path: 16-byte vectors of kElemBytes-sized elements, 4 vectors in flight per thread.
template <int kElemBytes>
__global__ void __launch_bounds__(256) cache_copy_vec_kernel(CacheWriteArgs a, int tok_per_cta, int64_t num_tokens) {
  pdl_launch_dependents();  // the attention launch behind us may start its prologue now (common.cuh)
  pdl_wait();               // ... and this grid may itself have been scheduled early: wait for the kernel in front
  constexpr int kVec = 16 / kElemBytes;
  __shared__ int64_t s_slot[kMaxTokPerCta];
  const int64_t t0 = (int64_t)blockIdx.x * tok_per_cta;
  const int ntok = (int)min((int64_t)tok_per_cta, num_tokens - t0);
  if (threadIdx.x < ntok) s_slot[threadIdx.x] = a.slot_mapping[t0 + threadIdx.x];
  __syncthreads();
  const int vec_per_head = a.head_size / kVec;
  const int n_vec = a.num_kv_heads * vec_per_head;  // vectors of one row
  const int total = ntok * 2 * n_vec;
  auto locate = [&](int u, const char*& src, char*& dst) {  // false: padded token
    const int tok = u / (2 * n_vec), r = u - tok * 2 * n_vec, which = r / n_vec, v = r - which * n_vec;
    const int64_t slot = s_slot[tok];
    if (slot < 0) return false;
    const int64_t blk = slot / a.block_size, off = slot % a.block_size;
    const int h = v / vec_per_head, dv = v - h * vec_per_head;
    src = static_cast<const char*>(a.src[which]) + (t0 + tok) * a.src_stride0[which] * kElemBytes + (int64_t)v * 16;
    dst = static_cast<char*>(a.dst[which]) +
          (blk * a.block_stride[which] + off * a.slot_stride[which] + (int64_t)h * a.head_stride[which]) * kElemBytes + (int64_t)dv * 16;
    return true;
  };
  for (int u0 = threadIdx.x; u0 < total; u0 += 4 * blockDim.x) {
    int4 x[4];
    char* dst[4];
    bool ok[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int u = u0 + i * blockDim.x;
      const char* src = nullptr;
      dst[i] = nullptr;
      ok[i] = u < total && locate(u, src, dst[i]);
      if (ok[i]) x[i] = ld_nc_v4(src);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (ok[i]) st_v4(dst[i], x[i]);
  }
}

// FP8 path, 16 elements per thread and unit: two 16-byte loads, one 16-byte store; 2 units in flight per thread.
template <typename SrcT>
__global__ void __launch_bounds__(256) cache_quant_vec16_kernel(CacheWriteArgs a, int tok_per_cta, int64_t num_tokens) {
  pdl_launch_dependents();  // the attention launch behind us may start its prologue now (common.cuh)
  pdl_wait();               // ... and this grid may itself have been scheduled early: wait for the kernel in front
  __shared__ int64_t s_slot[kMaxTokPerCta];
  const int64_t t0 = (int64_t)blockIdx.x * tok_per_cta;
  const int ntok = (int)min((int64_t)tok_per_cta, num_tokens - t0);
  if (threadIdx.x < ntok) s_slot[threadIdx.x] = a.slot_mapping[t0 + threadIdx.x];
  __syncthreads();
  const float scale[2] = {*a.scale[0], *a.scale[1]};
  const int vec_per_head = a.head_size / 16;
  const int n_vec = a.num_kv_heads * vec_per_head;
  const int total = ntok * 2 * n_vec;
  for (int u0 = threadIdx.x; u0 < total; u0 += 2 * blockDim.x) {
    int4 lo[2], hi[2];
    uint8_t* dst[2];
    float sc[2];
    bool ok[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int u = u0 + i * blockDim.x;
      ok[i] = false;
      if (u < total) {
        const int tok = u / (2 * n_vec), r = u - tok * 2 * n_vec, which = r / n_vec, v = r - which * n_vec;
        const int64_t slot = s_slot[tok];
        if (slot >= 0) {
          const int64_t blk = slot / a.block_size, off = slot % a.block_size;
          const int h = v / vec_per_head, dv = v - h * vec_per_head;
          const SrcT* src = static_cast<const SrcT*>(a.src[which]) + (t0 + tok) * a.src_stride0[which] + (int64_t)v * 16;
          dst[i] = static_cast<uint8_t*>(a.dst[which]) + blk * a.block_stride[which] + off * a.slot_stride[which] +
                   (int64_t)h * a.head_stride[which] + dv * 16;
          lo[i] = ld_nc_v4(src), hi[i] = ld_nc_v4(src + 8);
          sc[i] = scale[which];
          ok[i] = true;
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      if (!ok[i]) continue;
      float f[16];
      unpack8<SrcT>(lo[i], *reinterpret_cast<float(*)[8]>(&f[0]));
      unpack8<SrcT>(hi[i], *reinterpret_cast<float(*)[8]>(&f[8]));
      uint32_t w[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        w[j] = (uint32_t)quant2(f[4 * j], f[4 * j + 1], sc[i]) | ((uint32_t)quant2(f[4 * j + 2], f[4 * j + 3], sc[i]) << 16);
      st_v4(dst[i], make_int4(w[0], w[1], w[2], w[3]));
    }
  }
}

// Element-wise net for shapes/alignments the vector kernels do not take.
template <typename SrcT, bool kQuant>
__global__ void __launch_bounds__(256) cache_scalar_kernel(CacheWriteArgs a) {
  pdl_launch_dependents();  // the attention launch behind us may start its prologue now (common.cuh)
  pdl_wait();               // ... and this grid may itself have been scheduled early: wait for the kernel in front
  const int64_t token = blockIdx.x;
  const int which = blockIdx.y;
  const int64_t slot = a.slot_mapping[token];
  if (slot < 0) return;
  const int64_t blk = slot / a.block_size, off = slot % a.block_size;
  const SrcT* src = static_cast<const SrcT*>(a.src[which]) + token * a.src_stride0[which];
  const int n = a.num_kv_heads * a.head_size;
  const int64_t base = blk * a.block_stride[which] + off * a.slot_stride[which];
  if constexpr (kQuant) {
    const float scale = *a.scale[which];
    uint8_t* dst = static_cast<uint8_t*>(a.dst[which]);
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
      const int h = e / a.head_size, d = e - h * a.head_size;
      float x = __fdiv_rn(to_float<SrcT>(src[e]), scale);
      dst[base + (int64_t)h * a.head_stride[which] + d] = (uint8_t)__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3);
    }
  } else {
    SrcT* dst = static_cast<SrcT*>(a.dst[which]);
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
      const int h = e / a.head_size, d = e - h * a.head_size;
      dst[base + (int64_t)h * a.head_stride[which] + d] = src[e];
    }
  }
}

// K4: out = e4m3(float(in) * (1/scale)), one CTA per row
// (reference vllm/csrc/quantization/w8a8/fp8/common.cu:11-29, common.cuh:40-60)
template <typename SrcT>
__global__ void __launch_bounds__(256) scaled_fp8_quant_kernel(uint8_t* out, const SrcT* in, int64_t row_elems,
                                                               int64_t in_row_stride, const float* scale) {
  const int64_t row = blockIdx.x;
  const float inv = 1.0f / *scale;
  const SrcT* src = in + row * in_row_stride;
  uint8_t* dst = out + row * row_elems;
  for (int64_t e = threadIdx.x; e < row_elems; e += blockDim.x) {
    float x = to_float<SrcT>(src[e]) * inv;
    x = fmaxf(-448.0f, fminf(448.0f, x));
    dst[e] = (uint8_t)__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3);
  }
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// tokens per CTA of the vector kernels: one while that still gives every SM a few CTAs, up to kMaxTokPerCta
inline int tok_per_cta(int64_t num_tokens) {
  const int64_t want = num_tokens / (4 * (device_info().num_sms > 0 ? device_info().num_sms : 148));
  return want < 1 ? 1 : (want > kMaxTokPerCta ? kMaxTokPerCta : (int)want);
}

}  // namespace

int launch_cache_write(const void* key, const void* value, void* k_cache, void* v_cache, const int64_t* slot_mapping,
                       int64_t num_tokens, int num_kv_heads, int head_size, int block_size, int64_t key_stride0,
                       int64_t value_stride0, int64_t kc_block_stride, int64_t kc_slot_stride, int64_t kc_head_stride,
                       int64_t vc_block_stride, int64_t vc_slot_stride, int64_t vc_head_stride, int src_dtype,
                       int kv_dtype, const float* k_scale, const float* v_scale, cudaStream_t stream) {
  if (num_tokens == 0) return B200ATTN_OK;
  if (!key || !value || !k_cache || !v_cache || !slot_mapping) {
    set_error("reshape_and_cache_flash: null pointer");
    return B200ATTN_E_NULL;
  }
  if (num_kv_heads <= 0 || head_size <= 0 || block_size <= 0 || num_tokens < 0) {
    set_error("reshape_and_cache_flash: bad shape (heads %d, head_size %d, block_size %d)", num_kv_heads, head_size,
              block_size);
    return B200ATTN_E_BAD_SHAPE;
  }
  const bool quant = (kv_dtype == B200ATTN_FP8_E4M3) && (src_dtype != B200ATTN_FP8_E4M3);
  if (!quant && kv_dtype != src_dtype) {
    set_error("reshape_and_cache_flash: kv_dtype %d must equal src_dtype %d or be fp8_e4m3", kv_dtype, src_dtype);
    return B200ATTN_E_BAD_DTYPE;
  }
  if (quant && (src_dtype != B200ATTN_BF16 && src_dtype != B200ATTN_F16 && src_dtype != B200ATTN_F32)) {
    set_error("reshape_and_cache_flash: unsupported source dtype %d", src_dtype);
    return B200ATTN_E_BAD_DTYPE;
  }
  if (quant && (!k_scale || !v_scale)) {
    set_error("reshape_and_cache_flash: fp8 cache needs k_scale and v_scale");
    return B200ATTN_E_NULL;
  }
  CacheWriteArgs a;
  a.src[0] = key, a.src[1] = value;
  a.dst[0] = k_cache, a.dst[1] = v_cache;
  a.slot_mapping = slot_mapping;
  a.src_stride0[0] = key_stride0, a.src_stride0[1] = value_stride0;
  a.block_stride[0] = kc_block_stride, a.block_stride[1] = vc_block_stride;
  a.slot_stride[0] = kc_slot_stride, a.slot_stride[1] = vc_slot_stride;
  a.head_stride[0] = kc_head_stride, a.head_stride[1] = vc_head_stride;
  a.scale[0] = k_scale, a.scale[1] = v_scale;
  a.num_kv_heads = num_kv_heads, a.head_size = head_size, a.block_size = block_size;

  const int sb = dtype_bytes(src_dtype), kb = dtype_bytes(kv_dtype);
  if (sb == 0 || kb == 0) {
    set_error("reshape_and_cache_flash: bad dtype");
    return B200ATTN_E_BAD_DTYPE;
  }
  auto mult = [](int64_t x, int64_t m) { return x % m == 0; };
  const dim3 grid((unsigned)num_tokens, 2);
  const int n_elems = num_kv_heads * head_size;
  const bool pdl = !(tuning().reserved & 512);  // common.cuh: launch_pdl

  if (!quant) {
    const int vec = 16 / sb;
    const bool vec_ok = mult(head_size, vec) && aligned16(key) && aligned16(value) && aligned16(k_cache) &&
                        aligned16(v_cache) && mult(key_stride0, vec) && mult(value_stride0, vec) &&
                        mult(kc_block_stride, vec) && mult(kc_slot_stride, vec) && mult(kc_head_stride, vec) &&
                        mult(vc_block_stride, vec) && mult(vc_slot_stride, vec) && mult(vc_head_stride, vec);
    if (vec_ok) {
      const int units = 2 * n_elems / vec;  // 16-byte vectors of one token (K and V)
      const int tpc = tok_per_cta(num_tokens);
      int threads = (tpc * units + 3) / 4;
      threads = threads < 32 ? 32 : (threads > 256 ? 256 : ((threads + 31) / 32) * 32);
      const dim3 vgrid((unsigned)((num_tokens + tpc - 1) / tpc));
      if (sb == 2) launch_pdl(cache_copy_vec_kernel<2>, vgrid, threads, 0, stream, pdl, a, tpc, num_tokens);
      else if (sb == 4) launch_pdl(cache_copy_vec_kernel<4>, vgrid, threads, 0, stream, pdl, a, tpc, num_tokens);
      else launch_pdl(cache_copy_vec_kernel<1>, vgrid, threads, 0, stream, pdl, a, tpc, num_tokens);
      note_launch("cache_copy_vec");
    } else {
      int threads = n_elems > 256 ? 256 : ((n_elems + 31) / 32) * 32;
      if (src_dtype == B200ATTN_BF16) launch_pdl(cache_scalar_kernel<__nv_bfloat16, false>, grid, threads, 0, stream, pdl, a);
      else if (src_dtype == B200ATTN_F16) launch_pdl(cache_scalar_kernel<__half, false>, grid, threads, 0, stream, pdl, a);
      else if (src_dtype == B200ATTN_F32) launch_pdl(cache_scalar_kernel<float, false>, grid, threads, 0, stream, pdl, a);
      else launch_pdl(cache_scalar_kernel<uint8_t, false>, grid, threads, 0, stream, pdl, a);
      note_launch("cache_copy_scalar");
    }
  } else {
    const bool vec_ok = sb == 2 && mult(head_size, 16) && aligned16(key) && aligned16(value) && aligned16(k_cache) &&
                        aligned16(v_cache) && mult(key_stride0, 8) && mult(value_stride0, 8) &&
                        mult(kc_block_stride, 16) && mult(kc_slot_stride, 16) && mult(kc_head_stride, 16) &&
                        mult(vc_block_stride, 16) && mult(vc_slot_stride, 16) && mult(vc_head_stride, 16);
    if (vec_ok) {
      const int units = 2 * n_elems / 16;  // 16-element groups of one token (K and V)
      const int tpc = tok_per_cta(num_tokens);
      int threads = (tpc * units + 1) / 2;
      threads = threads < 32 ? 32 : (threads > 256 ? 256 : ((threads + 31) / 32) * 32);
      const dim3 vgrid((unsigned)((num_tokens + tpc - 1) / tpc));
      if (src_dtype == B200ATTN_BF16) launch_pdl(cache_quant_vec16_kernel<__nv_bfloat16>, vgrid, threads, 0, stream, pdl, a, tpc, num_tokens);
      else launch_pdl(cache_quant_vec16_kernel<__half>, vgrid, threads, 0, stream, pdl, a, tpc, num_tokens);
      note_launch("cache_quant_vec16");
    } else {
      int threads = n_elems > 256 ? 256 : ((n_elems + 31) / 32) * 32;
      if (src_dtype == B200ATTN_BF16) launch_pdl(cache_scalar_kernel<__nv_bfloat16, true>, grid, threads, 0, stream, pdl, a);
      else if (src_dtype == B200ATTN_F16) launch_pdl(cache_scalar_kernel<__half, true>, grid, threads, 0, stream, pdl, a);
      else launch_pdl(cache_scalar_kernel<float, true>, grid, threads, 0, stream, pdl, a);
      note_launch("cache_quant_scalar");
    }
  }
  return check_launch("reshape_and_cache_flash");
}

// ---- K1 with dynamic per-(token, head) quantisation -------------------------------------------------------------
// One warp per (token, kv head, K|V) row: the row is read once into registers (lane l holds elements 4l .. 4l+3 of
// every 128-element group), the absmax goes through five shuffles, and the row leaves as 4-byte stores of e4m3 or
// int8 codes plus one fp32 scale.  Arithmetic as the upstream Triton kernel states it (vLLM 0.22
// vllm/v1/attention/ops/triton_reshape_and_cache_flash.py:186-236; oracle/cache_write.py per_token_head_quant_ref):
// scale = max(absmax / QMAX, 1e-6), code = Q(clamp(x * (1 / scale), QMIN, QMAX)) with IEEE division, e4m3 by
// round-to-nearest-even, int8 by truncation toward zero.
// R rows per warp, all of their loads issued before the first reduction (one row per warp left the kernel waiting on
// a single 8-byte load per lane: 8192 tokens took 34 us, twice the per-tensor quantising write).
template <typename SrcT, bool kInt8, bool kVec, int R>
__global__ void __launch_bounds__(256) cache_quant_per_token_head_kernel(const PerTokenHeadWriteArgs a) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  // 32-bit index arithmetic (the launcher checks that the row count fits): with 64-bit divisions for the token, the
  // head and the page of every row, the index arithmetic was five times the quantisation itself
  const uint32_t row0 = (blockIdx.x * 8u + (threadIdx.x >> 5)) * (uint32_t)R;  // (token, head, K|V), K|V fastest
  const uint32_t rows = (uint32_t)(a.num_tokens * a.num_kv_heads * 2);
  const uint32_t hkv = (uint32_t)a.num_kv_heads;
  const int D = a.head_size;
  constexpr float kQMax = kInt8 ? 127.f : 448.f, kQMin = kInt8 ? -128.f : -448.f;
  constexpr int kGroups = 2;  // head_size <= 256
  float x[R][kGroups][4];
  uint8_t* dst[R];
  float* sdst[R];
  bool live[R];
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const uint32_t row = row0 + j;
    live[j] = false, dst[j] = nullptr, sdst[j] = nullptr;
#pragma unroll
    for (int g = 0; g < kGroups; ++g) x[j][g][0] = x[j][g][1] = x[j][g][2] = x[j][g][3] = 0.f;
    if (row >= rows) continue;
    const int w = (int)(row & 1u);
    const uint32_t t = (row >> 1) / hkv;
    const int h = (int)((row >> 1) - t * hkv);
    const int64_t slot = a.slot_mapping[t];
    if (slot < 0) continue;  // padding token (:171-173)
    live[j] = true;
    int64_t blk, off;
    if (a.bs_shift >= 0) blk = slot >> a.bs_shift, off = slot & (a.block_size - 1);
    else if (slot < (int64_t(1) << 31)) blk = (uint32_t)slot / (uint32_t)a.block_size, off = (uint32_t)slot - (uint32_t)blk * (uint32_t)a.block_size;
    else blk = slot / a.block_size, off = slot % a.block_size;
    const SrcT* src = static_cast<const SrcT*>(a.src[w]) + (int64_t)t * a.src_stride0[w] + (int64_t)h * a.src_stride1[w];
    dst[j] = static_cast<uint8_t*>(a.dst[w]) + blk * a.block_stride[w] + off * a.slot_stride[w] + (int64_t)h * a.head_stride[w];
    sdst[j] = a.scales[w] + blk * a.s_block_stride[w] + off * a.s_slot_stride[w] + (int64_t)h * a.s_head_stride[w];
#pragma unroll
    for (int g = 0; g < kGroups; ++g) {
      if constexpr (kVec) {
        const int d0 = (g * 32 + lane) * 4;
        if (d0 < D) {
          if constexpr (sizeof(SrcT) == 2) {
            const uint2 raw = *reinterpret_cast<const uint2*>(src + d0);
            const SrcT* e = reinterpret_cast<const SrcT*>(&raw);
#pragma unroll
            for (int i = 0; i < 4; ++i) x[j][g][i] = to_float<SrcT>(e[i]);
          } else {
            const float4 raw = *reinterpret_cast<const float4*>(src + d0);
            x[j][g][0] = raw.x, x[j][g][1] = raw.y, x[j][g][2] = raw.z, x[j][g][3] = raw.w;
          }
        }
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int d = (g * 4 + i) * 32 + lane;
          if (d < D) x[j][g][i] = to_float<SrcT>(src[d]);
        }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < R; ++j) {
    float amax = 0.f;
#pragma unroll
    for (int g = 0; g < kGroups; ++g)
#pragma unroll
      for (int i = 0; i < 4; ++i) amax = fmaxf(amax, fabsf(x[j][g][i]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if (!live[j]) continue;  // (warp-uniform)
    const float scale = fmaxf(__fdiv_rn(amax, kQMax), 1e-6f);
    const float inv = __fdiv_rn(1.0f, scale);
    if (lane == 0) *sdst[j] = scale;
    auto code = [&](float v) -> uint32_t {
      const float y = fminf(fmaxf(v * inv, kQMin), kQMax);
      if constexpr (kInt8) return (uint32_t)(uint8_t)(int8_t)__float2int_rz(y);
      else return (uint32_t)fp8_out(y);
    };
#pragma unroll
    for (int g = 0; g < kGroups; ++g) {
      if constexpr (kVec) {
        const int d0 = (g * 32 + lane) * 4;
        if (d0 < D) {
          uint32_t w4;
          if constexpr (kInt8) {
            w4 = code(x[j][g][0]) | (code(x[j][g][1]) << 8) | (code(x[j][g][2]) << 16) | (code(x[j][g][3]) << 24);
          } else {
            // e4m3: the saturating conversion IS the clamp to +-448 (for every finite value; a NaN stays a NaN, as
            // in the oracle), two values per conversion
            const uint32_t lo = __nv_cvt_float2_to_fp8x2(make_float2(x[j][g][0] * inv, x[j][g][1] * inv), __NV_SATFINITE, __NV_E4M3);
            const uint32_t hi = __nv_cvt_float2_to_fp8x2(make_float2(x[j][g][2] * inv, x[j][g][3] * inv), __NV_SATFINITE, __NV_E4M3);
            w4 = lo | (hi << 16);
          }
          *reinterpret_cast<uint32_t*>(dst[j] + d0) = w4;
        }
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int d = (g * 4 + i) * 32 + lane;
          if (d < D) dst[j][d] = (uint8_t)code(x[j][g][i]);
        }
      }
    }
  }
}

template <typename SrcT, int R>
static void launch_pth_r(const PerTokenHeadWriteArgs& a, bool vec, cudaStream_t stream) {
  const int64_t rows = a.num_tokens * a.num_kv_heads * 2;
  const dim3 grid((unsigned)((rows + 8 * R - 1) / (8 * R)));
  const bool pdl = !(tuning().reserved & 512);
  const bool i8 = a.kv_dtype == B200ATTN_INT8;
  if (i8 && vec) launch_pdl(cache_quant_per_token_head_kernel<SrcT, true, true, R>, grid, 256, 0, stream, pdl, a);
  else if (i8) launch_pdl(cache_quant_per_token_head_kernel<SrcT, true, false, R>, grid, 256, 0, stream, pdl, a);
  else if (vec) launch_pdl(cache_quant_per_token_head_kernel<SrcT, false, true, R>, grid, 256, 0, stream, pdl, a);
  else launch_pdl(cache_quant_per_token_head_kernel<SrcT, false, false, R>, grid, 256, 0, stream, pdl, a);
}

template <typename SrcT>
static void launch_pth(const PerTokenHeadWriteArgs& a, bool vec, cudaStream_t stream) {
  // decode-sized writes (a few hundred rows) keep one row per warp: they are latency, and more warps hide more of it
  const int64_t rows = a.num_tokens * a.num_kv_heads * 2;
  // (one, two and four rows per warp measure the same at prompt sizes - 30.3 / 29.6 / 31.9 us for 8192 tokens: the kernel is
  // bound by its instruction count, not by load latency)
  if (rows >= 4 * 8 * 2 * (int64_t)device_info().num_sms) launch_pth_r<SrcT, 2>(a, vec, stream);
  else launch_pth_r<SrcT, 1>(a, vec, stream);
}

int launch_cache_write_per_token_head(const PerTokenHeadWriteArgs& a_in, cudaStream_t stream) {
  PerTokenHeadWriteArgs a = a_in;
  a.bs_shift = -1;
  if (a.block_size > 0 && (a.block_size & (a.block_size - 1)) == 0)
    for (a.bs_shift = 0; (1 << a.bs_shift) < a.block_size; ++a.bs_shift) {}
  if (a.num_tokens == 0) return B200ATTN_OK;
  if (!a.src[0] || !a.src[1] || !a.dst[0] || !a.dst[1] || !a.scales[0] || !a.scales[1] || !a.slot_mapping) {
    set_error("reshape_and_cache_flash_per_token_head: null pointer argument");
    return B200ATTN_E_NULL;
  }
  if (a.num_kv_heads <= 0 || a.head_size <= 0 || a.head_size > 256 || a.block_size <= 0 || a.num_tokens < 0) {
    set_error("reshape_and_cache_flash_per_token_head: bad shape (Hkv=%d D=%d bs=%d; head_size <= 256)", a.num_kv_heads,
              a.head_size, a.block_size);
    return B200ATTN_E_BAD_SHAPE;
  }
  if (a.kv_dtype != B200ATTN_FP8_E4M3 && a.kv_dtype != B200ATTN_INT8) {
    set_error("reshape_and_cache_flash_per_token_head: cache dtype %d must be fp8_e4m3 or int8", a.kv_dtype);
    return B200ATTN_E_BAD_DTYPE;
  }
  const int sb = dtype_bytes(a.src_dtype);
  if (a.src_dtype != B200ATTN_BF16 && a.src_dtype != B200ATTN_F16 && a.src_dtype != B200ATTN_F32) {
    set_error("reshape_and_cache_flash_per_token_head: unsupported source dtype %d", a.src_dtype);
    return B200ATTN_E_BAD_DTYPE;
  }
  if (a.num_tokens * a.num_kv_heads * 2 >= (int64_t(1) << 31)) {
    set_error("reshape_and_cache_flash_per_token_head: %lld tokens x %d heads exceed 2^30 rows per call", (long long)a.num_tokens,
              a.num_kv_heads);
    return B200ATTN_E_BAD_SHAPE;
  }
  bool vec = a.head_size % 4 == 0;
  for (int w = 0; w < 2; ++w) {
    vec = vec && (reinterpret_cast<uintptr_t>(a.src[w]) % (4 * sb) == 0) && a.src_stride0[w] % 4 == 0 && a.src_stride1[w] % 4 == 0 &&
          (reinterpret_cast<uintptr_t>(a.dst[w]) % 4 == 0) && a.block_stride[w] % 4 == 0 && a.slot_stride[w] % 4 == 0 &&
          a.head_stride[w] % 4 == 0;
  }
  if (a.src_dtype == B200ATTN_BF16) launch_pth<__nv_bfloat16>(a, vec, stream);
  else if (a.src_dtype == B200ATTN_F16) launch_pth<__half>(a, vec, stream);
  else launch_pth<float>(a, vec, stream);
  note_launch(vec ? "cache_quant_per_token_head_vec" : "cache_quant_per_token_head");
  return check_launch("reshape_and_cache_flash_per_token_head");
}

int launch_scaled_fp8_quant(void* out, const void* in, int64_t rows, int64_t row_elems, int64_t in_row_stride,
                            int src_dtype, const float* scale, cudaStream_t stream) {
  if (rows == 0 || row_elems == 0) return B200ATTN_OK;
  if (!out || !in || !scale) {
    set_error("scaled_fp8_quant: null pointer");
    return B200ATTN_E_NULL;
  }
  int threads = row_elems > 256 ? 256 : (int)((row_elems + 31) / 32) * 32;
  uint8_t* o = static_cast<uint8_t*>(out);
  if (src_dtype == B200ATTN_BF16)
    scaled_fp8_quant_kernel<__nv_bfloat16><<<(unsigned)rows, threads, 0, stream>>>(
        o, static_cast<const __nv_bfloat16*>(in), row_elems, in_row_stride, scale);
  else if (src_dtype == B200ATTN_F16)
    scaled_fp8_quant_kernel<__half><<<(unsigned)rows, threads, 0, stream>>>(o, static_cast<const __half*>(in),
                                                                            row_elems, in_row_stride, scale);
  else if (src_dtype == B200ATTN_F32)
    scaled_fp8_quant_kernel<float><<<(unsigned)rows, threads, 0, stream>>>(o, static_cast<const float*>(in), row_elems,
                                                                           in_row_stride, scale);
  else {
    set_error("scaled_fp8_quant: unsupported source dtype %d", src_dtype);
    return B200ATTN_E_BAD_DTYPE;
  }
  note_launch("scaled_fp8_quant");
  return check_launch("scaled_fp8_quant");
}

}  // namespace b200attn

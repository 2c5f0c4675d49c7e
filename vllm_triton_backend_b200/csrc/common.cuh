// Shared device/host helpers for the b200attn kernels (sm_100a only).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200attn.h"

namespace b200attn {

// ---- host-side state shared by the C ABI (defined in cabi.cu) ---------------
void set_error(const char* fmt, ...);
void note_launch(const char* kernel_name);
int check_launch(const char* what);  // cudaGetLastError -> status

struct Tuning {
  int decode_heads_per_cta = 0;
  int decode_ctas_per_sm = 0;
  int decode_stages = 0;
  int reserved = 0;
};
Tuning& tuning();
// true while the current b200attn_unified_attention call is being captured into a CUDA graph: launchers then size
// grids from T and B only (max_seqlen_q / max_seqlen_k describe the captured batch, not the replayed ones)
bool launch_is_captured();
unsigned long long* decode_trace_buffer();  // debug (b200attn_debug_set_decode_trace)

struct DeviceInfo {
  int num_sms = 0;
  int max_smem_optin = 0;
  int cc_major = 0, cc_minor = 0;
  int device = 0;
};
const DeviceInfo& device_info();  // of the calling thread's current device
// cudaFuncAttributeMaxDynamicSharedMemorySize, set once per (device, kernel)
int ensure_dynamic_smem(const void* kernel, int bytes, const char* what);
constexpr int kMaxDecodeCtas = 160;  // decode grid = min(#SMs, this); the split-KV workspace is sized for it

struct SideStream {
  cudaStream_t stream = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
  bool ready = false, failed = false;
};
SideStream* side_stream(bool may_create);
// CTAs the split-KV kernel gets for the single-token rows of a mixed batch when they run beside the prompts
// (b200attn_unified_attention); == the full grid when running beside them is not worth it.  Host ints only.
int mixed_decode_ctas(const b200attn_attention_params& p, int grid);

// cuTensorMapEncodeTiled resolved through cudaGetDriverEntryPoint (no -lcuda, so
// the library also loads on machines without a driver).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn();
// TMA descriptors (tma_maps.cu); small per-thread cache keyed by pointer + strides
int make_kv_map(CUtensorMap* out, const void* base, int esz, int D, int bs, int hkv, int64_t nb, int64_t s_block,
                int64_t s_slot, int64_t s_head, int box_slots, int box_heads, int box_dc, int64_t pair_bytes = 0);
int make_thd_map(CUtensorMap* out, const void* base, int D, int heads, int64_t tokens, int64_t s_token, int64_t s_head,
                 int box_heads, int box_tokens);

// Which sequences of a batch a launch is responsible for: those with
// min_q_len <= query_len <= max_q_len.  A mixed batch is covered by up to three
// launches (decode rows -> split-KV kernel, short extends -> mma.sync kernel,
// long prefills -> tcgen05 kernel); every grid is a host-int upper bound and
// CTAs whose sequence belongs to another launch exit at once.
struct QLenRange {
  int min_q_len = 0;
  int max_q_len = 0x7fffffff;
};

// FP8 prefill staging (kv_stage.cu): a 16-bit copy of the pages of the sequences the tcgen05 kernel
// takes, carved out of the workspace past the decode kernel's region.
struct KvStage {
  int32_t* staged_base = nullptr;  // [num_seqs]: first staging page of a sequence, or -1
  void* k_stage = nullptr;
  void* v_stage = nullptr;
  int cap_pages = 0;
};
bool plan_kv_stage(const b200attn_attention_params& p, KvStage* st);
b200attn_attention_params staged_params(const b200attn_attention_params& p, const KvStage& st);
int launch_kv_stage(const b200attn_attention_params& p, const KvStage& st, QLenRange range, cudaStream_t stream);
uint64_t stage_bytes_for_pages(int num_seqs, int64_t pages, int block_size, int num_kv_heads, int head_size);

// kernel launchers (one per .cu file)
int launch_cache_write(const void* key, const void* value, void* k_cache, void* v_cache, const int64_t* slot_mapping,
                       int64_t num_tokens, int num_kv_heads, int head_size, int block_size, int64_t key_stride0,
                       int64_t value_stride0, int64_t kc_block_stride, int64_t kc_slot_stride, int64_t kc_head_stride,
                       int64_t vc_block_stride, int64_t vc_slot_stride, int64_t vc_head_stride, int src_dtype,
                       int kv_dtype, const float* k_scale, const float* v_scale, cudaStream_t stream);
struct PerTokenHeadWriteArgs {
  const void* src[2];  // key, value
  void* dst[2];        // k_cache, v_cache
  float* scales[2];    // k_scale_cache, v_scale_cache
  const int64_t* slot_mapping;
  int64_t num_tokens;
  int64_t src_stride0[2], src_stride1[2];
  int64_t block_stride[2], slot_stride[2], head_stride[2];
  int64_t s_block_stride[2], s_slot_stride[2], s_head_stride[2];
  int num_kv_heads, head_size, block_size, src_dtype, kv_dtype;
  int bs_shift;  // log2(block_size) when it is a power of two, else -1 (filled in by the launcher)
};
int launch_cache_write_per_token_head(const PerTokenHeadWriteArgs& a, cudaStream_t stream);
int launch_scaled_fp8_quant(void* out, const void* in, int64_t rows, int64_t row_elems, int64_t in_row_stride,
                            int src_dtype, const float* scale, cudaStream_t stream);
int launch_rope_cache(const b200attn_rope_cache_params& p, cudaStream_t stream);
int launch_attn_generic(const b200attn_attention_params& p, cudaStream_t stream);
int launch_attn_decode(const b200attn_attention_params& p, cudaStream_t stream, bool only_single_token_seqs = false);
bool decode_supported(const b200attn_attention_params& p, const char** why, bool mixed_batch = false);
uint64_t decode_workspace_bytes(int num_seqs, int num_q_heads, int num_kv_heads, int head_size);
uint64_t decode_plan_bytes(int num_seqs);
int build_decode_plan(const b200attn_attention_params& p, void* plan, uint64_t plan_bytes, uint64_t* key_out, cudaStream_t stream);
int launch_attn_prefill(const b200attn_attention_params& p, cudaStream_t stream, QLenRange range = QLenRange(),
                        const int32_t* skip_staged = nullptr);
bool prefill_supported(const b200attn_attention_params& p, const char** why);
int launch_attn_prefill_tc(const b200attn_attention_params& p, cudaStream_t stream, QLenRange range = QLenRange(),
                           const int32_t* staged_base = nullptr);
bool prefill_tc_supported(const b200attn_attention_params& p, const char** why);
// head size 256 (attn_prefill_tc256.cu); launch_attn_prefill_tc / prefill_tc_supported route to these
int launch_attn_prefill_tc256(const b200attn_attention_params& p, cudaStream_t stream, QLenRange range, const int32_t* staged_base);
bool prefill_tc256_supported(const b200attn_attention_params& p, const char** why);
int prefill_tc_min_q_len(const b200attn_attention_params& p);  // shortest query the tcgen05 tile is worth using for

#ifdef __CUDACC__
// Launch with (or without) programmatic stream serialisation: the grid may be scheduled while the kernel in front
// of it on the stream is still running, provided that kernel has released its dependents; the launched kernel
// must execute pdl_wait() before its first global-memory access AND in every CTA before it exits (a CTA that
// left without waiting would let the grid complete, and the next dependent start, too early).
template <typename... KArgs, typename... Args>
inline void launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, bool pdl, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid, cfg.blockDim = block, cfg.dynamicSmemBytes = smem, cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr, cfg.numAttrs = pdl ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

// per-(token, head) scales (b200attn.h: k_scale_cache / v_scale_cache)
static inline bool per_token_head(const b200attn_attention_params& p) { return p.k_scale_cache != nullptr; }

static inline int dtype_bytes(int dt) {
  switch (dt) {
    case B200ATTN_BF16:
    case B200ATTN_F16: return 2;
    case B200ATTN_F32: return 4;
    case B200ATTN_FP8_E4M3:
    case B200ATTN_INT8: return 1;
    default: return 0;
  }
}

#ifdef __CUDACC__
// ---- device helpers ---------------------------------------------------------
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

template <typename T> __device__ __forceinline__ float to_float(T x);
template <> __device__ __forceinline__ float to_float<float>(float x) { return x; }
template <> __device__ __forceinline__ float to_float<__half>(__half x) { return __half2float(x); }
template <> __device__ __forceinline__ float to_float<__nv_bfloat16>(__nv_bfloat16 x) { return __bfloat162float(x); }
template <> __device__ __forceinline__ float to_float<__nv_fp8_e4m3>(__nv_fp8_e4m3 x) { return float(x); }
template <> __device__ __forceinline__ float to_float<int8_t>(int8_t x) { return (float)x; }

template <typename T> __device__ __forceinline__ T from_float(float x);
template <> __device__ __forceinline__ float from_float<float>(float x) { return x; }
template <> __device__ __forceinline__ __half from_float<__half>(float x) { return __float2half_rn(x); }
template <> __device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float x) { return __float2bfloat16_rn(x); }

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// cap * tanh(x / cap), the overflow-free form of apply_softcap
// (reference triton_unified_attention.py:24-29).
__device__ __forceinline__ float softcap_fn(float x, float cap, float inv_cap) {
  float e = __expf(2.0f * x * inv_cap);
  // tanh(y) = 1 - 2/(e^{2y}+1); e=inf -> 1, e=0 -> -1
  return cap * (1.0f - __fdividef(2.0f, e + 1.0f));
}

// one MUFU op (relative error 2^-11): the prefill kernels' softcap, cap * tanh_approx(x / cap)
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Attention sink (upstream `init_softmax_M`, vllm/v1/attention/ops/triton_attention_helpers.py):
// one extra key per head with logit `sink` (natural units, not scaled) and a zero value row.
// Applied once, after the last KV tile: `m` is the final running max (`m_is_log2`: in log2
// units) and `l` the running sum; returns the factor for the un-normalised accumulator.
template <bool m_is_log2> __device__ __forceinline__ float sink_rescale(float sink, float& m, float& l) {
  const float s = m_is_log2 ? sink * kLog2e : sink;
  const float mx = fmaxf(m, s);
  if (mx == -INFINITY) {
    l = 0.f;
    return 0.f;
  }
  const float f = (m == -INFINITY) ? 0.f : (m_is_log2 ? exp2f(m - mx) : __expf(m - mx));
  l = l * f + (m_is_log2 ? exp2f(s - mx) : __expf(s - mx));
  m = mx;  // (m, l) describe the row with the sink in it from here on
  return f;
}

// Natural-log sum of exponentials of a row's logits (b200attn_attention_params::lse) from the running maximum
// (log2 units when m_is_log2) and the exp-sum relative to it, as they stand after the last KV tile (and the sink).
// Nothing visible: -inf.
template <bool m_is_log2> __device__ __forceinline__ float row_lse(float m, float l) {
  if (!(l > 0.f)) return -INFINITY;
  return (m_is_log2 ? m * 0.6931471805599453f : m) + __logf(l);
}

// ---- fused FP8 output quantisation (b200attn_attention_params::out_scale) -----
// Upstream: `acc = acc * tl.load(out_scale)` with out_scale = 1/output_scale, clamp to the e4m3
// range, store as e4m3 (vllm/v1/attention/ops/triton_unified_attention.py:608-610, 722-724).
// Saturating round-to-nearest-even conversion == clamp + rn.
__device__ __forceinline__ float out_quant_factor(const float* out_scale) { return out_scale ? 1.0f / __ldg(out_scale) : 0.f; }
__device__ __forceinline__ uint32_t fp8x2_out(float lo, float hi) {
  return (uint32_t)__nv_cvt_float2_to_fp8x2(make_float2(lo, hi), __NV_SATFINITE, __NV_E4M3);
}
__device__ __forceinline__ uint32_t fp8x4_out(float a, float b, float c, float d) { return fp8x2_out(a, b) | (fp8x2_out(c, d) << 16); }
__device__ __forceinline__ uint8_t fp8_out(float x) { return (uint8_t)__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3); }

// ---- programmatic dependent launch (griddepcontrol) ---------------------------
// The cache write (K1 / K5) is always followed on the stream by the attention launch of the same layer.  The
// writers release their dependents as soon as they start; the decode kernel is launched with
// cudaLaunchAttributeProgrammaticStreamSerialization and sets up shared memory, barriers and descriptors while
// the writer is still running, then waits (full completion + memory flush of everything before it) in front
// of its first global-memory read.  After any other kernel (which never triggers) the launch behaves as usual.
// It works the other way round as well: the decode kernel releases ITS dependents when it starts and the cache
// writers are launched with the attribute and wait first thing, so in a loop of (cache write, attention) pairs
// - or behind a previous decode launch - the next launch's latency hides under the running kernel.  Waits are
// transitive (each kernel of the chain waits for its predecessor's completion before it completes itself).
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---- mbarrier / TMA / ldmatrix / mma wrappers --------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a pipeline bug must trap, not hang the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 26)) __trap();
  }
}

// 5-D tiled TMA load, completion on an mbarrier (cp.async.bulk.tensor -> UTMALDG)
__device__ __forceinline__ void tma_load_5d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1,
                                            int c2, int c3, int c4, uint64_t cache_policy) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2], %8;"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2),
      "r"(c3), "r"(c4), "l"(cache_policy)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x2(uint32_t (&r)[2], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x2_trans(uint32_t (&r)[2], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
}
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}

// D(16x8,f32) += A(16x16) * B(16x8); bf16 or f16 operands
template <bool kBf16>
__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  if constexpr (kBf16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
}

template <bool kBf16> __device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  if constexpr (kBf16) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  } else {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
}

// two int8 bytes -> f16x2 (exact): 0x6400 | (b ^ 0x80) is the half 1152 + b
__device__ __forceinline__ uint32_t s8x2_to_f16x2(uint16_t x) {
  const uint32_t t = __byte_perm((uint32_t)x ^ 0x8080u, 0x64646464u, 0x4140);
  const __half2 bias = __halves2half2(__ushort_as_half((unsigned short)0x6480), __ushort_as_half((unsigned short)0x6480));
  const __half2 r = __hsub2(*reinterpret_cast<const __half2*>(&t), bias);
  return *reinterpret_cast<const uint32_t*>(&r);
}
template <bool kInt8> __device__ __forceinline__ uint32_t kv8x2_to_f16x2(uint16_t x);
// two e4m3 bytes (low 16 bits of x) -> f16x2 (exact)
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2(uint16_t x) {
  uint32_t r;
  asm volatile("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(r) : "h"(x));
  return r;
}
template <> __device__ __forceinline__ uint32_t kv8x2_to_f16x2<false>(uint16_t x) { return e4m3x2_to_f16x2(x); }
template <> __device__ __forceinline__ uint32_t kv8x2_to_f16x2<true>(uint16_t x) { return s8x2_to_f16x2(x); }
// Four one-byte elements (b0 b1 b2 b3 of w, e4m3 or int8) -> two f16x2, exact.  kTranspose = false: lo = (b0, b1),
// hi = (b2, b3); true: lo = (b0, b2), hi = (b1, b3) - the V^T fragments of the decode kernel.  int8 costs five
// instructions per word this way (one XOR, two PRMT that also do the transposition, two HADD2) where converting
// 16 bits at a time cost ten.
template <bool kInt8, bool kTranspose> __device__ __forceinline__ void kv8x4_to_f16x2(uint32_t w, uint32_t& lo, uint32_t& hi) {
  if constexpr (kInt8) {
    const uint32_t x = w ^ 0x80808080u;
    const uint32_t l = __byte_perm(x, 0x64646464u, kTranspose ? 0x4240 : 0x4140);
    const uint32_t h = __byte_perm(x, 0x64646464u, kTranspose ? 0x4341 : 0x4342);
    const __half2 bias = __halves2half2(__ushort_as_half((unsigned short)0x6480), __ushort_as_half((unsigned short)0x6480));
    const __half2 lr = __hsub2(*reinterpret_cast<const __half2*>(&l), bias), hr = __hsub2(*reinterpret_cast<const __half2*>(&h), bias);
    lo = *reinterpret_cast<const uint32_t*>(&lr), hi = *reinterpret_cast<const uint32_t*>(&hr);
  } else {
    const uint32_t x = kTranspose ? __byte_perm(w, 0, 0x3120) : w;
    lo = e4m3x2_to_f16x2((uint16_t)(x & 0xffffu)), hi = e4m3x2_to_f16x2((uint16_t)(x >> 16));
  }
}

__device__ __forceinline__ void cp_async_16(uint32_t smem_addr, const void* gptr, bool pred) {
  int sz = pred ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_addr), "l"(gptr), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
#endif  // __CUDACC__

}  // namespace b200attn

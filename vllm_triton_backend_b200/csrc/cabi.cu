// extern "C" surface of libb200attn.so (declared in include/b200attn.h).
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include "common.cuh"

namespace b200attn {

namespace {
thread_local char g_error[512] = "";
thread_local uint64_t g_launches = 0;
thread_local const char* g_last_kernel = "";
Tuning g_tuning;  // benchmark A/B switches; set before launching, not while other threads launch
thread_local bool g_capturing = false;
unsigned long long* g_decode_trace = nullptr;
}  // namespace

unsigned long long* decode_trace_buffer() { return g_decode_trace; }

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

void note_launch(const char* kernel_name) {
  ++g_launches;
  g_last_kernel = kernel_name;
}

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: CUDA launch failed: %s", what, cudaGetErrorString(e));
    return B200ATTN_E_CUDA;
  }
  return B200ATTN_OK;
}

Tuning& tuning() { return g_tuning; }
bool launch_is_captured() { return g_capturing; }

// Per-device attributes and per-(device, kernel) launch attributes: a process may drive several GPUs from one
// thread (cudaSetDevice between calls), so nothing here is cached "for the first device seen".
constexpr int kMaxDevices = 64;

const DeviceInfo& device_info() {
  static DeviceInfo info[kMaxDevices];
  static std::once_flag once[kMaxDevices];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) dev = 0;
  std::call_once(once[dev], [dev] {
    DeviceInfo& d = info[dev];
    cudaDeviceGetAttribute(&d.num_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&d.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaDeviceGetAttribute(&d.cc_major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&d.cc_minor, cudaDevAttrComputeCapabilityMinor, dev);
    d.device = dev;
  });
  return info[dev];
}

int ensure_dynamic_smem(const void* kernel, int bytes, const char* what) {
  static std::mutex mu;
  static std::unordered_map<uint64_t, int> done;  // (kernel, device) -> bytes granted
  const uint64_t key = reinterpret_cast<uint64_t>(kernel) * 64u + (uint64_t)device_info().device;
  {
    std::lock_guard<std::mutex> lock(mu);
    auto it = done.find(key);
    if (it != done.end() && it->second >= bytes) return B200ATTN_OK;
  }
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) {
    set_error("%s: cudaFuncSetAttribute(%d bytes of shared memory) failed: %s", what, bytes, cudaGetErrorString(e));
    return B200ATTN_E_CUDA;
  }
  std::lock_guard<std::mutex> lock(mu);
  done[key] = bytes;
  return B200ATTN_OK;
}

// A second stream per device for the HBM-bound half of a mixed batch (see b200attn_unified_attention): created on
// first use, never destroyed (a per-device resource like the attributes above).  nullptr if it cannot be had.
SideStream* side_stream(bool may_create) {
  static SideStream streams[kMaxDevices];
  static std::mutex mu;
  const int dev = device_info().device;
  SideStream& ss = streams[dev];
  if (ss.ready) return &ss;
  if (!may_create) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  if (ss.ready) return &ss;
  if (ss.failed) return nullptr;
  int lo = 0, hi = 0;
  cudaStream_t st = nullptr;
  cudaEvent_t e1 = nullptr, e2 = nullptr;
  if (cudaDeviceGetStreamPriorityRange(&lo, &hi) != cudaSuccess ||
      cudaStreamCreateWithPriority(&st, cudaStreamNonBlocking, hi) != cudaSuccess ||
      cudaEventCreateWithFlags(&e1, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&e2, cudaEventDisableTiming) != cudaSuccess) {
    (void)cudaGetLastError();
    ss.failed = true;
    return nullptr;
  }
  ss.stream = st, ss.fork = e1, ss.join = e2, ss.ready = true;
  return &ss;
}

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

namespace {

int validate_attention(const b200attn_attention_params& p) {
  if (!p.out || !p.q || !p.k_cache || !p.v_cache || !p.block_tables || !p.seq_lens || !p.cu_seqlens_q) {
    set_error("unified_attention: null pointer argument");
    return B200ATTN_E_NULL;
  }
  if (p.num_seqs < 0 || p.num_tokens < 0 || p.num_q_heads <= 0 || p.num_kv_heads <= 0 || p.head_size <= 0 ||
      p.block_size <= 0 || p.num_q_heads % p.num_kv_heads != 0) {
    set_error("unified_attention: bad shape (T=%d B=%d Hq=%d Hkv=%d D=%d bs=%d)", p.num_tokens, p.num_seqs,
              p.num_q_heads, p.num_kv_heads, p.head_size, p.block_size);
    return B200ATTN_E_BAD_SHAPE;
  }
  if (p.q_dtype != B200ATTN_BF16 && p.q_dtype != B200ATTN_F16 && p.q_dtype != B200ATTN_F32) {
    // the reference rejects q_descale and we keep queries unquantised
    set_error("unified_attention: query dtype %d not supported (bf16, f16, f32)", p.q_dtype);
    return B200ATTN_E_BAD_DTYPE;
  }
  if ((p.k_scale_cache == nullptr) != (p.v_scale_cache == nullptr)) {
    set_error("unified_attention: k_scale_cache and v_scale_cache come together");
    return B200ATTN_E_NULL;
  }
  if (per_token_head(p)) {
    if (p.kv_dtype != B200ATTN_FP8_E4M3 && p.kv_dtype != B200ATTN_INT8) {
      set_error("unified_attention: per-(token, head) scales need an fp8_e4m3 or int8 cache (kv dtype %d)", p.kv_dtype);
      return B200ATTN_E_BAD_DTYPE;
    }
  } else if (p.kv_dtype != p.q_dtype && p.kv_dtype != B200ATTN_FP8_E4M3) {
    set_error("unified_attention: kv dtype %d must equal the query dtype or be fp8_e4m3 (int8: with scale caches)", p.kv_dtype);
    return B200ATTN_E_BAD_DTYPE;
  }
  if (p.head_size > 256) {
    set_error("unified_attention: head_size %d > 256", p.head_size);
    return B200ATTN_E_BAD_SHAPE;
  }
  if (p.out_scale && ((reinterpret_cast<uintptr_t>(p.out) & 7) || p.out_stride0 % 8 || p.out_stride1 % 8 || p.head_size % 8)) {
    set_error("unified_attention: fp8 output rows must be 8-byte aligned (out=%p strides %lld,%lld)", p.out,
              (long long)p.out_stride0, (long long)p.out_stride1);
    return B200ATTN_E_ALIGNMENT;
  }
  return B200ATTN_OK;
}

}  // namespace
}  // namespace b200attn

using namespace b200attn;

extern "C" {

int b200attn_abi_version(void) { return B200ATTN_ABI_VERSION; }
uint64_t b200attn_params_size(int32_t which) {
  return which == 0 ? sizeof(b200attn_attention_params) : which == 1 ? sizeof(b200attn_rope_cache_params) : 0;
}
const char* b200attn_last_error(void) { return g_error; }
uint64_t b200attn_launch_count(void) { return g_launches; }
void b200attn_reset_launch_count(void) { g_launches = 0; }
const char* b200attn_last_kernel(void) { return g_last_kernel; }

void b200attn_set_tuning(int32_t decode_heads_per_cta, int32_t decode_pool_sixteenths, int32_t decode_stages,
                         int32_t reserved) {
  g_tuning.decode_heads_per_cta = decode_heads_per_cta;
  g_tuning.decode_ctas_per_sm = decode_pool_sixteenths;
  g_tuning.decode_stages = decode_stages;
  g_tuning.reserved = reserved;
}

void b200attn_debug_set_decode_trace(void* device_buffer) {
  g_decode_trace = static_cast<unsigned long long*>(device_buffer);
}

int b200attn_reshape_and_cache_flash(const void* key, const void* value, void* k_cache, void* v_cache,
                                     const int64_t* slot_mapping, int64_t num_tokens, int32_t num_kv_heads,
                                     int32_t head_size, int32_t block_size, int64_t key_stride0, int64_t value_stride0,
                                     int64_t kc_block_stride, int64_t kc_slot_stride, int64_t kc_head_stride,
                                     int64_t vc_block_stride, int64_t vc_slot_stride, int64_t vc_head_stride,
                                     int32_t src_dtype, int32_t kv_dtype, const float* k_scale, const float* v_scale,
                                     b200attn_stream_t stream) {
  return launch_cache_write(key, value, k_cache, v_cache, slot_mapping, num_tokens, num_kv_heads, head_size,
                            block_size, key_stride0, value_stride0, kc_block_stride, kc_slot_stride, kc_head_stride,
                            vc_block_stride, vc_slot_stride, vc_head_stride, src_dtype, kv_dtype, k_scale, v_scale,
                            reinterpret_cast<cudaStream_t>(stream));
}

int b200attn_reshape_and_cache_flash_per_token_head(
    const void* key, const void* value, void* k_cache, void* v_cache, float* k_scale_cache, float* v_scale_cache,
    const int64_t* slot_mapping, int64_t num_tokens, int32_t num_kv_heads, int32_t head_size, int32_t block_size,
    int64_t key_stride0, int64_t key_stride1, int64_t value_stride0, int64_t value_stride1, int64_t kc_block_stride,
    int64_t kc_slot_stride, int64_t kc_head_stride, int64_t vc_block_stride, int64_t vc_slot_stride, int64_t vc_head_stride,
    int64_t ks_block_stride, int64_t ks_slot_stride, int64_t ks_head_stride, int64_t vs_block_stride, int64_t vs_slot_stride,
    int64_t vs_head_stride, int32_t src_dtype, int32_t kv_dtype, b200attn_stream_t stream) {
  PerTokenHeadWriteArgs a;
  a.src[0] = key, a.src[1] = value, a.dst[0] = k_cache, a.dst[1] = v_cache;
  a.scales[0] = k_scale_cache, a.scales[1] = v_scale_cache;
  a.slot_mapping = slot_mapping, a.num_tokens = num_tokens;
  a.src_stride0[0] = key_stride0, a.src_stride0[1] = value_stride0, a.src_stride1[0] = key_stride1, a.src_stride1[1] = value_stride1;
  a.block_stride[0] = kc_block_stride, a.block_stride[1] = vc_block_stride;
  a.slot_stride[0] = kc_slot_stride, a.slot_stride[1] = vc_slot_stride;
  a.head_stride[0] = kc_head_stride, a.head_stride[1] = vc_head_stride;
  a.s_block_stride[0] = ks_block_stride, a.s_block_stride[1] = vs_block_stride;
  a.s_slot_stride[0] = ks_slot_stride, a.s_slot_stride[1] = vs_slot_stride;
  a.s_head_stride[0] = ks_head_stride, a.s_head_stride[1] = vs_head_stride;
  a.num_kv_heads = num_kv_heads, a.head_size = head_size, a.block_size = block_size;
  a.src_dtype = src_dtype, a.kv_dtype = kv_dtype;
  return launch_cache_write_per_token_head(a, reinterpret_cast<cudaStream_t>(stream));
}

int b200attn_rope_and_cache_flash(const b200attn_rope_cache_params* p, b200attn_stream_t stream) {
  if (!p) {
    set_error("rope_and_cache_flash: null params");
    return B200ATTN_E_NULL;
  }
  return launch_rope_cache(*p, reinterpret_cast<cudaStream_t>(stream));
}

int b200attn_scaled_fp8_quant(void* out_fp8, const void* in, int64_t num_rows, int64_t row_elems,
                              int64_t in_row_stride, int32_t src_dtype, const float* scale,
                              b200attn_stream_t stream) {
  return launch_scaled_fp8_quant(out_fp8, in, num_rows, row_elems, in_row_stride, src_dtype, scale,
                                 reinterpret_cast<cudaStream_t>(stream));
}

uint64_t b200attn_workspace_bytes(int32_t num_tokens, int32_t num_seqs, int32_t num_q_heads, int32_t num_kv_heads,
                                  int32_t head_size) {
  (void)num_tokens;
  return decode_workspace_bytes(num_seqs, num_q_heads, num_kv_heads, head_size);
}

uint64_t b200attn_decode_plan_bytes(int32_t num_seqs) { return decode_plan_bytes(num_seqs); }

int b200attn_decode_plan(const b200attn_attention_params* pp, void* plan, uint64_t plan_bytes, uint64_t* key_out_host,
                         b200attn_stream_t s) {
  if (key_out_host) *key_out_host = 0;
  if (!pp || !pp->seq_lens || !pp->cu_seqlens_q) {
    set_error("decode_plan: null params / seq_lens / cu_seqlens_q");
    return B200ATTN_E_NULL;
  }
  if (pp->num_seqs <= 0 || pp->num_q_heads <= 0 || pp->num_kv_heads <= 0 || pp->num_q_heads % pp->num_kv_heads) return B200ATTN_OK;
  return build_decode_plan(*pp, plan, plan_bytes, key_out_host, reinterpret_cast<cudaStream_t>(s));
}

int b200attn_unified_attention(const b200attn_attention_params* pp, b200attn_stream_t s) {
  if (!pp) {
    set_error("unified_attention: null params");
    return B200ATTN_E_NULL;
  }
  const b200attn_attention_params& p = *pp;
  if (p.num_tokens == 0 || p.num_seqs == 0) return B200ATTN_OK;
  int rc = validate_attention(p);
  if (rc) return rc;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(s);
  switch (p.path) {
    case B200ATTN_PATH_GENERIC: return launch_attn_generic(p, stream);
    case B200ATTN_PATH_PREFILL: return launch_attn_prefill(p, stream);
    case B200ATTN_PATH_DECODE: return launch_attn_decode(p, stream);
    case B200ATTN_PATH_PREFILL_TC: return launch_attn_prefill_tc(p, stream);
    case B200ATTN_PATH_AUTO: break;
    default:
      set_error("unified_attention: unknown path %d", p.path);
      return B200ATTN_E_BAD_SHAPE;
  }
  // AUTO: reference dispatch (max_seqlen_q > 1 -> 2D else 3D,
  // triton_unified_attention.py:884); shapes a tensor-core kernel does not take
  // run the SIMT kernel (still CUDA; there is no CPU path).
  //
  // A launch that is being captured into a CUDA graph is replayed with other batch CONTENTS (and, in vLLM's mixed
  // FULL graphs, other query lengths) behind the same host ints, so under capture the set of launches depends on
  // capability and shape class only - the reference's two kernels each take any batch, and so must the set
  // recorded here: no shortcut that assumes every sequence looks like the captured ones, and grids are sized
  // from T and B alone (launch_is_captured() in the launchers).
  cudaStreamCaptureStatus cap_status = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &cap_status) != cudaSuccess) cap_status = cudaStreamCaptureStatusNone, (void)cudaGetLastError();
  g_capturing = cap_status != cudaStreamCaptureStatusNone;
  const bool captured = g_capturing;
  const char* why = nullptr;
  if (p.max_seqlen_q <= 1) {
    if (decode_supported(p, &why) && p.workspace &&
        p.workspace_bytes >= decode_workspace_bytes(p.num_seqs, p.num_q_heads, p.num_kv_heads, p.head_size))
      return launch_attn_decode(p, stream);
    if (prefill_supported(p, &why)) return launch_attn_prefill(p, stream);
    return launch_attn_generic(p, stream);
  }
  // A batch with multi-token queries.  Long queries go to the tcgen05 kernel;
  // what it leaves (short extends, decode rows of a mixed batch) goes to the
  // mma.sync kernel and, for single-token rows, to the split-KV decode kernel.
  // Which sequences are which is device data, so every launch carries a
  // query-length range and skips the others' sequences.
  if (!prefill_supported(p, &why)) return launch_attn_generic(p, stream);
  const bool ws_ok = p.workspace && p.workspace_bytes >= decode_workspace_bytes(p.num_seqs, p.num_q_heads, p.num_kv_heads, p.head_size);
  const bool dec_ok = ws_ok && decode_supported(p, &why, /*mixed_batch=*/true);
  // every sequence has exactly max_seqlen_q tokens <=> T == B * max_seqlen_q (true of THIS batch only: not used
  // when the launches are being captured)
  const bool uniform = !captured && (int64_t)p.num_tokens == (int64_t)p.num_seqs * p.max_seqlen_q;
  // The single-token rows are HBM-bound, the prompts tensor-bound: run one beside the other.  The split-KV launch
  // goes FIRST, on a second (high-priority) stream forked from this one, with its grid capped at the rows' share
  // of the step (mixed_decode_ctas); the tcgen05 grid is persistent (one CTA per free SM, the rest of its tiles
  // taken by cluster launch control), so it fills the SMs the rows leave and gets the others as the rows finish.
  // The reference covers the whole mixed batch with one launch of its 2D kernel
  // (triton_unified_attention.py:275-523); round 1 ran its three launches back to back, 0.533 ms where the
  // tensor-bound share alone takes 0.256 (C3).  Fork and join are events, so the pattern is graph-capturable.
  // tuning().reserved bit 12 keeps the launches on one stream.
  bool dec_due = dec_ok && !uniform;  // a launch for the single-token rows is still to come
  SideStream* ss = nullptr;
  int rc_side = B200ATTN_OK;
  if (dec_due && !(g_tuning.reserved & 4096) && mixed_decode_ctas(p, device_info().num_sms) < device_info().num_sms &&
      (ss = side_stream(/*may_create=*/!captured)) != nullptr) {
    if (cudaEventRecord(ss->fork, stream) == cudaSuccess && cudaStreamWaitEvent(ss->stream, ss->fork, 0) == cudaSuccess) {
      rc_side = launch_attn_decode(p, ss->stream, /*only_single_token_seqs=*/true);
      dec_due = false;
    } else {
      (void)cudaGetLastError();
      ss = nullptr;
    }
  } else {
    ss = nullptr;
  }
  // every way out of the mixed path: wait for the forked launch, then the rows' launch if it was not forked
  auto finish = [&](int rc_main) {
    if (ss) {
      cudaEventRecord(ss->join, ss->stream);
      cudaStreamWaitEvent(stream, ss->join, 0);
    }
    if (rc_main) return rc_main;
    if (rc_side) return rc_side;
    if (dec_due) return launch_attn_decode(p, stream, /*only_single_token_seqs=*/true);
    return (int)B200ATTN_OK;
  };
  if (dtype_bytes(p.kv_dtype) == 1 && !(g_tuning.reserved & 256)) {
    // FP8 cache: the sequences with long queries are dequantised once into the staging pool (the part of
    // the workspace past the decode region) and run the tcgen05 kernel on it; what does not fit the pool,
    // short extends and decode rows keep reading the FP8 cache.  (tuning().reserved bit 8 switches it off.)
    KvStage st;
    if (plan_kv_stage(p, &st)) {
      const b200attn_attention_params s16 = staged_params(p, st);
      if (prefill_tc_supported(s16, &why) && (captured || p.max_seqlen_q >= prefill_tc_min_q_len(s16))) {
        QLenRange r;
        r.min_q_len = prefill_tc_min_q_len(s16);
        rc = launch_kv_stage(p, st, r, stream);
        if (rc) return finish(rc);
        rc = launch_attn_prefill_tc(s16, stream, r, st.staged_base);
        if (rc) return finish(rc);
        QLenRange rest;  // every multi-token query that was not staged
        rest.min_q_len = dec_ok ? 2 : 0;
        // a pool that holds every sequence at full length cannot overflow: only short queries are left
        const int64_t every = (int64_t)p.num_seqs * ((p.max_seqlen_k + p.block_size - 1) / p.block_size);
        const bool fits = !captured && st.cap_pages >= every;
        if (fits) rest.max_q_len = r.min_q_len - 1;
        if (rest.max_q_len >= rest.min_q_len && !(uniform && fits)) {
          rc = launch_attn_prefill(p, stream, rest, st.staged_base);
          if (rc) return finish(rc);
        }
        return finish(B200ATTN_OK);
      }
    }
  }
  // Long queries: tcgen05.  What it leaves - short extends - runs the mma.sync kernel; like the rows they are short
  // on arithmetic, so when the rows were forked they go with them (their launch needs nothing from this stream).
  const bool tc = prefill_tc_supported(p, &why) && (captured || p.max_seqlen_q >= prefill_tc_min_q_len(p));
  if (!tc && uniform) return finish(launch_attn_prefill(p, stream));
  QLenRange r_short;
  r_short.min_q_len = dec_ok ? 2 : 0;
  r_short.max_q_len = tc ? prefill_tc_min_q_len(p) - 1 : 0x7fffffff;
  const bool short_due = r_short.max_q_len >= r_short.min_q_len && !(tc && uniform);
  if (short_due && ss) {
    rc = launch_attn_prefill(p, ss->stream, r_short);
    if (rc) return finish(rc);
  }
  if (tc) {
    QLenRange r;
    r.min_q_len = prefill_tc_min_q_len(p);
    rc = launch_attn_prefill_tc(p, stream, r);
    if (rc) return finish(rc);
  }
  if (short_due && !ss) {
    rc = launch_attn_prefill(p, stream, r_short);
    if (rc) return finish(rc);
  }
  return finish(B200ATTN_OK);
}

}  // extern "C"

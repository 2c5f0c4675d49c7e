// b200attn_decode_attention_static_scratch: the split-KV decode launch for callers that cannot allocate device
// memory (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart
// in the reference.
//
// The decode kernel needs scratch: ticket counters (zero between calls), the partial (max, exp-sum, accumulator)
// records of items split across CTAs, and - when the caller brings no plan - the step's work plan.  Through
// b200attn_unified_attention that scratch is the caller's `workspace`.  Here it is a file-scope device array of
// B200ATTN_STATIC_SCRATCH_BYTES that every GPU which loads this module holds until the process exits.
//
// In HOST code the array's name is a host-side stand-in, not a device address (a kernel handed `g_scratch` itself
// would fault, or read host memory where the GPU can reach it).  The device address comes from
// cudaGetSymbolAddress, asked once per device with that device current, and that pointer is what goes to the
// kernel (as the `workspace` of the decode launcher, which carves it exactly as it carves a caller's buffer).
//
// One array per GPU, shared by every call: it keeps the last call's contents.  The only part that has to start at
// zero is the first MiB (ticket counters and the ranges' claim words); the kernels leave it at zero when a launch
// completes, but a launch that was aborted would not, so every call clears it on its stream in front of the launch
// (cudaMemsetAsync: no allocation).  Everything behind it is rewritten before it is read.  Calls that overlap on
// the GPU (two streams, two host threads) race on the array - serialise them, or give the concurrent ones a
// workspace of their own through b200attn_unified_attention.
#include <mutex>

#include "common.cuh"

namespace b200attn {
namespace {

constexpr uint64_t kScratchBytes = B200ATTN_STATIC_SCRATCH_BYTES;
constexpr uint64_t kZeroedBytes = 1ull << 20;  // attn_decode.cu: the counter region at the start of a decode workspace
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char g_scratch[kScratchBytes];

std::mutex g_mu;
void* g_scratch_ptr[kMaxDevices];  // device address of g_scratch on each device, nullptr until asked

// Device address of this device's copy of the array (nullptr + error set on failure).
void* scratch_address() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    (void)cudaGetLastError();
    set_error("decode_attention_static_scratch: no current CUDA device (or device index >= %d)", kMaxDevices);
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_ptr[dev]) {
    void* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_scratch);  // `dev` is current: its copy of the symbol
    if (e != cudaSuccess || !ptr) {
      (void)cudaGetLastError();
      set_error("decode_attention_static_scratch: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return nullptr;
    }
    g_scratch_ptr[dev] = ptr;
  }
  return g_scratch_ptr[dev];
}

}  // namespace
}  // namespace b200attn

using namespace b200attn;

extern "C" int b200attn_decode_attention_static_scratch(const b200attn_attention_params* pp, b200attn_stream_t s) {
  if (!pp) {
    set_error("decode_attention_static_scratch: null params");
    return B200ATTN_E_NULL;
  }
  if (pp->num_tokens == 0 || pp->num_seqs == 0) return B200ATTN_OK;
  if (!pp->out || !pp->q || !pp->k_cache || !pp->v_cache || !pp->block_tables || !pp->seq_lens || !pp->cu_seqlens_q) {
    set_error("decode_attention_static_scratch: null pointer argument");
    return B200ATTN_E_NULL;
  }
  if (pp->num_seqs < 0 || pp->num_tokens < 0 || pp->num_q_heads <= 0 || pp->num_kv_heads <= 0 || pp->head_size <= 0 ||
      pp->block_size <= 0 || pp->num_q_heads % pp->num_kv_heads != 0) {
    set_error("decode_attention_static_scratch: bad shape (T=%d B=%d Hq=%d Hkv=%d D=%d bs=%d)", pp->num_tokens,
              pp->num_seqs, pp->num_q_heads, pp->num_kv_heads, pp->head_size, pp->block_size);
    return B200ATTN_E_BAD_SHAPE;
  }
  if (pp->max_seqlen_q != 1) {
    set_error("decode_attention_static_scratch: decode batches only (max_seqlen_q = %d, must be 1)", pp->max_seqlen_q);
    return B200ATTN_E_UNSUPPORTED;
  }
  const char* why = nullptr;
  if (!decode_supported(*pp, &why)) {
    set_error("decode_attention_static_scratch: the split-KV kernel does not take this configuration (%s)", why ? why : "?");
    return B200ATTN_E_UNSUPPORTED;
  }
  const uint64_t need = decode_workspace_bytes(pp->num_seqs, pp->num_q_heads, pp->num_kv_heads, pp->head_size);
  if (need > kScratchBytes) {
    set_error("decode_attention_static_scratch: this shape needs %llu bytes of scratch, the static array holds %llu "
              "(B200ATTN_STATIC_SCRATCH_BYTES); use b200attn_unified_attention with a caller-owned workspace",
              (unsigned long long)need, (unsigned long long)kScratchBytes);
    return B200ATTN_E_WORKSPACE;
  }
  void* ptr = scratch_address();
  if (!ptr) return B200ATTN_E_CUDA;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(s);
  if (cudaMemsetAsync(ptr, 0, kZeroedBytes, stream) != cudaSuccess) {
    set_error("decode_attention_static_scratch: clearing the counters failed: %s", cudaGetErrorString(cudaGetLastError()));
    return B200ATTN_E_CUDA;
  }
  b200attn_attention_params p = *pp;
  p.workspace = ptr;  // the address cudaGetSymbolAddress returned, never the symbol
  p.workspace_bytes = kScratchBytes;
  p.path = B200ATTN_PATH_DECODE;
  return launch_attn_decode(p, stream);
}

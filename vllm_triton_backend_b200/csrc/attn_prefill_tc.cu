// K3 (fast path): tcgen05 / TMEM flash-attention forward for paged KV.
// Placeholder until the tensor-core kernel lands: reports "unsupported" so the
// dispatcher uses the mma.sync kernel in attn_prefill.cu.
#include "common.cuh"

namespace b200attn {

bool prefill_tc_supported(const b200attn_attention_params&, const char** why) {
  if (why) *why = "tcgen05 prefill kernel not built";
  return false;
}

int launch_attn_prefill_tc(const b200attn_attention_params&, cudaStream_t) {
  set_error("prefill_tc: tcgen05 prefill kernel not built");
  return B200ATTN_E_UNSUPPORTED;
}

}  // namespace b200attn

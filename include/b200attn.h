/* b200attn — C ABI of the B200-native paged-KV attention path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch types, no C++
 * exceptions.  Every entry point replaces one piece of the reference
 * (foundation-model-stack/vllm-triton-backend) hot path; the reference location
 * each one stands in for is cited beside it (paths relative to the reference
 * checkout).  The reference-side binding a maintainer would add is shown in
 * INTEGRATION.md (ctypes, because the reference host code is Python).
 *
 * Conventions
 *  - all data pointers are DEVICE pointers unless the name ends in `_host`;
 *  - strides are in ELEMENTS of the tensor's dtype;
 *  - the library allocates nothing, never synchronises, never reads device
 *    data on the host and launches only on the stream it is given, so every
 *    call can be captured in a CUDA graph (grids depend on host ints only);
 *  - return value: 0 on success, a B200ATTN_E_* code otherwise;
 *    b200attn_last_error() returns a thread-local description.
 */
#ifndef B200ATTN_H_
#define B200ATTN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200ATTN_ABI_VERSION 5

#if defined(__GNUC__)
#define B200ATTN_API __attribute__((visibility("default")))
#else
#define B200ATTN_API
#endif

typedef struct CUstream_st* b200attn_stream_t; /* == cudaStream_t */

enum b200attn_dtype {
  B200ATTN_BF16 = 0,
  B200ATTN_F16 = 1,
  B200ATTN_F32 = 2,
  B200ATTN_FP8_E4M3 = 3,
  B200ATTN_INT8 = 4 /* KV cache only, and only with per-(token, head) scales (k_scale_cache / v_scale_cache) */
};

enum b200attn_status {
  B200ATTN_OK = 0,
  B200ATTN_E_BAD_DTYPE = 1,
  B200ATTN_E_BAD_SHAPE = 2,
  B200ATTN_E_ALIGNMENT = 3,
  B200ATTN_E_WORKSPACE = 4,
  B200ATTN_E_UNSUPPORTED = 5,
  B200ATTN_E_CUDA = 6,
  B200ATTN_E_NULL = 7
};

/* Which kernel the unified entry point runs.  AUTO mirrors the reference
 * dispatch `max_seqlen_q > 1 -> 2D else 3D`
 * (ibm-triton-lib/ibm_triton_lib/kernels/triton_unified_attention.py:884). */
enum b200attn_path {
  B200ATTN_PATH_AUTO = 0,
  B200ATTN_PATH_GENERIC = 1, /* SIMT kernel: any head size / group size (correctness net) */
  B200ATTN_PATH_PREFILL = 2, /* "2D": varlen / prefix / chunked / mixed batches        */
  B200ATTN_PATH_DECODE = 3,  /* "3D": split-KV GQA-packed decode + segment reduce       */
  B200ATTN_PATH_PREFILL_TC = 4 /* tcgen05/TMEM flash-attention (bf16/f16, D=128)       */
};

/* ---- K1: KV-cache scatter write -------------------------------------------
 * Replaces torch.ops._C_cache_ops.reshape_and_cache_flash as called at
 * ibm-triton-lib/ibm_triton_lib/backend/triton_attn.py:396-405
 * (kernel: vllm/csrc/cache_kernels.cu:286-358, host wrapper :700-738).
 *
 * For i in [0, num_tokens), slot = slot_mapping[i], slot >= 0:
 *   k_cache[slot / block_size][slot % block_size][h][:] = f(key[i][h][:])
 * f = copy when kv_dtype == src_dtype, or e4m3_rn_satfinite(float(x) / *scale)
 * when kv_dtype == B200ATTN_FP8_E4M3.  num_tokens is slot_mapping's length
 * (cache_kernels.cu:709-719).  k_scale / v_scale: device fp32 scalars (may be
 * NULL when not quantising).
 */
B200ATTN_API int b200attn_reshape_and_cache_flash(
    const void* key, const void* value, void* k_cache, void* v_cache,
    const int64_t* slot_mapping, int64_t num_tokens, int32_t num_kv_heads,
    int32_t head_size, int32_t block_size, int64_t key_stride0,
    int64_t value_stride0, int64_t kc_block_stride, int64_t kc_slot_stride,
    int64_t kc_head_stride, int64_t vc_block_stride, int64_t vc_slot_stride,
    int64_t vc_head_stride, int32_t src_dtype, int32_t kv_dtype,
    const float* k_scale, const float* v_scale, b200attn_stream_t stream);

/* K1 with dynamic per-(token, head) quantisation.  Replaces
 * triton_reshape_and_cache_flash_per_token_head_quant (vLLM 0.22
 * vllm/v1/attention/ops/triton_reshape_and_cache_flash.py:138-236, called at
 * vllm/v1/attention/backends/triton_attn.py:733-748), the cache write of kv_cache_dtype
 * "fp8_per_token_head" / "int8_per_token_head".  For every token i with slot >= 0 and every kv head h:
 *   s = max(absmax(key[i][h][:]) / QMAX, 1e-6)           fp32; QMAX = 448 (FP8_E4M3) or 127 (INT8)
 *   k_scale_cache[slot / bs][slot % bs][h] = s
 *   k_cache[...][h][d] = Q(clamp(float(key[i][h][d]) * (1.0f / s), QMIN, QMAX))
 * Q = round to nearest even to e4m3, or truncation toward zero to int8 (what storing an fp32 value through an
 * int8 pointer does in the upstream kernel); QMIN = -448 / -128.  The same for value.  Strides in elements
 * (scale caches: in floats); key / value rows innermost-contiguous. */
B200ATTN_API int b200attn_reshape_and_cache_flash_per_token_head(
    const void* key, const void* value, void* k_cache, void* v_cache,
    float* k_scale_cache, float* v_scale_cache,
    const int64_t* slot_mapping, int64_t num_tokens, int32_t num_kv_heads,
    int32_t head_size, int32_t block_size,
    int64_t key_stride0, int64_t key_stride1, int64_t value_stride0, int64_t value_stride1,
    int64_t kc_block_stride, int64_t kc_slot_stride, int64_t kc_head_stride,
    int64_t vc_block_stride, int64_t vc_slot_stride, int64_t vc_head_stride,
    int64_t ks_block_stride, int64_t ks_slot_stride, int64_t ks_head_stride,
    int64_t vs_block_stride, int64_t vs_slot_stride, int64_t vs_head_stride,
    int32_t src_dtype, int32_t kv_dtype, b200attn_stream_t stream);

/* ---- K2/K3: paged attention -----------------------------------------------
 * Replaces `unified_attention` (triton_unified_attention.py:839-1030): the 2D
 * kernel (:275-523), the 3D split-KV kernel (:526-754) and `reduce_segments`
 * (:757-836).  Field meanings follow the reference launcher's arguments.
 */
typedef struct b200attn_attention_params {
  void* out;                   /* [T, Hq, D]  q_dtype (FP8-e4m3 when out_scale != NULL) */
  const void* q;               /* [T, Hq, D]  innermost contiguous                 */
  const void* k_cache;         /* [NB, bs, Hkv, D] view, innermost contiguous      */
  const void* v_cache;
  const int32_t* block_tables; /* [B, >=ceil(max_seqlen_k/bs)]                     */
  const int32_t* seq_lens;     /* [B]  context + new tokens ("seqused_k")          */
  const int32_t* cu_seqlens_q; /* [B+1]                                            */
  const float* alibi_slopes;   /* [Hq] or NULL                                     */
  const float* k_scale;        /* device scalar ("k_descale" element 0) or NULL=1  */
  const float* v_scale;
  const float* sinks;          /* [Hq] fp32 attention-sink logits or NULL: one extra
                                  key per head that only joins the softmax denominator
                                  (upstream addition to the same kernel, `init_softmax_M`
                                  in vllm/v1/attention/ops/triton_attention_helpers.py) */
  float softmax_scale;
  float softcap;               /* 0 = off                                          */
  int32_t sliding_window;      /* 0 = off; = 1 + window_size[0] (:915)             */
  int32_t num_tokens;          /* T                                                */
  int32_t num_seqs;            /* B                                                */
  int32_t num_q_heads;
  int32_t num_kv_heads;
  int32_t head_size;
  int32_t block_size;
  int32_t max_seqlen_q;
  int32_t max_seqlen_k;
  int32_t q_dtype;             /* enum b200attn_dtype: BF16 / F16 (F32: generic)   */
  int32_t kv_dtype;            /* == q_dtype, or FP8_E4M3; INT8 with scale caches  */
  int32_t path;                /* enum b200attn_path                               */
  int32_t reserved0;
  int64_t num_blocks;          /* NB (extent of the cache views)                   */
  int64_t q_stride0, q_stride1;
  int64_t out_stride0, out_stride1;
  int64_t bt_stride0;
  int64_t kc_block_stride, kc_slot_stride, kc_head_stride;
  int64_t vc_block_stride, vc_slot_stride, vc_head_stride;
  void* workspace;             /* split-KV partials + ticket counters; caller-owned,
                                  zero-filled ONCE when allocated.  The first 1 MiB
                                  holds the counters (the kernels leave them at zero
                                  after every call) whatever the shape of the call, so
                                  one buffer serves batches of every size; everything
                                  behind it is scratch that is rewritten before use.
                                  One workspace per stream with calls in flight.    */
  uint64_t workspace_bytes;
  const float* out_scale;      /* NULL: `out` has q_dtype.  Else a device scalar ("output_scale" of
                                  AttentionImpl.forward when fused_output_quant_supported, vLLM 0.22
                                  vllm/v1/attention/backends/triton_attn.py:453-454,663; upstream kernel
                                  `acc * (1/output_scale)`, clamp +-448, store e4m3): `out` is then
                                  FP8-e4m3 [T, Hq, D], out_stride* in 1-byte elements, rows 8-byte
                                  aligned, and every kernel of the path quantises in its epilogue:
                                  out = e4m3_rn_satfinite(o_fp32 * (1.0f / *out_scale))               */
  const float* qq_bias;        /* NULL, or a [q_len_max, qq_bias_stride0] fp32 matrix (upstream `qq_bias`, tree /
                                  speculative-decoding masks; vllm/v1/attention/ops/triton_unified_attention.py:376,
                                  526-528, triton_attention_helpers.py:344-357): after masking and ALiBi,
                                  S[query i of a sequence][key j] += qq_bias[i][j - context_len] for
                                  0 <= j - context_len < qq_bias_stride0.  Every kernel applies it.               */
  int64_t qq_bias_stride0;
  float* lse;                  /* NULL, or [T, lse_stride0 >= Hq] fp32: every kernel also writes the natural-log sum of
                                  exponentials of each (token, head) row's logits (sink included; -inf when no key is
                                  visible, sink or not) - what a caller needs to combine partial attention over disjoint key sets
                                  (decode context parallelism: AttentionImpl.can_return_lse_for_decode,
                                  vllm/vllm/attention/backends/abstract.py:130-153).  The kernels hold (max, exp-sum)
                                  per row in their epilogues; the reference's own kernels do not export it.        */
  int64_t lse_stride0;
  const void* decode_plan;     /* NULL, or a plan written by b200attn_decode_plan for THIS step's seq_lens /
                                  cu_seqlens_q (all layers of a step share it; the metadata builder makes it once).
                                  Without one the split-KV launch plans for itself (one more small launch).   */
  uint64_t decode_plan_key;    /* the key b200attn_decode_plan returned; a plan whose key does not match the
                                  launch's own geometry (other head counts, window, ...) is ignored            */
  const float* k_scale_cache;  /* both NULL, or fp32 views [NB, bs, Hkv] (strides below, in floats): one scale per cached
                                  (token, kv head), written by b200attn_reshape_and_cache_flash_per_token_head - upstream's
                                  KVQuantMode.FP8_PER_TOKEN_HEAD / INT8_PER_TOKEN_HEAD (vLLM 0.22
                                  vllm/v1/attention/ops/triton_unified_attention.py:469-489, 505-508, 539-544):
                                  K and V are cast to the query dtype as they are, the logit of key j is
                                  dot(q, K_j) * (softmax_scale * k_scale_cache[j]) and P_ij * v_scale_cache[j], rounded to
                                  the query dtype, multiplies V_j.  kv_dtype is FP8_E4M3 or INT8; k_scale / v_scale are
                                  not used.                                                                     */
  const float* v_scale_cache;
  int64_t ks_block_stride, ks_slot_stride, ks_head_stride;
  int64_t vs_block_stride, vs_slot_stride, vs_head_stride;
} b200attn_attention_params;

/* Bytes of workspace b200attn_unified_attention may need for this shape
 * (replaces the per-call torch.empty of segm_output/segm_max/segm_expsum,
 * triton_unified_attention.py:950-971).  Host ints only. */
B200ATTN_API uint64_t b200attn_workspace_bytes(int32_t num_tokens, int32_t num_seqs,
                                  int32_t num_q_heads, int32_t num_kv_heads,
                                  int32_t head_size);

B200ATTN_API int b200attn_unified_attention(const b200attn_attention_params* p,
                               b200attn_stream_t stream);

/* Work plan of the split-KV decode launches of one step.  The reference cuts every sequence into 16 fixed
 * segments on the host side of the launch (triton_unified_attention.py:931-948); here the (sequence, kv-head
 * group, KV step) space is cut on the device from seq_lens: one static range per CTA plus a pool of chunks
 * the CTAs share (csrc/attn_decode.cu).  That cut depends on seq_lens / cu_seqlens_q only, not on the layer,
 * so the metadata builder computes it once per step (replaces nothing in the reference: its builder computes
 * nothing of the kind, triton_attn.py:130-227) and passes it in `decode_plan`.
 *   p:          the step's parameters (seq_lens, cu_seqlens_q, num_seqs, head counts, head_size, block_size,
 *               dtypes, sliding_window, max_seqlen_q; data pointers other than the two metadata arrays unused)
 *   plan:       device buffer of b200attn_decode_plan_bytes(num_seqs) bytes, 16-byte aligned, caller-owned
 *   key_out:    host: the key to put in decode_plan_key; 0 = this shape does not use the split-KV kernel */
B200ATTN_API uint64_t b200attn_decode_plan_bytes(int32_t num_seqs);
B200ATTN_API int b200attn_decode_plan(const b200attn_attention_params* p, void* plan, uint64_t plan_bytes,
                                      uint64_t* key_out_host, b200attn_stream_t stream);

/* Decode attention for callers that cannot allocate device memory (code under CUDA-graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arithmetic and kernel as b200attn_unified_attention with path = B200ATTN_PATH_DECODE, but the
 * split-KV scratch (ticket counters, partial records, the work plan when `decode_plan` is not given) lives in a
 * file-scope device array of B200ATTN_STATIC_SCRATCH_BYTES inside the library - every GPU that loads the library
 * holds it until the process exits - instead of the caller's `workspace` (`workspace`, `workspace_bytes` and `path`
 * of `p` are ignored).  The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine): it
 * clears the first MiB of the array on `stream` and launches.
 *   - decode batches only: max_seqlen_q == 1 and a configuration the split-KV kernel takes
 *     (B200ATTN_E_UNSUPPORTED otherwise);
 *   - limit: b200attn_workspace_bytes(...) of the shape must not exceed B200ATTN_STATIC_SCRATCH_BYTES (24 MiB:
 *     e.g. up to 32 query heads at head size 128 with 8 kv heads, or a 1-2 kv-head shard of a 64-head model);
 *     larger shapes return B200ATTN_E_WORKSPACE with a message naming both sizes;
 *   - the array is shared by all calls on a GPU and keeps the last call's contents; calls that overlap on the GPU
 *     (other streams, other threads) race on it: serialise them or use b200attn_unified_attention with a
 *     workspace per stream. */
#define B200ATTN_STATIC_SCRATCH_BYTES (24u << 20)
B200ATTN_API int b200attn_decode_attention_static_scratch(const b200attn_attention_params* p,
                                                          b200attn_stream_t stream);

/* ---- K4 (optional): static per-tensor FP8 quantisation of Q ----------------
 * Replaces ops.scaled_fp8_quant(query, layer._q_scale) at triton_attn.py:417
 * (vllm/csrc/quantization/w8a8/fp8/common.cu:11-29):
 *   out[i] = e4m3_rn_satfinite(float(in[i]) * (1 / *scale)).
 */
B200ATTN_API int b200attn_scaled_fp8_quant(void* out_fp8, const void* in, int64_t num_rows,
                              int64_t row_elems, int64_t in_row_stride,
                              int32_t src_dtype, const float* scale,
                              b200attn_stream_t stream);

/* ---- K5: rotary embedding of q and k (in place) + KV-cache write, one launch ----
 * Replaces the two launches vLLM issues before attention:
 * rotary_embedding (vllm/csrc/pos_encoding_kernels.cu) and reshape_and_cache_flash
 * (vllm/csrc/cache_kernels.cu:286-358), behind
 * AttentionImpl.do_rope_and_kv_cache_update (vLLM 0.22, vllm/v1/attention/backend.py:823-841).
 *   for t < rope_tokens, every q and k head, pair (x, y) of the first rot_dim dims
 *   (neox: (j, j + rot_dim/2); else (2j, 2j+1)), c = cos_sin_cache[positions[t]][j],
 *   s = cos_sin_cache[positions[t]][rot_dim/2 + j]:
 *       x' = R(x c - y s),  y' = R(y c + x s)    fp32 arithmetic, R = one rounding to the 16-bit type
 *       (what vLLM 0.22's compiled rotary_embedding produces, bit for bit)
 *   for t < cache_tokens with slot_mapping[t] >= 0: the rotated k row and the v row go to
 *   the cache exactly as b200attn_reshape_and_cache_flash writes them.
 * Strides in elements; rows must be 16-byte aligned (B200ATTN_E_ALIGNMENT otherwise). */
typedef struct b200attn_rope_cache_params {
  void* q;                      /* [T, Hq, D]  rotated in place                     */
  void* k;                      /* [T, Hkv, D] rotated in place                     */
  const void* v;                /* [T, Hkv, D]                                      */
  const int64_t* positions;     /* [rope_tokens]                                    */
  const void* cos_sin_cache;    /* [max_position, rot_dim], dtype of q              */
  void* k_cache;                /* [NB, bs, Hkv, D] views                           */
  void* v_cache;
  const int64_t* slot_mapping;  /* [cache_tokens], -1 = padding                     */
  const float* k_scale;         /* device scalars, fp8 cache only                   */
  const float* v_scale;
  int64_t rope_tokens, cache_tokens;
  int64_t q_stride0, q_stride1, k_stride0, k_stride1, v_stride0, v_stride1;
  int64_t kc_block_stride, kc_slot_stride, kc_head_stride;
  int64_t vc_block_stride, vc_slot_stride, vc_head_stride;
  int32_t num_q_heads, num_kv_heads, head_size, rot_dim, block_size;
  int32_t is_neox;              /* 1: GPT-NeoX halves, 0: GPT-J interleaved pairs   */
  int32_t src_dtype;            /* BF16 / F16                                       */
  int32_t kv_dtype;             /* == src_dtype or FP8_E4M3                         */
} b200attn_rope_cache_params;

B200ATTN_API int b200attn_rope_and_cache_flash(const b200attn_rope_cache_params* p, b200attn_stream_t stream);

/* ---- introspection ---------------------------------------------------------*/
B200ATTN_API int b200attn_abi_version(void);
/* sizeof the parameter blocks as this library was compiled (0: b200attn_attention_params,
 * 1: b200attn_rope_cache_params; else 0): a binding checks its own struct layout against it. */
B200ATTN_API uint64_t b200attn_params_size(int32_t which);
B200ATTN_API const char* b200attn_last_error(void);
/* Number of kernels this library launched on the calling thread since the last
 * reset (bench.py reports it as `gpu_launches`). */
B200ATTN_API uint64_t b200attn_launch_count(void);
B200ATTN_API void b200attn_reset_launch_count(void);
/* Name of the kernel variant the last attention call on this thread ran. */
B200ATTN_API const char* b200attn_last_kernel(void);
/* Tuning knobs (process-wide; 0 = library default).  Used by the benchmark
 * sweeps; not needed for correctness.  decode_pool_sixteenths: 1..12 lets idle CTAs of the split-KV
 * decode kernel steal that share (in 1/16) of every other CTA's range (default: off, static split,
 * bit-reproducible results). */
B200ATTN_API void b200attn_set_tuning(int32_t decode_heads_per_cta, int32_t decode_pool_sixteenths,
                         int32_t decode_stages, int32_t reserved);

/* Debug: when `device_buffer` is not NULL the split-KV decode kernel writes 16 uint64
 * per CTA (globaltimer at entry, then %clock64 at: entry, first run known, first KV
 * stage landed, last stage consumed, exit; globaltimer at exit; runs finalised; then cycle
 * sums: consumers in run epilogues, consumers waiting for data, producer waiting for a free
 * stage, producer between ranges; 4 unused).
 * Process-wide; NULL switches it off.  Used by scripts/kbench.py decode --trace. */
B200ATTN_API void b200attn_debug_set_decode_trace(void* device_buffer);

#ifdef __cplusplus
}
#endif
#endif /* B200ATTN_H_ */

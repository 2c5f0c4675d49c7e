"""vLLM V1 attention backend over the b200attn CUDA library.

Host-side mirror of the reference plugin's backend module
(ibm-triton-lib/ibm_triton_lib/backend/triton_attn.py): same three classes —
metadata + builder (:60-233), backend (:236-285), impl (:288-470) — with the same
argument meanings and error behaviour, written against the interface of the
vLLM that is installed here (0.22: `vllm.v1.attention.backend`, split
`do_kv_cache_update`, builder `(kv_cache_spec, layer_names, vllm_config, device)`).

Differences from the reference, all deliberate:
  * the cache write and the attention run in libb200attn.so (no Triton, no
    `_C_cache_ops`); a missing library is an ImportError, never a fallback;
  * split-KV workspaces are allocated once by the builder instead of per call
    (reference kernels/triton_unified_attention.py:950-971);
  * Q stays bf16/fp16 with an FP8 KV cache (the reference's fp8-Q branch drops
    k_scale/v_scale), so `supports_quant_query_input=False`.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import ClassVar, Optional

import torch

from vllm.v1.attention.backend import (
    AttentionBackend,
    AttentionCGSupport,
    AttentionImpl,
    AttentionMetadataBuilder,
    AttentionType,
    CommonAttentionMetadata,
    MultipleOf,
)

from .. import ops

# vLLM resolves `AttentionBackendEnum[backend.get_name()]` when it builds each Attention layer
# (site-packages/vllm/model_executor/layers/attention/attention.py:386), so an out-of-tree
# backend has to carry the name of the enum member it is registered under.
BACKEND_NAME = "CUSTOM"
BACKEND_PATH = "vllm_triton_backend_b200.backend.attn.B200AttentionBackend"


@dataclass
class B200AttentionMetadata:
    # |---------- context_len ----------|-- query_len ---|   (seq_len = both)
    num_actual_tokens: int  # tokens excluding padding
    max_query_len: int
    query_start_loc: torch.Tensor  # int32 [B+1] device
    max_seq_len: int
    seq_lens: torch.Tensor  # int32 [B] device
    block_table: torch.Tensor  # int32 [B, MB] device
    slot_mapping: torch.Tensor  # int64 [T] device, padding = -1
    workspace: torch.Tensor  # split-KV partials, allocated once by the builder
    decode_plan: Optional["ops.DecodePlan"] = None  # this step's split-KV work plan, built once for all layers

    # cascade attention is never used (reference triton_attn.py:279-281)
    use_cascade: bool = False
    common_prefix_len: int = 0
    cu_prefix_query_lens: Optional[torch.Tensor] = None
    prefix_kv_lens: Optional[torch.Tensor] = None
    suffix_kv_lens: Optional[torch.Tensor] = None


class B200AttentionMetadataBuilder(AttentionMetadataBuilder[B200AttentionMetadata]):
    # full CUDA graphs: grids depend on host ints only (reference :107, :229-233)
    _cudagraph_support: ClassVar[AttentionCGSupport] = AttentionCGSupport.ALWAYS

    def __init__(self, kv_cache_spec, layer_names, vllm_config, device):
        super().__init__(kv_cache_spec, layer_names, vllm_config, device)
        self.block_size = kv_cache_spec.block_size
        model_config = vllm_config.model_config
        parallel_config = vllm_config.parallel_config
        self.num_heads_q = model_config.get_num_attention_heads(parallel_config)
        self.num_heads_kv = model_config.get_num_kv_heads(parallel_config)
        self.headdim = model_config.get_head_size()
        max_num_seqs = vllm_config.scheduler_config.max_num_seqs
        self._ws_seqs = max_num_seqs
        nbytes = ops.workspace_bytes(max_num_seqs, max_num_seqs, self.num_heads_q, self.num_heads_kv, self.headdim)
        cache_dtype = str(getattr(vllm_config.cache_config, "cache_dtype", "auto"))
        if cache_dtype.startswith("fp8") or cache_dtype in PER_TOKEN_HEAD_DTYPES:
            # FP8 cache: staging pool that lets long prefills run the tcgen05 kernel (csrc/kv_stage.cu).  It sits
            # outside vLLM's KV-cache budget (gpu_memory_utilization does not see it), so it is bounded: one
            # sequence of max_model_len, at most B200ATTN_FP8_STAGING_MIB (default 256) MiB; what does not fit
            # runs the mma.sync kernel.  B200ATTN_FP8_STAGING_TOKENS overrides the token count (0 = no staging).
            import os

            per_token = 2 * self.num_heads_kv * self.headdim * 2  # K + V, 16-bit
            cap_tokens = (int(os.environ.get("B200ATTN_FP8_STAGING_MIB", 256)) << 20) // per_token
            tokens = int(os.environ.get("B200ATTN_FP8_STAGING_TOKENS", min(int(model_config.max_model_len), cap_tokens)))
            stage = ops.fp8_staging_bytes(max_num_seqs, tokens, self.block_size, self.num_heads_kv, self.headdim)
            nbytes += stage
            import logging

            logging.getLogger("vllm_triton_backend_b200").info(
                "b200attn: FP8 prefill staging pool %.1f MiB (%d tokens), outside the KV-cache budget", stage / 2**20, tokens)
        # zeroed once: holds the decode kernel's self-cleaning ticket counters
        self.workspace = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        self._ws_extra = nbytes - ops.workspace_bytes(max_num_seqs, max_num_seqs, self.num_heads_q, self.num_heads_kv,
                                                      self.headdim)
        # The split-KV work plan depends on seq_lens only, and all layers of a step share them: one plan kernel per
        # step here instead of a partition inside every layer's launch.  A layer whose geometry differs from what
        # is assumed here (another sliding window) ignores it and plans for itself (the key does not match).
        q_dtype = model_config.dtype
        kv_dtype = torch.float8_e4m3fn if cache_dtype.startswith("fp8") else q_dtype
        if cache_dtype in PER_TOKEN_HEAD_DTYPES:
            kv_dtype = PER_TOKEN_HEAD_DTYPES[cache_dtype]
        self.decode_plan = None
        if q_dtype in (torch.bfloat16, torch.float16):
            self.decode_plan = ops.DecodePlan(device, max_num_seqs, self.num_heads_q, self.num_heads_kv, self.headdim,
                                              self.block_size, q_dtype, kv_dtype,
                                              int(getattr(kv_cache_spec, "sliding_window", 0) or 0))

    def _fit_workspace(self, num_seqs: int) -> None:
        """Chunked local attention (vLLM wraps this builder and hands it virtual batches,
        `make_local_attention_virtual_batches`) can present more sequences than max_num_seqs; the split-KV
        records are sized per sequence, so grow once (eager steps only: captured sizes never exceed
        max_num_seqs)."""
        if num_seqs <= self._ws_seqs:
            return
        self._ws_seqs = max(num_seqs, 2 * self._ws_seqs)
        nbytes = ops.workspace_bytes(self._ws_seqs, self._ws_seqs, self.num_heads_q, self.num_heads_kv, self.headdim)
        # captured CUDA graphs hold the old buffer's address: keep it alive
        self._retired_workspaces = getattr(self, "_retired_workspaces", []) + [self.workspace]
        self.workspace = torch.zeros(nbytes + self._ws_extra, dtype=torch.uint8, device=self.workspace.device)

    def build_for_cudagraph_capture(self, common_attn_metadata: CommonAttentionMetadata) -> B200AttentionMetadata:
        attn_metadata = self.build(0, common_attn_metadata)
        # capture with seq_lens = 1 so it is fast (reference :120-128); the captured
        # grid does not depend on the lengths.
        attn_metadata.seq_lens.fill_(1)
        if self.decode_plan is not None:  # the plan follows seq_lens
            self.decode_plan.build(attn_metadata.query_start_loc, attn_metadata.seq_lens, attn_metadata.max_query_len,
                                   attn_metadata.num_actual_tokens)
        return attn_metadata

    def build(self, common_prefix_len: int, common_attn_metadata: CommonAttentionMetadata,
              fast_build: bool = False) -> B200AttentionMetadata:
        m = common_attn_metadata
        self._fit_workspace(m.seq_lens.shape[0])
        if self.decode_plan is not None:
            self.decode_plan.build(m.query_start_loc, m.seq_lens, m.max_query_len, m.num_actual_tokens)
        return B200AttentionMetadata(
            num_actual_tokens=m.num_actual_tokens,
            max_query_len=m.max_query_len,
            query_start_loc=m.query_start_loc,
            max_seq_len=m.max_seq_len,
            seq_lens=m.seq_lens,
            block_table=m.block_table_tensor,
            slot_mapping=m.slot_mapping,
            workspace=self.workspace,
            decode_plan=self.decode_plan,
            use_cascade=False,
            common_prefix_len=common_prefix_len,
        )


class B200AttentionBackend(AttentionBackend):
    accept_output_buffer: bool = True
    supported_dtypes: ClassVar[list] = [torch.float16, torch.bfloat16]
    supported_kv_cache_dtypes: ClassVar[list] = ["auto", "float16", "bfloat16", "fp8", "fp8_e4m3", "fp8_per_token_head",
                                                 "int8_per_token_head"]
    # vLLM >= 0.11 calls do_kv_cache_update before forward
    forward_includes_kv_cache_update: bool = False

    @staticmethod
    def get_supported_kernel_block_sizes():
        return [MultipleOf(16)]

    @classmethod
    def supports_block_size(cls, block_size) -> bool:
        return block_size is None or block_size % 16 == 0

    @classmethod
    def get_supported_head_sizes(cls) -> list[int]:
        return [32, 64, 96, 128, 160, 192, 224, 256]  # reference :241-242

    @classmethod
    def validate_head_size(cls, head_size: int) -> None:
        supported = cls.get_supported_head_sizes()
        if head_size not in supported:
            raise ValueError(  # reference :245-254
                f"Head size {head_size} is not supported by {cls.__name__.removesuffix('Backend')}. "
                f"Supported head sizes are: {supported}."
            )

    @staticmethod
    def get_name() -> str:
        return BACKEND_NAME

    @staticmethod
    def get_impl_cls():
        return B200AttentionImpl

    @staticmethod
    def get_metadata_cls():
        return B200AttentionMetadata

    @staticmethod
    def get_builder_cls():
        return B200AttentionMetadataBuilder

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int,
                           cache_dtype_str: str = "auto") -> tuple[int, ...]:
        if block_size % 16 != 0:
            raise ValueError("Block size must be a multiple of 16.")  # reference :275-276
        # K and V planes are separate: each (page, kv-head-group) tile the decode
        # kernel streams is one contiguous TMA box, and K/V of a page do not share
        # DRAM rows.  (The reference plugin used this layout too, :277.)
        if cache_dtype_str in PER_TOKEN_HEAD_DTYPES:
            # the bytes vLLM budgets for this mode: every (token, head) row plus one fp32 scale
            # (AttentionSpec.page_size_bytes; upstream's backend declares the same padded shape,
            # vllm/v1/attention/backends/triton_attn.py:319-333).  Inside a block the data comes first and the
            # scales behind it (ops.per_token_head_cache_views), not inline behind every row.
            return (2, num_blocks, block_size, num_kv_heads, head_size + ops.PER_TOKEN_HEAD_SCALE_PAD_BYTES)
        return (2, num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def get_kv_cache_stride_order(include_num_layers_dimension: bool = False) -> tuple[int, ...]:
        """Permutation from `get_kv_cache_shape` to the memory order (vLLM's VLLM_KV_CACHE_LAYOUT /
        KV-connector preference; same contract as flash_attn.py:152-170).  The kernels take the cache as
        strided [NB, bs, Hkv, D] views, so NHD and HND are both just strides."""
        from vllm.v1.attention.backends.utils import get_kv_cache_layout

        layout = get_kv_cache_layout()
        if layout not in ("NHD", "HND"):
            raise ValueError(f"Unknown cache layout format {layout}.")
        if include_num_layers_dimension:
            # physical (num_blocks, [num_kv_heads,] num_layers, 2, block_size, [num_kv_heads,] head_size)
            return (2, 0, 1, 3, 4, 5) if layout == "NHD" else (2, 4, 0, 1, 3, 5)
        return (0, 1, 2, 3, 4) if layout == "NHD" else (0, 1, 3, 2, 4)

    @staticmethod
    def use_cascade_attention(*args, **kwargs) -> bool:
        return False

    @classmethod
    def supports_attn_type(cls, attn_type: str) -> bool:
        return attn_type == AttentionType.DECODER

    @classmethod
    def supports_compute_capability(cls, capability) -> bool:
        return capability.major == 10  # sm_100a cubins only (they do not load on sm_12x)

    @classmethod
    def supports_sink(cls) -> bool:
        return True  # every kernel applies the sink in its epilogue (decode: direct epilogue or the merge)


# kv_cache_dtype strings of upstream's dynamically quantised caches (vllm/v1/kv_cache_interface.py:57-64) -> element type
PER_TOKEN_HEAD_DTYPES = {"fp8_per_token_head": torch.float8_e4m3fn, "int8_per_token_head": torch.int8}


def split_kv_cache(kv_cache: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    """K/V views [NB, bs, Hkv, D] of either allocation vLLM may hand us:
    (2, NB, bs, Hkv, D) — ours / the reference plugin's — or (NB, 2, bs, Hkv, D)
    (stock vLLM >= 0.11, block stride 2*bs*Hkv*D)."""
    if kv_cache.dim() == 5 and kv_cache.shape[0] == 2 and kv_cache.shape[1] == 2:
        # NB == 2 makes the two allocations look alike; this backend declares (2, NB, ...) (get_kv_cache_shape),
        # so that is what a cache vLLM allocated for it is
        return kv_cache.unbind(0)
    if kv_cache.dim() == 5 and kv_cache.shape[1] == 2:
        return kv_cache.unbind(1)
    return kv_cache.unbind(0)


class B200AttentionImpl(AttentionImpl):
    # Every kernel can write the softmax log-sum-exp of its rows (`softmax_lse=` of ops.unified_attention / of the
    # per-layer plan; vllm/attention/backends/abstract.py:130-153).  The collective that decode context parallelism
    # builds on it is outside this path: forward() refuses to run with dcp_world_size > 1 rather than return
    # un-combined partial attention.
    can_return_lse_for_decode: bool = True

    def __init__(
        self,
        num_heads: int,
        head_size: int,
        scale: float,
        num_kv_heads: Optional[int] = None,
        alibi_slopes: Optional[list] = None,
        sliding_window: Optional[int] = None,
        kv_cache_dtype: str = "auto",
        logits_soft_cap: Optional[float] = None,
        attn_type: str = AttentionType.DECODER,
        kv_sharing_target_layer_name: Optional[str] = None,
        blocksparse_params=None,
        sinks: Optional[torch.Tensor] = None,
        use_alibi_sqrt: bool = False,  # vLLM 0.22 impl arguments that change the result and are not built here:
        chunk_lookback: int = -1,      # a non-default value raises (upstream triton_attn.py TritonAttentionImpl.__init__)
        **extra_impl_args,
    ) -> None:
        if use_alibi_sqrt:
            raise NotImplementedError("B200AttentionImpl: use_alibi_sqrt is not implemented")
        if chunk_lookback is not None and int(chunk_lookback) >= 0:
            raise NotImplementedError("B200AttentionImpl: chunk_lookback is not implemented")
        if extra_impl_args:
            # an argument this backend does not know may change what the layer computes: refuse, never ignore
            raise NotImplementedError(f"B200AttentionImpl: unsupported impl arguments {sorted(extra_impl_args)}")
        if blocksparse_params is not None:
            raise ValueError("B200Attention does not support block-sparse attention.")  # reference :305-306
        if sinks is not None and sinks.shape[0] != num_heads:
            raise ValueError(  # same check as upstream's launcher (triton_unified_attention.py: "Sinks must be ...")
                f"Sinks must have the same number of heads as the number of heads in the layer. "
                f"Expected {num_heads}, but got {sinks.shape[0]}."
            )
        self.sinks = sinks  # converted to fp32 when the plan is made (weights are loaded by then)
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        self.num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        self._alibi_slopes_host = None if alibi_slopes is None else torch.tensor(alibi_slopes, dtype=torch.float32)
        self.alibi_slopes = None  # moved to the device on first use
        self.sliding_window = (-1, -1) if sliding_window is None else (sliding_window - 1, 0)  # reference :314-317
        self.kv_cache_dtype = kv_cache_dtype
        self.logits_soft_cap = 0 if logits_soft_cap is None else logits_soft_cap  # 0 = off (:319-322)
        self.kv_sharing_target_layer_name = kv_sharing_target_layer_name
        self.num_queries_per_kv = self.num_heads // self.num_kv_heads
        B200AttentionBackend.validate_head_size(head_size)
        if attn_type != AttentionType.DECODER:
            raise NotImplementedError(  # reference :331-337
                "Encoder self-attention and encoder/decoder cross-attention are not implemented for B200AttentionImpl"
            )
        self.attn_type = attn_type
        self.supports_quant_query_input = False
        self._plans: dict = {}
        self._per_token_head = kv_cache_dtype in PER_TOKEN_HEAD_DTYPES
        self._pth_filled: set = set()
        if kv_cache_dtype.startswith("fp8") and kv_cache_dtype not in ("fp8", "fp8_e4m3") and not self._per_token_head:
            raise NotImplementedError(f"kv_cache_dtype {kv_cache_dtype!r}: only fp8_e4m3 is implemented")

    def _views(self, kv_cache: torch.Tensor):
        key_cache, value_cache = split_kv_cache(kv_cache)
        if self._per_token_head:
            raise RuntimeError("per-token-head caches are read through _per_token_head_views")
        if self.kv_cache_dtype.startswith("fp8") and key_cache.dtype != torch.float8_e4m3fn:
            key_cache = key_cache.view(torch.float8_e4m3fn)  # reference :407-409
            value_cache = value_cache.view(torch.float8_e4m3fn)
        return key_cache, value_cache

    def _per_token_head_views(self, kv_cache: torch.Tensor):
        """(K data, V data, K scales, V scales) of a cache allocated with the padded shape of get_kv_cache_shape."""
        k_plane, v_plane = split_kv_cache(kv_cache)
        dt = PER_TOKEN_HEAD_DTYPES[self.kv_cache_dtype]
        key_cache, k_scales = ops.per_token_head_cache_views(k_plane, self.head_size, dt)
        value_cache, v_scales = ops.per_token_head_cache_views(v_plane, self.head_size, dt)
        return key_cache, value_cache, k_scales, v_scales

    def _plans_for(self, layer, kv_cache: torch.Tensor, q_dtype: torch.dtype):
        """(cache-write plan, attention plan) of this layer's cache tensor: the argument
        lists of the two C calls with everything that does not change from step to step
        filled in once (ops.AttentionPlan / ops.CacheWritePlan)."""
        key = (kv_cache.data_ptr(), q_dtype)
        plans = self._plans.get(key)
        if plans is None and self._per_token_head:
            if self._alibi_slopes_host is not None and self.alibi_slopes is None:
                self.alibi_slopes = self._alibi_slopes_host.to(kv_cache.device)
            key_cache, value_cache, k_scales, v_scales = self._per_token_head_views(kv_cache)
            # Upstream fills its scale views with 1.0 when it first sees a cache (triton_attn.py:441,452).  The kernels
            # here never read the scale of a slot that has not been written (they stop at seq_len), so this is
            # hygiene only - and done once per allocation: a re-made plan must not wipe live scales.
            if kv_cache.data_ptr() not in self._pth_filled:
                self._pth_filled.add(kv_cache.data_ptr())
                k_scales.fill_(1.0)
                v_scales.fill_(1.0)
            plans = (
                ops.PerTokenHeadCacheWritePlan(key_cache, value_cache, k_scales, v_scales, q_dtype),
                ops.AttentionPlan(key_cache, value_cache, q_dtype, self.num_heads, self.head_size, self.scale,
                                  self.sliding_window, self.logits_soft_cap, self.alibi_slopes, None, None,
                                  sinks=self.sinks, k_scale_cache=k_scales, v_scale_cache=v_scales),
            )
            self._plans[key] = plans
        if plans is None:
            key_cache, value_cache = self._views(kv_cache)
            if self._alibi_slopes_host is not None and self.alibi_slopes is None:
                self.alibi_slopes = self._alibi_slopes_host.to(kv_cache.device)
            fp8 = self.kv_cache_dtype.startswith("fp8")
            plans = (
                ops.CacheWritePlan(key_cache, value_cache, self.kv_cache_dtype, q_dtype, layer._k_scale, layer._v_scale),
                ops.AttentionPlan(key_cache, value_cache, q_dtype, self.num_heads, self.head_size, self.scale,
                                  self.sliding_window, self.logits_soft_cap, self.alibi_slopes,
                                  layer._k_scale if fp8 else None, layer._v_scale if fp8 else None,  # element 0 is read (:438)
                                  sinks=self.sinks),
            )
            if len(self._plans) > 8:  # a layer has one cache tensor; bound the dict if a caller keeps reallocating
                self._plans.clear()
            self._plans[key] = plans
        return plans

    def do_kv_cache_update(self, layer, key, value, kv_cache, slot_mapping) -> None:
        """Cache write, split out of forward by vLLM >= 0.11 (reference did it
        inside forward, triton_attn.py:393-405)."""
        if self.kv_sharing_target_layer_name is not None:
            return
        self._plans_for(layer, kv_cache, key.dtype)[0](key, value, slot_mapping)

    def fused_output_quant_supported(self, quant_key) -> bool:
        """Static per-tensor symmetric FP8 of the output, done in every kernel's epilogue
        (vLLM 0.22 vllm/v1/attention/backends/triton_attn.py:453-454)."""
        try:
            from vllm.model_executor.layers.quantization.utils.quant_utils import kFp8StaticTensorSym
        except Exception:  # pragma: no cover - depends on the installed vLLM
            return False
        return quant_key == kFp8StaticTensorSym

    def fused_rope_kvcache_supported(self) -> bool:
        """RoPE + cache write in one launch (ops.rope_and_cache_flash); vLLM's
        RopeKVCacheFusionPass asks before rewriting the graph."""
        return self.kv_sharing_target_layer_name is None and self.head_size % 8 == 0 and not self._per_token_head

    def do_rope_and_kv_cache_update(self, layer, query, key, value, positions, cos_sin_cache, is_neox, kv_cache,
                                    layer_slot_mapping) -> None:
        key_cache, value_cache = self._views(kv_cache)
        ops.rope_and_cache_flash(query, key, value, positions, cos_sin_cache, is_neox, key_cache, value_cache,
                                 layer_slot_mapping, self.kv_cache_dtype, layer._k_scale, layer._v_scale)

    def forward(
        self,
        layer,
        query: torch.Tensor,  # [num_tokens, num_heads, head_size]
        key: torch.Tensor,
        value: torch.Tensor,
        kv_cache: torch.Tensor,
        attn_metadata: B200AttentionMetadata,
        output: Optional[torch.Tensor] = None,
        output_scale: Optional[torch.Tensor] = None,
        output_block_scale: Optional[torch.Tensor] = None,
    ) -> torch.Tensor:
        assert output is not None, "Output tensor must be provided."  # reference :366
        if output_block_scale is not None:
            raise NotImplementedError(  # vLLM 0.22 triton_attn.py:557-561 (reference :368-372 rejected both)
                "fused block_scale output quantization is not yet supported for B200AttentionImpl"
            )
        if attn_metadata is None:
            return output.fill_(0)  # profiling run (reference :374-376)
        assert attn_metadata.use_cascade is False
        if getattr(self, "dcp_world_size", 1) > 1:
            raise NotImplementedError("B200AttentionImpl: decode context parallelism (the cross-rank LSE combine) is not "
                                      "implemented; the kernels export the LSE (softmax_lse=) for a caller that combines")

        n = attn_metadata.num_actual_tokens
        q = query[:n].view(-1, self.num_heads, self.head_size)
        out = output[:n].view(-1, self.num_heads, self.head_size)
        # reference launcher arguments (triton_attn.py:446-468): causal, q_descale=None (:465),
        # window/softcap/alibi/scales are fixed per layer and live in the plan
        self._plans_for(layer, kv_cache, query.dtype)[1](
            q, out, attn_metadata.query_start_loc, attn_metadata.max_query_len, attn_metadata.seq_lens,
            attn_metadata.max_seq_len, attn_metadata.block_table, attn_metadata.workspace, output_scale,
            attn_metadata.decode_plan,
        )
        return output

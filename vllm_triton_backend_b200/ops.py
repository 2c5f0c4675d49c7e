"""Host-side operators: the reference's call signatures over the C ABI.

* `reshape_and_cache_flash(...)` has the argument list of
  `torch.ops._C_cache_ops.reshape_and_cache_flash` as the reference calls it
  (ibm-triton-lib/ibm_triton_lib/backend/triton_attn.py:396-405).
* `unified_attention(...)` has the keyword names of the reference launcher
  (ibm-triton-lib/ibm_triton_lib/kernels/triton_unified_attention.py:839-860) so
  `scripts/callers/unified_triton.py:61-82` and the upstream tests can call it
  unchanged, and mirrors its checks and its 2D/3D dispatch (:861-867, :884).

PyTorch is used for device memory and streams only; the arithmetic runs in
libb200attn.so.  Nothing here falls back to PyTorch or the CPU.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from . import _cabi
from ._cabi import AttentionParams

_DTYPE = {
    torch.bfloat16: _cabi.BF16,
    torch.float16: _cabi.F16,
    torch.float32: _cabi.F32,
    torch.float8_e4m3fn: _cabi.FP8_E4M3,
    torch.uint8: _cabi.FP8_E4M3,  # vLLM allocates fp8 caches as uint8
    torch.int8: _cabi.INT8,  # KV cache only, with per-(token, head) scale caches
}

# upstream KVQuantMode (vllm/v1/kv_cache_interface.py): NONE = 0 and FP8_PER_TENSOR = 1 take the per-tensor
# k_descale / v_descale; INT8_PER_TOKEN_HEAD = 2 and FP8_PER_TOKEN_HEAD = 3 take k_scale_cache / v_scale_cache
_PER_TENSOR_KV_MODES = (0, 1)
_PER_TOKEN_HEAD_KV_MODES = (2, 3)

_workspaces: dict = {}
_retired_workspaces: list = []
_scalar_cache: dict = {}


def _stream_ptr(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "b200attn operators run on CUDA tensors only (there is no CPU fallback); got a tensor on " f"{t.device}"
            )


def _scalar_tensor(value, device) -> torch.Tensor:
    """fp32 device scalar for a host float (cached; device tensors pass through)."""
    if isinstance(value, torch.Tensor):
        if value.dtype != torch.float32:
            raise ValueError("scales must be float32 tensors")
        return value
    key = (str(device), float(value))
    t = _scalar_cache.get(key)
    if t is None:
        t = torch.tensor([float(value)], dtype=torch.float32, device=device)
        _scalar_cache[key] = t
    return t


def workspace_bytes(num_tokens: int, num_seqs: int, num_q_heads: int, num_kv_heads: int, head_size: int) -> int:
    return int(_cabi.load_library().b200attn_workspace_bytes(num_tokens, num_seqs, num_q_heads, num_kv_heads, head_size))


def fp8_staging_bytes(num_seqs: int, tokens: int, block_size: int, num_kv_heads: int, head_size: int) -> int:
    """Extra workspace (past `workspace_bytes(...)`) that lets FP8-KV prefill run the tcgen05 kernel:
    a 16-bit staging pool for `tokens` cached tokens (csrc/kv_stage.cu).  Sequences that do not fit
    stay on the mma.sync kernel, so any size is valid; 0 disables staging."""
    if tokens <= 0:
        return 0
    pages = (tokens + block_size - 1) // block_size + num_seqs
    index = (num_seqs * 4 + 255) // 256 * 256
    return 256 + index + 2 * pages * block_size * num_kv_heads * head_size * 2 + 512


class DecodePlan:
    """The split-KV work plan of one step (`b200attn_decode_plan`): where every CTA and every pool chunk of
    the decode launches starts, computed on the device from `seq_lens` / `cu_seqlens_q`.  All layers of a step
    share those arrays, so the metadata builder builds it once and every layer's launch reads it; a launch
    without one plans for itself (one more small kernel per call).

    One buffer, rebuilt in place every step (its address is what a captured CUDA graph holds)."""

    __slots__ = ("buf", "key", "_p", "_key_out", "_fn", "_keep", "max_seqs")

    def __init__(self, device, max_num_seqs: int, num_q_heads: int, num_kv_heads: int, head_size: int, block_size: int,
                 q_dtype: torch.dtype, kv_dtype: torch.dtype, sliding_window: int = 0):
        lib = _cabi.load_library()
        self.max_seqs = int(max_num_seqs)
        self.buf = torch.zeros(int(lib.b200attn_decode_plan_bytes(self.max_seqs)), dtype=torch.uint8, device=device)
        p = AttentionParams()
        p.num_q_heads, p.num_kv_heads, p.head_size, p.block_size = num_q_heads, num_kv_heads, head_size, block_size
        p.q_dtype, p.kv_dtype = _DTYPE[q_dtype], _DTYPE[kv_dtype]
        p.sliding_window = int(sliding_window or 0)
        # alignment / stride fields decode_supported() looks at: dense views
        p.kc_block_stride = p.vc_block_stride = block_size * num_kv_heads * head_size
        p.kc_slot_stride = p.vc_slot_stride = num_kv_heads * head_size
        p.kc_head_stride = p.vc_head_stride = head_size
        p.q_stride0, p.q_stride1 = num_q_heads * head_size, head_size
        self._p, self._key_out, self._fn, self.key, self._keep = p, ctypes.c_uint64(0), lib.b200attn_decode_plan, 0, None

    def build(self, cu_seqlens_q: torch.Tensor, seqused_k: torch.Tensor, max_seqlen_q: int, num_tokens: int = 0) -> int:
        """Launches the plan kernel on the current stream; returns the key (0: nothing to plan for this shape).
        `num_tokens` (query tokens of the step) matters for mixed batches only: it sizes the share of the SMs the
        single-token rows get beside the prompts."""
        n = seqused_k.shape[0]
        if n > self.max_seqs:
            if torch.cuda.is_current_stream_capturing():
                self.key = 0
                return 0
            self.max_seqs = max(n, 2 * self.max_seqs)
            self._keep = (self._keep, self.buf)  # a captured graph may hold the old address
            self.buf = torch.zeros(int(_cabi.load_library().b200attn_decode_plan_bytes(self.max_seqs)), dtype=torch.uint8,
                                   device=self.buf.device)
        p = self._p
        p.seq_lens, p.cu_seqlens_q, p.num_seqs = seqused_k.data_ptr(), cu_seqlens_q.data_ptr(), n
        p.max_seqlen_q = int(max_seqlen_q)
        p.num_tokens = int(num_tokens) if num_tokens else n
        rc = self._fn(ctypes.byref(p), self.buf.data_ptr(), self.buf.numel(), ctypes.byref(self._key_out),
                      torch.cuda.current_stream(self.buf.device).cuda_stream)
        if rc:
            _cabi.check(rc)
        self.key = int(self._key_out.value)
        return self.key


def get_workspace(device, nbytes: int) -> torch.Tensor:
    """Grow-only per-device split-KV workspace (replaces the reference's per-call
    torch.empty of segm_output/segm_max/segm_expsum, triton_unified_attention.py:950-971).
    Zero-filled once: it starts with the kernels' self-cleaning ticket counters."""
    key = str(device)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("b200attn workspace must be allocated before CUDA-graph capture (pass workspace=...)")
        if ws is not None:
            # a CUDA graph captured earlier may still hold its address; the buffer at least doubles, so the
            # retired ones add up to less than the live one
            _retired_workspaces.append(ws)
            nbytes = max(nbytes, 2 * ws.numel())
        ws = torch.zeros(max(nbytes, 4 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def reshape_and_cache_flash(
    key: torch.Tensor,  # [>=T, Hkv, D]
    value: torch.Tensor,
    key_cache: torch.Tensor,  # [NB, bs, Hkv, D] view (any block stride; NHD or HND strides)
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,  # int64 [T]
    kv_cache_dtype: str,
    k_scale,
    v_scale,
) -> None:
    _require_cuda(key, value, key_cache, value_cache, slot_mapping)
    if slot_mapping.dtype != torch.int64:
        raise ValueError("slot_mapping must be int64")
    if key.dim() != 3 or key_cache.dim() != 4:
        raise ValueError("key must be [T, Hkv, D] and key_cache [NB, bs, Hkv, D]")
    if key.stride(2) != 1 or key.stride(1) != key.shape[2] or value.stride(2) != 1 or value.stride(1) != value.shape[2]:
        raise ValueError("key/value rows must be contiguous in (head, dim)")
    if key_cache.stride(3) != 1 or value_cache.stride(3) != 1:
        raise ValueError("cache head_dim must be the contiguous dimension")
    src_dt = _DTYPE.get(key.dtype)
    fp8 = kv_cache_dtype.startswith("fp8")
    if fp8:
        if kv_cache_dtype not in ("fp8", "fp8_e4m3"):
            raise ValueError(f"unsupported kv_cache_dtype {kv_cache_dtype!r} (fp8_e4m3 only)")
        if key_cache.element_size() != 1:
            raise ValueError("fp8 kv cache must be a 1-byte tensor")
        kv_dt = _cabi.FP8_E4M3
        ks = _scalar_tensor(k_scale, key.device)
        vs = _scalar_tensor(v_scale, key.device)
        ks_ptr, vs_ptr = ks.data_ptr(), vs.data_ptr()
    else:
        if key_cache.dtype != key.dtype:
            raise ValueError(f"kv cache dtype {key_cache.dtype} != key dtype {key.dtype} with kv_cache_dtype={kv_cache_dtype!r}")
        kv_dt = src_dt
        ks_ptr = vs_ptr = None
    if src_dt is None:
        raise ValueError(f"unsupported key dtype {key.dtype}")
    lib = _cabi.load_library()
    _, bs, hkv, d = key_cache.shape
    _cabi.check(
        lib.b200attn_reshape_and_cache_flash(
            key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), slot_mapping.data_ptr(),
            slot_mapping.shape[0], hkv, d, bs, key.stride(0), value.stride(0),
            key_cache.stride(0), key_cache.stride(1), key_cache.stride(2),
            value_cache.stride(0), value_cache.stride(1), value_cache.stride(2),
            src_dt, kv_dt, ks_ptr, vs_ptr, _stream_ptr(key),
        )
    )


def reshape_and_cache_flash_per_token_head(
    key: torch.Tensor,  # [>=T, Hkv, D]
    value: torch.Tensor,
    key_cache: torch.Tensor,  # [NB, bs, Hkv, D] float8_e4m3fn / uint8 (e4m3) or int8 view
    value_cache: torch.Tensor,
    k_scale_cache: torch.Tensor,  # [NB, bs, Hkv] float32 view (any strides)
    v_scale_cache: torch.Tensor,
    slot_mapping: torch.Tensor,  # int64 [T]
) -> None:
    """Cache write of kv_cache_dtype "fp8_per_token_head" / "int8_per_token_head": every (token, head) row is quantised
    with its own scale absmax / QMAX, which goes to the scale cache (upstream
    `triton_reshape_and_cache_flash_per_token_head_quant`, vLLM 0.22
    vllm/v1/attention/ops/triton_reshape_and_cache_flash.py:247-320; same argument order)."""
    _require_cuda(key, value, key_cache, value_cache, k_scale_cache, v_scale_cache, slot_mapping)
    if slot_mapping.dtype != torch.int64:
        raise ValueError("slot_mapping must be int64")
    if key.dim() != 3 or key_cache.dim() != 4 or k_scale_cache.dim() != 3 or v_scale_cache.dim() != 3:
        raise ValueError("key must be [T, Hkv, D], key_cache [NB, bs, Hkv, D] and the scale caches [NB, bs, Hkv]")
    if key.stride(2) != 1 or value.stride(2) != 1 or key_cache.stride(3) != 1 or value_cache.stride(3) != 1:
        raise ValueError("head_dim must be the contiguous dimension of key, value and the caches")
    if k_scale_cache.dtype != torch.float32 or v_scale_cache.dtype != torch.float32:
        raise ValueError("scale caches must be float32")
    if key_cache.dtype != value_cache.dtype or key_cache.element_size() != 1:
        raise ValueError(f"per-token-head caches are 1-byte tensors (float8_e4m3fn / uint8, or int8); got {key_cache.dtype}")
    src_dt, kv_dt = _DTYPE.get(key.dtype), _DTYPE[key_cache.dtype]
    if src_dt not in (_cabi.BF16, _cabi.F16, _cabi.F32) or value.dtype != key.dtype:
        raise ValueError(f"unsupported key/value dtype {key.dtype}/{value.dtype}")
    _, bs, hkv, d = key_cache.shape
    if tuple(k_scale_cache.shape) != tuple(key_cache.shape[:3]) or tuple(v_scale_cache.shape) != tuple(value_cache.shape[:3]):
        raise ValueError("scale caches must be [num_blocks, block_size, num_kv_heads] like their caches")
    _cabi.check(
        _cabi.load_library().b200attn_reshape_and_cache_flash_per_token_head(
            key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), k_scale_cache.data_ptr(),
            v_scale_cache.data_ptr(), slot_mapping.data_ptr(), slot_mapping.shape[0], hkv, d, bs,
            key.stride(0), key.stride(1), value.stride(0), value.stride(1),
            key_cache.stride(0), key_cache.stride(1), key_cache.stride(2),
            value_cache.stride(0), value_cache.stride(1), value_cache.stride(2),
            k_scale_cache.stride(0), k_scale_cache.stride(1), k_scale_cache.stride(2),
            v_scale_cache.stride(0), v_scale_cache.stride(1), v_scale_cache.stride(2),
            src_dt, kv_dt, _stream_ptr(key),
        )
    )


def scaled_fp8_quant(x: torch.Tensor, scale: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Static per-tensor FP8 quantisation (`ops.scaled_fp8_quant(x, scale)`,
    reference triton_attn.py:417). x: [rows, n]."""
    _require_cuda(x, scale)
    if x.dim() != 2 or x.stride(1) != 1:
        raise ValueError("x must be 2-D with a contiguous last dimension")
    out = torch.empty(x.shape, dtype=torch.float8_e4m3fn, device=x.device)
    _cabi.check(
        _cabi.load_library().b200attn_scaled_fp8_quant(
            out.data_ptr(), x.data_ptr(), x.shape[0], x.shape[1], x.stride(0), _DTYPE[x.dtype], scale.data_ptr(), _stream_ptr(x)
        )
    )
    return out, scale


def _check_out(out, q_dtype, output_scale) -> None:
    """`output_scale` (upstream keyword: fused static per-tensor FP8 quantisation of the output,
    vllm/v1/attention/ops/triton_unified_attention.py:786,986) switches `out` to float8_e4m3fn."""
    if output_scale is None:
        if out.dtype != q_dtype:
            raise ValueError("out dtype must equal q dtype")
        return
    if out.dtype not in (torch.float8_e4m3fn, torch.uint8):
        raise ValueError("with output_scale the output must be float8_e4m3fn")
    if output_scale.dtype != torch.float32 or output_scale.numel() != 1 or not output_scale.is_cuda:
        raise ValueError("output_scale must be a one-element float32 CUDA tensor")


def build_attention_params(
    q, k, v, out, cu_seqlens_q, max_seqlen_q, seqused_k, max_seqlen_k, softmax_scale, window_size, block_table,
    softcap, k_descale, v_descale, alibi_slopes, path, workspace, sinks=None, output_scale=None, qq_bias=None,
    softmax_lse=None, k_scale_cache=None, v_scale_cache=None,
) -> AttentionParams:
    if q.dim() != 3 or k.dim() != 4 or v.dim() != 4:
        raise ValueError("q must be [T, Hq, D]; k, v must be [NB, bs, Hkv, D]")
    if q.stride(2) != 1 or out.stride(2) != 1 or k.stride(3) != 1 or v.stride(3) != 1:
        raise ValueError("innermost dimension of q/out/k/v must be contiguous")
    if block_table.dtype != torch.int32 or seqused_k.dtype != torch.int32 or cu_seqlens_q.dtype != torch.int32:
        raise ValueError("block_table, seqused_k and cu_seqlens_q must be int32")
    _check_out(out, q.dtype, output_scale)
    q_dt, kv_dt = _DTYPE.get(q.dtype), _DTYPE.get(k.dtype)
    if q_dt is None or q_dt == _cabi.FP8_E4M3:
        raise ValueError(f"unsupported query dtype {q.dtype}")
    if kv_dt is None or v.dtype != k.dtype:
        raise ValueError(f"unsupported kv cache dtypes {k.dtype}/{v.dtype}")
    p = AttentionParams()
    p.out, p.q, p.k_cache, p.v_cache = out.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr()
    p.block_tables, p.seq_lens, p.cu_seqlens_q = block_table.data_ptr(), seqused_k.data_ptr(), cu_seqlens_q.data_ptr()
    p.alibi_slopes = alibi_slopes.data_ptr() if alibi_slopes is not None else None
    p.k_scale = k_descale.data_ptr() if k_descale is not None else None
    p.v_scale = v_descale.data_ptr() if v_descale is not None else None
    if sinks is not None:
        if sinks.shape[0] != q.shape[1]:
            raise ValueError("Sinks must be num_query_heads size")  # upstream's assertion
        if sinks.dtype != torch.float32 or not sinks.is_contiguous():
            sinks = sinks.detach().to(torch.float32).contiguous()  # upstream takes the model dtype; the kernels read fp32
        p._keep_sinks = sinks  # the struct holds a raw pointer
    p.sinks = sinks.data_ptr() if sinks is not None else None
    p.out_scale = output_scale.data_ptr() if output_scale is not None else None
    if qq_bias is not None:
        if qq_bias.dtype != torch.float32 or qq_bias.dim() != 2 or qq_bias.stride(1) != 1 or not qq_bias.is_cuda:
            raise ValueError("qq_bias must be a 2-D float32 CUDA tensor with a contiguous last dimension")
        if qq_bias.shape[0] < int(max_seqlen_q):
            raise ValueError("qq_bias needs one row per query position of the longest sequence")
        p.qq_bias, p.qq_bias_stride0 = qq_bias.data_ptr(), qq_bias.stride(0)
    if softmax_lse is not None:
        if (softmax_lse.dtype != torch.float32 or softmax_lse.dim() != 2 or softmax_lse.stride(1) != 1 or not softmax_lse.is_cuda
                or softmax_lse.shape[0] < q.shape[0] or softmax_lse.shape[1] < q.shape[1]):
            raise ValueError("softmax_lse must be a float32 CUDA tensor [num_tokens, num_q_heads] with a contiguous last dimension")
        p.lse, p.lse_stride0 = softmax_lse.data_ptr(), softmax_lse.stride(0)
    p.softmax_scale = float(softmax_scale)
    p.softcap = float(softcap or 0.0)
    p.sliding_window = int(1 + window_size[0]) if window_size is not None else 0  # reference :915
    p.num_tokens, p.num_seqs = q.shape[0], seqused_k.shape[0]
    p.num_q_heads, p.num_kv_heads, p.head_size = q.shape[1], k.shape[2], q.shape[2]
    p.block_size = v.shape[1]
    p.max_seqlen_q, p.max_seqlen_k = int(max_seqlen_q), int(max_seqlen_k)
    p.q_dtype, p.kv_dtype, p.path = q_dt, kv_dt, path
    p.num_blocks = k.shape[0]
    p.q_stride0, p.q_stride1 = q.stride(0), q.stride(1)
    p.out_stride0, p.out_stride1 = out.stride(0), out.stride(1)
    p.bt_stride0 = block_table.stride(0)
    p.kc_block_stride, p.kc_slot_stride, p.kc_head_stride = k.stride(0), k.stride(1), k.stride(2)
    p.vc_block_stride, p.vc_slot_stride, p.vc_head_stride = v.stride(0), v.stride(1), v.stride(2)
    if workspace is not None:
        p.workspace, p.workspace_bytes = workspace.data_ptr(), workspace.numel() * workspace.element_size()
    else:
        p.workspace, p.workspace_bytes = None, 0
    _set_scale_caches(p, k, v, k_scale_cache, v_scale_cache)
    return p


def _set_scale_caches(p: AttentionParams, k, v, k_scale_cache, v_scale_cache) -> None:
    """Per-(token, head) scale caches (upstream `k_scale_cache` / `v_scale_cache`): [NB, bs, Hkv] float32 views."""
    if k_scale_cache is None and v_scale_cache is None:
        if k.dtype == torch.int8:
            raise ValueError("an int8 kv cache needs k_scale_cache / v_scale_cache (kv_quant_mode INT8_PER_TOKEN_HEAD)")
        return
    if k_scale_cache is None or v_scale_cache is None:
        raise ValueError("k_scale_cache and v_scale_cache come together")
    if k.element_size() != 1:
        raise ValueError("per-token-head scale caches need a 1-byte kv cache (float8_e4m3fn / uint8 or int8)")
    for sc, c in ((k_scale_cache, k), (v_scale_cache, v)):
        if sc.dtype != torch.float32 or not sc.is_cuda or tuple(sc.shape) != tuple(c.shape[:3]):
            raise ValueError("scale caches must be float32 CUDA tensors [num_blocks, block_size, num_kv_heads]")
    p.k_scale, p.v_scale = None, None  # not used in this mode (b200attn.h)
    p.k_scale_cache, p.v_scale_cache = k_scale_cache.data_ptr(), v_scale_cache.data_ptr()
    p.ks_block_stride, p.ks_slot_stride, p.ks_head_stride = k_scale_cache.stride()
    p.vs_block_stride, p.vs_slot_stride, p.vs_head_stride = v_scale_cache.stride()


def unified_attention(
    q: torch.Tensor,  # [T, Hq, D]
    k: torch.Tensor,  # key cache   [NB, bs, Hkv, D]
    v: torch.Tensor,  # value cache [NB, bs, Hkv, D]
    out: torch.Tensor,  # [T, Hq, D]
    cu_seqlens_q: torch.Tensor,  # int32 [B+1]
    max_seqlen_q: int,
    seqused_k: torch.Tensor,  # int32 [B]
    max_seqlen_k: int,
    softmax_scale: float,
    causal: bool,
    window_size,  # (sliding_window - 1, 0) or (-1, -1)
    block_table: torch.Tensor,  # int32 [B, MB]
    softcap: float,
    q_descale,
    k_descale: Optional[torch.Tensor],
    v_descale: Optional[torch.Tensor],
    alibi_slopes: Optional[torch.Tensor] = None,
    avg_seqlen_q=None,  # tuning inputs of the reference's autotuner; unused here
    avg_seqlen_k=None,
    force_selection: Optional[int] = None,  # None, 2 (2D kernel), 3 (3D kernel) as in the reference; 1 = SIMT net
    workspace: Optional[torch.Tensor] = None,
    sinks: Optional[torch.Tensor] = None,  # [Hq] float32 (upstream keyword)
    output_scale: Optional[torch.Tensor] = None,  # upstream keyword: `out` is e4m3, out = quant(o / output_scale)
    qq_bias: Optional[torch.Tensor] = None,  # upstream keyword: [q, q] fp32 bias between a step's query tokens (SIMT kernel)
    decode_plan: Optional["DecodePlan"] = None,  # this step's split-KV plan (DecodePlan.build), shared by all layers
    softmax_lse: Optional[torch.Tensor] = None,  # out: fp32 [T, Hq], natural-log sum of exponentials of every row's logits
    # Pure tuning inputs of upstream's launcher (vLLM 0.22 vllm/v1/attention/ops/triton_unified_attention.py):
    # where its split-KV partials live and when it splits.  They do not change the result; this library sizes
    # and owns its workspace itself, so they are accepted and unused.
    seq_threshold_3D=None,
    num_par_softmax_segments=None,
    softmax_segm_output=None,
    softmax_segm_max=None,
    softmax_segm_expsum=None,
    use_td: bool = False,
    # Result-changing upstream keywords that are not built here: a non-default value raises instead of
    # silently computing plain causal attention.
    mm_prefix_range=None,
    use_alibi_sqrt: bool = False,
    chunk_lookback: int = -1,
    # upstream's per-(token, head) quantised caches: kv_quant_mode INT8_PER_TOKEN_HEAD (2) / FP8_PER_TOKEN_HEAD (3)
    # with one fp32 scale per cached (token, kv head) in k_scale_cache / v_scale_cache [NB, bs, Hkv]
    kv_quant_mode=None,
    k_scale_cache: Optional[torch.Tensor] = None,
    v_scale_cache: Optional[torch.Tensor] = None,
) -> None:
    if mm_prefix_range is not None:
        raise NotImplementedError("b200attn: mm_prefix_range (PrefixLM bidirectional ranges) is not implemented")
    if use_alibi_sqrt:
        raise NotImplementedError("b200attn: use_alibi_sqrt is not implemented")
    if chunk_lookback is not None and int(chunk_lookback) >= 0:
        raise NotImplementedError("b200attn: chunk_lookback (chunked attention with lookback) is not implemented; "
                                  "chunked local attention runs through vLLM's virtual batches")
    mode = 0 if kv_quant_mode is None else int(getattr(kv_quant_mode, "value", kv_quant_mode))
    if mode in _PER_TOKEN_HEAD_KV_MODES:
        if k_scale_cache is None or v_scale_cache is None:
            raise AssertionError(f"kv_quant_mode={kv_quant_mode!r} requires k_scale_cache / v_scale_cache")  # upstream :814
        if (mode == 2) != (k.dtype == torch.int8):
            raise ValueError(f"kv_quant_mode={kv_quant_mode!r} does not match the cache dtype {k.dtype}")
    elif mode in _PER_TENSOR_KV_MODES:
        if kv_quant_mode is not None and (k_scale_cache is not None or v_scale_cache is not None):
            raise ValueError(f"scale caches passed with kv_quant_mode={kv_quant_mode!r}")
    else:
        raise NotImplementedError(f"b200attn: kv_quant_mode={kv_quant_mode!r} is not implemented")
    if not causal:
        raise AssertionError("Only causal attention is supported")  # reference :861
    if q_descale is not None:
        raise AssertionError("Q scales not supported")  # reference :862
    _require_cuda(q, k, v, out, cu_seqlens_q, seqused_k, block_table, sinks)
    if q.element_size() < 2:
        # the reference's fp8-Q branch ignores k/v scales; not offered
        raise ValueError("fp8 queries are not supported; keep Q in bf16/fp16")

    max_seqlen_q = int(max_seqlen_q)
    # reference dispatch (:884): any batch with a multi-token query runs the 2D kernel
    if force_selection == 1:
        path = _cabi.PATH_GENERIC
    elif force_selection == 4:
        path = _cabi.PATH_PREFILL_TC
    elif max_seqlen_q > 1:
        path = _cabi.PATH_AUTO if force_selection != 22 else _cabi.PATH_PREFILL
    elif force_selection == 2:
        path = _cabi.PATH_PREFILL
    else:
        path = _cabi.PATH_AUTO

    if workspace is None and path in (_cabi.PATH_AUTO, _cabi.PATH_DECODE):
        # decode batches, and the single-token rows of a mixed batch, use the split-KV kernel
        need = workspace_bytes(q.shape[0], seqused_k.shape[0], q.shape[1], k.shape[2], q.shape[2])
        if k.element_size() == 1 and max_seqlen_q > 1:
            # FP8 prefill: room to stage the batch (bounded at 128k tokens - 512 MiB at 8 kv heads, D=128; what does
            # not fit runs mma.sync).  Callers that care pass their own workspace.
            tokens = min(seqused_k.shape[0] * int(max_seqlen_k), 1 << 17)
            need += fp8_staging_bytes(seqused_k.shape[0], tokens, v.shape[1], k.shape[2], q.shape[2])
        workspace = get_workspace(q.device, need)
    p = build_attention_params(
        q, k, v, out, cu_seqlens_q, max_seqlen_q, seqused_k, max_seqlen_k, softmax_scale, window_size, block_table,
        softcap, k_descale, v_descale, alibi_slopes, path, workspace, sinks, output_scale, qq_bias, softmax_lse,
        k_scale_cache, v_scale_cache,
    )
    if decode_plan is not None and decode_plan.key:
        p.decode_plan, p.decode_plan_key = decode_plan.buf.data_ptr(), decode_plan.key
    _cabi.check(_cabi.load_library().b200attn_unified_attention(ctypes.byref(p), _stream_ptr(q)))


# ---------------------------------------------------------------------------------
# Per-layer plans: the argument marshalling above costs ~20 us of Python per call
# (every .stride()/.data_ptr() is a C++ round trip).  Inside vLLM the same layer
# calls with the same cache tensor and head geometry on every step, and in
# piecewise-CUDA-graph mode the attention op runs eagerly once per layer per step,
# so the backend keeps the invariant part filled in and only writes the step's
# pointers and sizes.  Same C entry points, same checks (done once, at plan time).
# ---------------------------------------------------------------------------------
class AttentionPlan:
    """`b200attn_attention_params` pre-filled for one (cache, head geometry, features)."""

    __slots__ = ("p", "_ref", "_fn", "_keep", "_ws_ptr", "num_q_heads", "head_size", "q_dtype")

    def __init__(self, k, v, q_dtype, num_q_heads, head_size, softmax_scale, window_size, softcap, alibi_slopes,
                 k_descale, v_descale, sinks=None, k_scale_cache=None, v_scale_cache=None):
        _require_cuda(k, v, alibi_slopes, k_descale, v_descale, sinks, k_scale_cache, v_scale_cache)
        if sinks is not None:
            if sinks.shape[0] != num_q_heads:
                raise ValueError("sinks must have one entry per query head")
            sinks = sinks.detach().to(torch.float32).contiguous()
        if k.dim() != 4 or v.dim() != 4 or k.stride(3) != 1 or v.stride(3) != 1:
            raise ValueError("k, v must be [NB, bs, Hkv, D] views with a contiguous innermost dimension")
        q_dt, kv_dt = _DTYPE.get(q_dtype), _DTYPE.get(k.dtype)
        if q_dt is None or q_dt == _cabi.FP8_E4M3:
            raise ValueError(f"unsupported query dtype {q_dtype}")
        if kv_dt is None or v.dtype != k.dtype:
            raise ValueError(f"unsupported kv cache dtypes {k.dtype}/{v.dtype}")
        p = AttentionParams()
        p.k_cache, p.v_cache = k.data_ptr(), v.data_ptr()
        p.alibi_slopes = alibi_slopes.data_ptr() if alibi_slopes is not None else None
        p.k_scale = k_descale.data_ptr() if k_descale is not None else None
        p.v_scale = v_descale.data_ptr() if v_descale is not None else None
        p.sinks = sinks.data_ptr() if sinks is not None else None
        p.softmax_scale = float(softmax_scale)
        p.softcap = float(softcap or 0.0)
        p.sliding_window = int(1 + window_size[0]) if window_size is not None else 0
        p.num_q_heads, p.num_kv_heads, p.head_size, p.block_size = num_q_heads, k.shape[2], head_size, v.shape[1]
        p.q_dtype, p.kv_dtype, p.path = q_dt, kv_dt, _cabi.PATH_AUTO
        p.num_blocks = k.shape[0]
        p.kc_block_stride, p.kc_slot_stride, p.kc_head_stride = k.stride(0), k.stride(1), k.stride(2)
        p.vc_block_stride, p.vc_slot_stride, p.vc_head_stride = v.stride(0), v.stride(1), v.stride(2)
        _set_scale_caches(p, k, v, k_scale_cache, v_scale_cache)
        self.p, self._ref = p, ctypes.byref(p)
        self._fn = _cabi.load_library().b200attn_unified_attention
        # the struct holds raw pointers into these
        self._keep = (k, v, alibi_slopes, k_descale, v_descale, sinks, k_scale_cache, v_scale_cache)
        self._ws_ptr = None
        self.num_q_heads, self.head_size, self.q_dtype = num_q_heads, head_size, q_dtype

    def __call__(self, q, out, cu_seqlens_q, max_seqlen_q, seqused_k, max_seqlen_k, block_table, workspace,
                 output_scale=None, decode_plan=None, softmax_lse=None) -> None:
        """q/out: [T, Hq, D] views (innermost contiguous); metadata tensors int32 on the device.
        `output_scale`: one-element fp32 device tensor => `out` is float8_e4m3fn (fused output quant)."""
        if q.dtype != self.q_dtype or not q.is_cuda:
            raise ValueError("query dtype or device differs from the plan")
        _check_out(out, self.q_dtype, output_scale)
        p = self.p
        p.q, p.out = q.data_ptr(), out.data_ptr()
        p.out_scale = output_scale.data_ptr() if output_scale is not None else None
        qs, os_ = q.stride(), out.stride()
        if qs[2] != 1 or os_[2] != 1:
            raise ValueError("innermost dimension of q/out must be contiguous")
        p.q_stride0, p.q_stride1, p.out_stride0, p.out_stride1 = qs[0], qs[1], os_[0], os_[1]
        p.num_tokens, p.num_seqs = q.shape[0], seqused_k.shape[0]
        p.block_tables, p.seq_lens, p.cu_seqlens_q = block_table.data_ptr(), seqused_k.data_ptr(), cu_seqlens_q.data_ptr()
        p.bt_stride0 = block_table.stride(0)
        p.max_seqlen_q, p.max_seqlen_k = max_seqlen_q, max_seqlen_k
        if decode_plan is not None and decode_plan.key:
            p.decode_plan, p.decode_plan_key = decode_plan.buf.data_ptr(), decode_plan.key
        else:
            p.decode_plan, p.decode_plan_key = None, 0
        if softmax_lse is not None:
            if softmax_lse.dtype != torch.float32 or softmax_lse.dim() != 2 or softmax_lse.stride(1) != 1 or \
                    softmax_lse.shape[0] < q.shape[0] or softmax_lse.shape[1] < self.num_q_heads:
                raise ValueError("softmax_lse must be float32 [num_tokens, num_q_heads]")
            p.lse, p.lse_stride0 = softmax_lse.data_ptr(), softmax_lse.stride(0)
        else:
            p.lse, p.lse_stride0 = None, 0
        wp = workspace.data_ptr()
        if wp != self._ws_ptr:
            p.workspace, p.workspace_bytes, self._ws_ptr = wp, workspace.numel() * workspace.element_size(), wp
        rc = self._fn(self._ref, torch.cuda.current_stream(q.device).cuda_stream)
        if rc:
            _cabi.check(rc)


class CacheWritePlan:
    """`b200attn_reshape_and_cache_flash` with the cache side of the argument list fixed."""

    __slots__ = ("_fn", "_tail", "_head", "_keep", "_src_dtype", "_src_dt")

    def __init__(self, key_cache, value_cache, kv_cache_dtype: str, src_dtype, k_scale, v_scale):
        _require_cuda(key_cache, value_cache)
        if key_cache.dim() != 4 or key_cache.stride(3) != 1 or value_cache.stride(3) != 1:
            raise ValueError("caches must be [NB, bs, Hkv, D] views with a contiguous head_dim")
        src_dt = _DTYPE.get(src_dtype)
        if src_dt is None or src_dt == _cabi.FP8_E4M3:
            raise ValueError(f"unsupported key dtype {src_dtype}")
        if kv_cache_dtype.startswith("fp8"):
            if kv_cache_dtype not in ("fp8", "fp8_e4m3"):
                raise ValueError(f"unsupported kv_cache_dtype {kv_cache_dtype!r} (fp8_e4m3 only)")
            if key_cache.element_size() != 1:
                raise ValueError("fp8 kv cache must be a 1-byte tensor")
            kv_dt = _cabi.FP8_E4M3
            ks, vs = _scalar_tensor(k_scale, key_cache.device), _scalar_tensor(v_scale, key_cache.device)
            ks_ptr, vs_ptr = ks.data_ptr(), vs.data_ptr()
        else:
            if key_cache.dtype != src_dtype:
                raise ValueError(f"kv cache dtype {key_cache.dtype} != key dtype {src_dtype} with kv_cache_dtype={kv_cache_dtype!r}")
            kv_dt, ks, vs, ks_ptr, vs_ptr = src_dt, None, None, None, None
        _, bs, hkv, d = key_cache.shape
        self._fn = _cabi.load_library().b200attn_reshape_and_cache_flash
        self._head = (key_cache.data_ptr(), value_cache.data_ptr())
        self._tail = (hkv, d, bs)
        self._keep = (key_cache, value_cache, ks, vs,
                      (key_cache.stride(0), key_cache.stride(1), key_cache.stride(2),
                       value_cache.stride(0), value_cache.stride(1), value_cache.stride(2), src_dt, kv_dt, ks_ptr, vs_ptr))
        self._src_dtype, self._src_dt = src_dtype, src_dt

    def __call__(self, key, value, slot_mapping) -> None:
        if key.dtype != self._src_dtype or not key.is_cuda or slot_mapping.dtype != torch.int64:
            raise ValueError("key dtype/device differs from the plan, or slot_mapping is not int64")
        ks, vs = key.stride(), value.stride()
        hkv, d, bs = self._tail
        if ks[2] != 1 or ks[1] != d or vs[2] != 1 or vs[1] != d:
            raise ValueError("key/value rows must be contiguous in (head, dim)")
        rc = self._fn(key.data_ptr(), value.data_ptr(), self._head[0], self._head[1], slot_mapping.data_ptr(),
                      slot_mapping.shape[0], hkv, d, bs, ks[0], vs[0], *self._keep[4],
                      torch.cuda.current_stream(key.device).cuda_stream)
        if rc:
            _cabi.check(rc)


class PerTokenHeadCacheWritePlan:
    """`b200attn_reshape_and_cache_flash_per_token_head` with the cache side of the argument list fixed."""

    __slots__ = ("_fn", "_head", "_shape", "_tail", "_keep", "_src_dtype")

    def __init__(self, key_cache, value_cache, k_scale_cache, v_scale_cache, src_dtype):
        _require_cuda(key_cache, value_cache, k_scale_cache, v_scale_cache)
        if key_cache.dim() != 4 or key_cache.stride(3) != 1 or value_cache.stride(3) != 1 or key_cache.element_size() != 1:
            raise ValueError("caches must be 1-byte [NB, bs, Hkv, D] views with a contiguous head_dim")
        if k_scale_cache.dtype != torch.float32 or v_scale_cache.dtype != torch.float32 or \
                tuple(k_scale_cache.shape) != tuple(key_cache.shape[:3]) or tuple(v_scale_cache.shape) != tuple(value_cache.shape[:3]):
            raise ValueError("scale caches must be float32 [NB, bs, Hkv]")
        src_dt = _DTYPE.get(src_dtype)
        if src_dt not in (_cabi.BF16, _cabi.F16, _cabi.F32):
            raise ValueError(f"unsupported key dtype {src_dtype}")
        _, bs, hkv, d = key_cache.shape
        self._fn = _cabi.load_library().b200attn_reshape_and_cache_flash_per_token_head
        self._head = (key_cache.data_ptr(), value_cache.data_ptr(), k_scale_cache.data_ptr(), v_scale_cache.data_ptr())
        self._shape = (hkv, d, bs)
        self._tail = (*key_cache.stride()[:3], *value_cache.stride()[:3], *k_scale_cache.stride(), *v_scale_cache.stride(),
                      src_dt, _DTYPE[key_cache.dtype])
        self._keep = (key_cache, value_cache, k_scale_cache, v_scale_cache)
        self._src_dtype = src_dtype

    def __call__(self, key, value, slot_mapping) -> None:
        if key.dtype != self._src_dtype or value.dtype != key.dtype or not key.is_cuda or slot_mapping.dtype != torch.int64:
            raise ValueError("key dtype/device differs from the plan, or slot_mapping is not int64")
        ks, vs = key.stride(), value.stride()
        if ks[2] != 1 or vs[2] != 1:
            raise ValueError("key/value rows must be contiguous in head_dim")
        rc = self._fn(key.data_ptr(), value.data_ptr(), *self._head, slot_mapping.data_ptr(), slot_mapping.shape[0],
                      *self._shape, ks[0], ks[1], vs[0], vs[1], *self._tail, torch.cuda.current_stream(key.device).cuda_stream)
        if rc:
            _cabi.check(rc)


PER_TOKEN_HEAD_SCALE_PAD_BYTES = 4  # one fp32 scale per (token, head) row, budgeted by vLLM as head_size + 4 / elem size


def per_token_head_cache_views(plane: torch.Tensor, head_size: int, kv_dtype: torch.dtype):
    """Data and scale views of one plane (K or V) of a per-(token, head) quantised cache.

    `plane`: [NB, bs, Hkv, head_size + 4] one-byte elements, each block contiguous - the bytes vLLM budgets for this
    mode (`AttentionSpec.page_size_bytes`, vllm/v1/kv_cache_interface.py:150-160: the data plus one fp32 per (token,
    head)).  Upstream's Triton backend keeps each row's scale INLINE behind the row (rows of head_size + 4 bytes,
    triton_attn.py:401-452): a 132-byte row pitch that no TMA descriptor can express (strides must be 16-byte
    multiples).  Here a block holds its data first, [bs][Hkv][head_size] dense as in the unquantised layout, and
    behind it the block's scales as [Hkv][bs] fp32 - the same bytes, TMA-able rows, and a warp of the decode kernel
    finds the 16 scales of its 16 tokens in one 64-byte line.

    Returns (data [NB, bs, Hkv, head_size] as `kv_dtype`, scales [NB, bs, Hkv] float32 with strides (block, 1, bs))."""
    if plane.dim() != 4 or plane.element_size() != 1 or plane.shape[3] != head_size + PER_TOKEN_HEAD_SCALE_PAD_BYTES:
        raise ValueError("plane must be [NB, bs, Hkv, head_size + 4] one-byte elements")
    nb, bs, hkv, dp = plane.shape
    blk = bs * hkv * dp
    if plane.stride() != (plane.stride(0), hkv * dp, dp, 1) or plane.stride(0) % 16 or (bs * hkv * head_size) % 16:
        raise ValueError("per-token-head cache blocks must be contiguous (NHD) with 16-byte aligned block strides")
    if plane.storage_offset() % 16:
        raise ValueError("per-token-head cache plane must start 16-byte aligned")
    u8 = plane.view(torch.uint8)
    data = u8.as_strided((nb, bs, hkv, head_size), (plane.stride(0), hkv * head_size, head_size, 1), plane.storage_offset())
    raw = torch.empty(0, dtype=torch.float32, device=plane.device).set_(plane.untyped_storage())
    scales = raw.as_strided((nb, bs, hkv), (plane.stride(0) // 4, 1, bs), (plane.storage_offset() + bs * hkv * head_size) // 4)
    assert blk >= bs * hkv * head_size + 4 * bs * hkv
    return data.view(kv_dtype), scales


def rope_and_cache_flash(
    query: torch.Tensor,  # [T, Hq, D] (any token stride: usually a view into the fused qkv projection)
    key: torch.Tensor,  # [T, Hkv, D]
    value: torch.Tensor,  # [T, Hkv, D]
    positions: torch.Tensor,  # int64 [T]
    cos_sin_cache: torch.Tensor,  # [max_position, rot_dim], query dtype
    is_neox: bool,
    key_cache: torch.Tensor,  # [NB, bs, Hkv, D] view
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,  # int64 [<=T]
    kv_cache_dtype: str,
    k_scale,
    v_scale,
) -> None:
    """Rotary embedding of `query`/`key` in place and the KV-cache write, in one launch
    (`AttentionImpl.do_rope_and_kv_cache_update`, vLLM 0.22 vllm/v1/attention/backend.py:823-841;
    unfused pair: `ops.rotary_embedding` + `reshape_and_cache_flash`, reference triton_attn.py:396-405)."""
    _require_cuda(query, key, value, positions, cos_sin_cache, key_cache, value_cache, slot_mapping)
    if query.dim() != 3 or key.dim() != 3 or value.dim() != 3 or key_cache.dim() != 4:
        raise ValueError("query/key/value must be [T, H, D] and the caches [NB, bs, Hkv, D]")
    if query.stride(2) != 1 or key.stride(2) != 1 or value.stride(2) != 1 or key_cache.stride(3) != 1 or value_cache.stride(3) != 1:
        raise ValueError("innermost dimension must be contiguous")
    if positions.dtype != torch.int64 or slot_mapping.dtype != torch.int64:
        raise ValueError("positions and slot_mapping must be int64")
    if positions.dim() != 1:
        raise ValueError("positions must be 1-D (mrope is not supported)")
    if cos_sin_cache.dtype != query.dtype or key.dtype != query.dtype or value.dtype != query.dtype:
        raise ValueError("query, key, value and cos_sin_cache must share one 16-bit dtype")
    src_dt = _DTYPE.get(query.dtype)
    if src_dt not in (_cabi.BF16, _cabi.F16):
        raise ValueError(f"unsupported dtype {query.dtype}")
    p = _cabi.RopeCacheParams()
    if kv_cache_dtype.startswith("fp8"):
        if kv_cache_dtype not in ("fp8", "fp8_e4m3") or key_cache.element_size() != 1:
            raise ValueError(f"unsupported kv_cache_dtype {kv_cache_dtype!r} / cache element size")
        ks, vs = _scalar_tensor(k_scale, query.device), _scalar_tensor(v_scale, query.device)
        p.kv_dtype, p.k_scale, p.v_scale = _cabi.FP8_E4M3, ks.data_ptr(), vs.data_ptr()
    else:
        if key_cache.dtype != query.dtype:
            raise ValueError(f"kv cache dtype {key_cache.dtype} != {query.dtype} with kv_cache_dtype={kv_cache_dtype!r}")
        p.kv_dtype, p.k_scale, p.v_scale = src_dt, None, None
    p.q, p.k, p.v = query.data_ptr(), key.data_ptr(), value.data_ptr()
    p.positions, p.cos_sin_cache = positions.data_ptr(), cos_sin_cache.data_ptr()
    p.k_cache, p.v_cache, p.slot_mapping = key_cache.data_ptr(), value_cache.data_ptr(), slot_mapping.data_ptr()
    p.rope_tokens, p.cache_tokens = min(positions.shape[0], query.shape[0]), min(slot_mapping.shape[0], key.shape[0])
    p.q_stride0, p.q_stride1 = query.stride(0), query.stride(1)
    p.k_stride0, p.k_stride1 = key.stride(0), key.stride(1)
    p.v_stride0, p.v_stride1 = value.stride(0), value.stride(1)
    p.kc_block_stride, p.kc_slot_stride, p.kc_head_stride = key_cache.stride(0), key_cache.stride(1), key_cache.stride(2)
    p.vc_block_stride, p.vc_slot_stride, p.vc_head_stride = value_cache.stride(0), value_cache.stride(1), value_cache.stride(2)
    p.num_q_heads, p.num_kv_heads, p.head_size = query.shape[1], key.shape[1], query.shape[2]
    p.rot_dim, p.block_size, p.is_neox, p.src_dtype = cos_sin_cache.shape[1], key_cache.shape[1], int(bool(is_neox)), src_dt
    _cabi.check(_cabi.load_library().b200attn_rope_and_cache_flash(ctypes.byref(p), _stream_ptr(query)))

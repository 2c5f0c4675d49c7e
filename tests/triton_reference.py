"""The reference Triton kernels, compiled natively on the GPU box (test/bench checker only).

Two sources, both run on the same device tensors our kernels read:

* `ReferenceFileKernels` - the reference's OWN file, ibm-triton-lib/ibm_triton_lib/kernels/
  triton_unified_attention.py, loaded by path from /root/reference or from the copy `__graft_entry__.build()`
  leaves in the git-ignored baseline/_ref/ (which travels to the GPU box), with the
  `triton_dejavu` autotuner stubbed to fixed tiles (it contributes no arithmetic).  This is the
  "reference Triton kernels" of BASELINE.json's tolerance clause.
* `TritonReference` - the upstream copy of the same kernel that the installed vLLM ships, which also carries the
  later additions (sinks, qq_bias, output_scale) the reference file does not have, and vLLM's own launch heuristics
  (what `scripts/kbench.py --impl triton` and bench.py's same-GPU comparator time).

`/root/reference` does not exist on the GPU box, but the vLLM installed in the
image ships the upstream copy of the same kernel the reference plugin carries
(`vllm/v1/attention/ops/triton_unified_attention.py` — the reference file
ibm-triton-lib/ibm_triton_lib/kernels/triton_unified_attention.py:275-836 was
upstreamed from it; the 2D/3D split, tile loop, softcap/ALiBi/window handling and
`reduce_segments` are the same code).  BASELINE.json asks for bf16/FP8 parity
"with the reference Triton kernels": this is them, run on the same tensors.

Never imported by the product; `scripts/kbench.py --impl triton` uses it to time
the reference kernels beside ours.
"""
from __future__ import annotations

import os
import sys

import torch

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
SOURCE = "vllm.v1.attention.ops.triton_unified_attention (upstream copy shipped by the installed vLLM)"


def reference_file_available() -> str | None:
    """None if the reference's own kernel file can run here, else the reason."""
    if not torch.cuda.is_available():
        return "no GPU"
    from oracle import reference_runner as rr

    if not rr.reference_available():
        return "reference kernel file neither under /root/reference nor in baseline/_ref (run __graft_entry__.build())"
    if os.environ.get("TRITON_INTERPRET") == "1":
        return "TRITON_INTERPRET=1 is set"
    return None


class ReferenceFileKernels:
    """`unified_attention` of the reference file (its launcher, :839-1030): 2D kernel for batches with a multi-token
    query, else the 3D kernel with 16 segments + reduce_segments."""

    def __init__(self, num_queries_per_kv: int, block_size: int):
        from oracle import reference_runner as rr

        bm = max(16, 1 << (max(num_queries_per_kv, 1) - 1).bit_length())
        self.mod = rr.load_reference_kernels(bm, max(32, block_size), interpret=False)
        self.source = "reference file " + rr._KERNEL_FILE

    def __call__(self, q, k_cache, v_cache, out, cu_seqlens_q, max_seqlen_q, seq_lens, max_seqlen_k, scale, block_table,
                 window=0, softcap=0.0, alibi_slopes=None, k_scale=None, v_scale=None):
        one = torch.ones(1, dtype=torch.float32, device=q.device)
        self.mod.unified_attention(
            q=q, k=k_cache, v=v_cache, out=out, cu_seqlens_q=cu_seqlens_q, max_seqlen_q=max_seqlen_q, seqused_k=seq_lens,
            max_seqlen_k=max_seqlen_k, avg_seqlen_q=max(1, q.shape[0] // max(1, seq_lens.shape[0])),
            avg_seqlen_k=max_seqlen_k, softmax_scale=scale, causal=True,
            window_size=(window - 1, 0) if window > 0 else (-1, -1), block_table=block_table, softcap=softcap,
            q_descale=None, k_descale=one if k_scale is None else k_scale, v_descale=one if v_scale is None else v_scale,
            alibi_slopes=alibi_slopes)
        return out


def available() -> str | None:
    """None if the Triton reference can run here, else the reason."""
    if not torch.cuda.is_available():
        return "no GPU"
    try:
        from vllm.v1.attention.ops.triton_unified_attention import unified_attention  # noqa: F401
    except Exception as e:  # pragma: no cover - depends on the image
        return f"vllm triton_unified_attention not importable: {type(e).__name__}: {e}"
    return None


class TritonReference:
    """Holds the 3D-path scratch tensors the way vLLM's builder allocates them
    (site-packages/vllm/v1/attention/backends/triton_attn.py:168-201)."""

    NUM_SEGMENTS = 16

    def __init__(self, max_seqs: int, num_q_heads: int, head_size: int, device="cuda"):
        pad = 1 << (head_size - 1).bit_length()
        self.max_seqs = max_seqs
        self.segm_output = torch.empty(max_seqs, num_q_heads, self.NUM_SEGMENTS, pad, dtype=torch.float32, device=device)
        self.segm_max = torch.empty(max_seqs, num_q_heads, self.NUM_SEGMENTS, dtype=torch.float32, device=device)
        self.segm_expsum = torch.empty_like(self.segm_max)

    def __call__(self, q, k_cache, v_cache, out, cu_seqlens_q, max_seqlen_q, seq_lens, max_seqlen_k, scale,
                 block_table, window=0, softcap=0.0, alibi_slopes=None, k_scale=None, v_scale=None, use_3d=True, sinks=None,
                 k_scale_cache=None, v_scale_cache=None):
        from vllm.v1.attention.ops.triton_unified_attention import unified_attention

        from vllm.v1.kv_cache_interface import KVQuantMode

        dev = q.device
        # fp8 caches: per-tensor dequant `fp8 -> f32 * scale -> q dtype` (upstream _cast_kv_tile; the
        # reference file does the same at triton_unified_attention.py:434-455 when Q is not fp8)
        mode = KVQuantMode.FP8_PER_TENSOR if k_cache.dtype == torch.float8_e4m3fn else KVQuantMode.NONE
        one = torch.ones(1, dtype=torch.float32, device=dev)
        kw = {}
        if k_scale_cache is not None:  # per-(token, head) scales: upstream modes 2 (int8) and 3 (e4m3)
            mode = KVQuantMode.INT8_PER_TOKEN_HEAD if k_cache.dtype == torch.int8 else KVQuantMode.FP8_PER_TOKEN_HEAD
            kw.update(k_scale_cache=k_scale_cache, v_scale_cache=v_scale_cache)
        if use_3d and max_seqlen_q == 1 and seq_lens.shape[0] <= self.max_seqs:
            kw.update(seq_threshold_3D=self.max_seqs, num_par_softmax_segments=self.NUM_SEGMENTS,
                      softmax_segm_output=self.segm_output, softmax_segm_max=self.segm_max,
                      softmax_segm_expsum=self.segm_expsum)
        unified_attention(
            q=q, k=k_cache, v=v_cache, out=out, cu_seqlens_q=cu_seqlens_q, max_seqlen_q=max_seqlen_q,
            seqused_k=seq_lens, max_seqlen_k=max_seqlen_k, softmax_scale=scale, causal=True,
            window_size=(window - 1, 0) if window > 0 else (-1, -1), block_table=block_table, softcap=softcap,
            q_descale=None, k_descale=one if k_scale is None else k_scale, v_descale=one if v_scale is None else v_scale,
            alibi_slopes=alibi_slopes, kv_quant_mode=mode, sinks=sinks, **kw,
        )
        return out

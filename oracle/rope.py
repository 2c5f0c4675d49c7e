"""CPU oracle (test infrastructure, not product): rotary embedding as vLLM applies it.

Restates `apply_token_rotary_embedding` (reference vllm/csrc/pos_encoding_kernels.cu:
x' = x*cos - y*sin, y' = y*cos + x*sin evaluated in `scalar_t`, i.e. every operation
rounds to the tensor's 16-bit type) and the torch-native `forward_static`
(reference vllm/vllm/model_executor/layers/rotary_embedding/base.py).  Used by the
tests of the fused rope + cache-write kernel (K5).

Parity status: pinned on the GPU box against `torch.ops._C.rotary_embedding` of the
installed vLLM (tests/test_gpu_rope_cache.py, bit-exact); no golden vector exists for
it in the reference repository (the reference path never fuses the two).
"""
from __future__ import annotations

import torch


def rotary_embedding_ref(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, cos_sin_cache: torch.Tensor,
                         is_neox: bool):
    """query [T, Hq, D], key [T, Hkv, D] -> rotated copies (same dtype); arithmetic in the tensors' dtype."""
    rot = cos_sin_cache.shape[1]
    half = rot // 2
    cs = cos_sin_cache[positions.long()]  # [T, rot]
    cos, sin = cs[:, None, :half], cs[:, None, half:]

    def apply(x: torch.Tensor) -> torch.Tensor:
        out = x.clone()
        if is_neox:
            a, b = x[..., :half], x[..., half:rot]
            out[..., :half] = a * cos - b * sin
            out[..., half:rot] = b * cos + a * sin
        else:
            a, b = x[..., 0:rot:2], x[..., 1:rot:2]
            out[..., 0:rot:2] = a * cos - b * sin
            out[..., 1:rot:2] = b * cos + a * sin
        return out

    return apply(query), apply(key)

"""CPU oracle (test infrastructure, not product): rotary embedding as vLLM applies it.

Restates `apply_token_rotary_embedding` (reference vllm/csrc/pos_encoding_kernels.cu:10-36:
x' = x*cos - y*sin, y' = y*cos + x*sin on pairs (j, j + rot/2) [NeoX] or (2j, 2j+1) [GPT-J])
with the arithmetic the vLLM this plugs into actually performs: fp32, rounded once to the
16-bit type (scripts/probe_rope_rounding.py, run on the GPU box against vLLM 0.22's
compiled op: 0 mismatches; per-operation rounding in `scalar_t`: 32 % mismatches).
Used by the tests of the fused rope + cache-write kernel (K5).

Parity status: pinned on the GPU box against `torch.ops._C.rotary_embedding` of the
installed vLLM (tests/test_gpu_rope_cache.py, bit-exact); no golden vector exists for
it in the reference repository (the reference path never fuses the two).
"""
from __future__ import annotations

import torch


def rotary_embedding_ref(positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor, cos_sin_cache: torch.Tensor,
                         is_neox: bool):
    """query [T, Hq, D], key [T, Hkv, D] -> rotated copies (same dtype); fp32 arithmetic (fp64 tensors
    stay fp64), one rounding."""
    rot = cos_sin_cache.shape[1]
    half = rot // 2
    cs = cos_sin_cache[positions.long()]  # [T, rot]
    wide = torch.float64 if query.dtype == torch.float64 else torch.float32
    cos, sin = cs[:, None, :half].to(wide), cs[:, None, half:].to(wide)

    def apply(x: torch.Tensor) -> torch.Tensor:
        out = x.clone()
        xw = x.to(wide)
        if is_neox:
            a, b = xw[..., :half], xw[..., half:rot]
            out[..., :half] = (a * cos - b * sin).to(x.dtype)
            out[..., half:rot] = (b * cos + a * sin).to(x.dtype)
        else:
            a, b = xw[..., 0:rot:2], xw[..., 1:rot:2]
            out[..., 0:rot:2] = (a * cos - b * sin).to(x.dtype)
            out[..., 1:rot:2] = (b * cos + a * sin).to(x.dtype)
        return out

    return apply(query), apply(key)

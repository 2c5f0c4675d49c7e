"""Algorithmic bytes / FLOPs of the hot path (host ints only), SURVEY.md §8(d).
These are the numerators of every roofline fraction bench.py reports."""
from __future__ import annotations

import math
from typing import Optional, Sequence


def decode_bytes(seq_lens: Sequence[int], num_q_heads: int, num_kv_heads: int, head_size: int, block_size: int,
                 kv_elem_bytes: int, q_elem_bytes: int = 2, num_tokens: Optional[int] = None) -> int:
    """sum_b seq_len_b*Hkv*D*(|K|+|V|) + T*Hq*D*|q| + T*Hq*D*2 (out) + block-table ints + lens.
    Split-KV partials are implementation overhead and are NOT counted."""
    B = len(seq_lens)
    T = B if num_tokens is None else num_tokens
    kv = sum(int(s) for s in seq_lens) * num_kv_heads * head_size * 2 * kv_elem_bytes
    qo = T * num_q_heads * head_size * (q_elem_bytes + 2)
    bt = sum(math.ceil(int(s) / block_size) for s in seq_lens) * 4
    return kv + qo + bt + (2 * B + 1) * 4


def cache_write_bytes(num_tokens: int, num_kv_heads: int, head_size: int, src_bytes: int, kv_bytes: int) -> int:
    """T*Hkv*D*(2*src + K + V) + 8T."""
    return num_tokens * num_kv_heads * head_size * (2 * src_bytes + 2 * kv_bytes) + 8 * num_tokens


def attention_flops(query_lens: Sequence[int], seq_lens: Sequence[int], num_q_heads: int, head_size: int) -> int:
    """sum_b 4*Hq*D*(q_b*ctx_b + q_b(q_b+1)/2): QK^T + PV, causal, 2 FLOP per MAC."""
    total = 0
    for ql, sl in zip(query_lens, seq_lens):
        ctx = sl - ql
        total += 4 * num_q_heads * head_size * (ql * ctx + ql * (ql + 1) // 2)
    return total

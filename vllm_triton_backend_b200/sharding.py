"""KV-head sharding of the attention path across the GPUs of one box.

The path shards the way vLLM tensor parallelism shards it (reference
vllm/vllm/config/model.py:1348-1363): rank r of G owns Hq/G query heads and
max(1, Hkv/G) KV heads plus its own KV-cache shard; block tables, sequence
lengths and slot mappings are replicated.  Attention needs no collective; an
all_gather of the per-rank outputs exists only so a harness can compare the
sharded result with the unsharded one.
"""
from __future__ import annotations

from typing import Tuple

import torch


def shard_heads(rank: int, world: int, num_q_heads: int, num_kv_heads: int) -> Tuple[slice, slice]:
    """(query-head slice, kv-head slice) owned by `rank`."""
    if num_q_heads % world:
        raise ValueError(f"{num_q_heads} query heads do not divide over {world} ranks")
    hq = num_q_heads // world
    q_slice = slice(rank * hq, (rank + 1) * hq)
    if num_kv_heads >= world:
        if num_kv_heads % world:
            raise ValueError(f"{num_kv_heads} kv heads do not divide over {world} ranks")
        hkv = num_kv_heads // world
        kv_slice = slice(rank * hkv, (rank + 1) * hkv)
    else:
        # fewer KV heads than ranks: each KV head is replicated over world/num_kv_heads ranks
        if world % num_kv_heads:
            raise ValueError(f"{world} ranks are not a multiple of {num_kv_heads} kv heads")
        h = rank // (world // num_kv_heads)
        kv_slice = slice(h, h + 1)
    return q_slice, kv_slice


def gather_head_shards(local_out: torch.Tensor, world: int) -> torch.Tensor:
    """[T, Hq/G, D] per rank -> [T, Hq, D] (verification only; NCCL or gloo)."""
    import torch.distributed as dist

    if world == 1:
        return local_out
    parts = [torch.empty_like(local_out) for _ in range(world)]
    dist.all_gather(parts, local_out.contiguous())
    return torch.cat(parts, dim=1)

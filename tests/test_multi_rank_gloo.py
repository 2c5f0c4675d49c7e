"""N>1 path on CPU: world_size-2 gloo ranks shard the heads, run the (oracle)
attention on their shard and all_gather for verification — the same plumbing
bench.py / the GPU harness use with NCCL.  The CUDA kernels are
not involved; this covers the host-side sharding logic."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import make_paged_inputs
from oracle.paged_attention import unified_attention_ref
from vllm_triton_backend_b200.sharding import gather_head_shards, shard_heads


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, hq, hkv, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        inp = make_paged_inputs([(1, 300), (1, 77), (5, 40)], hq, hkv, 64, 16, dtype=torch.float32, seed=4)
        full = unified_attention_ref(inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"],
                                     inp["block_table"], 0.125)
        qs, ks = shard_heads(rank, world, hq, hkv)
        local = unified_attention_ref(inp["q"][:, qs].contiguous(), inp["k_cache"][:, :, ks].contiguous(),
                                      inp["v_cache"][:, :, ks].contiguous(), inp["cu_seqlens_q"], inp["seq_lens"],
                                      inp["block_table"], 0.125)
        gathered = gather_head_shards(local, world)
        ok = torch.equal(gathered, full)  # per-head math is unchanged by sharding: bit-identical
        flag = torch.tensor([1 if ok else 0])
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            ret.put(int(flag))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("hq,hkv", [(8, 2), (8, 1)])
def test_head_sharded_attention_equals_unsharded(hq, hkv):
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, hq, hkv, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
        assert p.exitcode == 0
    assert ret.get(timeout=5) == 1

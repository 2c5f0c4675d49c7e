#!/usr/bin/env python
"""Multi-GPU check of the head-sharded path, run under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 \
        --master-port 29511 scripts/verify_head_shards.py

Every rank builds the same synthetic C4-shaped batch (FP8 KV, 64q/8kv heads, small
context), runs the CUDA path on ITS heads only (cache write + decode attention, then
a mixed prefill batch), the outputs are all_gathered over NCCL and rank 0 compares
them with the unsharded run of the same kernels and with the CPU oracle.  No
collective is part of the data path; NCCL is used for verification only.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from helpers import BF16_ATOL, BF16_RTOL, FP8_ATOL, make_paged_inputs  # noqa: E402
from vllm_triton_backend_b200 import ops  # noqa: E402
from vllm_triton_backend_b200.sharding import gather_head_shards, shard_heads  # noqa: E402


def run(inp, hq_slice, hkv_slice, dev, fp8, ks, vs, scale):
    q = inp["q"][:, hq_slice].contiguous().to(dev)
    kc = inp["k_cache"][:, :, hkv_slice].contiguous().to(dev)
    vc = inp["v_cache"][:, :, hkv_slice].contiguous().to(dev)
    out = torch.empty_like(q)
    sc = (torch.tensor([ks], dtype=torch.float32, device=dev), torch.tensor([vs], dtype=torch.float32, device=dev))
    ops.unified_attention(
        q=q, k=kc, v=vc, out=out, cu_seqlens_q=inp["cu_seqlens_q"].to(dev), max_seqlen_q=max(inp["q_lens"]),
        seqused_k=inp["seq_lens"].to(dev), max_seqlen_k=max(inp["kv_lens"]), softmax_scale=scale, causal=True,
        window_size=(-1, -1), block_table=inp["block_table"].to(dev), softcap=0, q_descale=None,
        k_descale=sc[0] if fp8 else None, v_descale=sc[1] if fp8 else None)
    torch.cuda.synchronize()
    return out


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    Hq, Hkv, D, bs = 64, 8, 128, 16
    scale = D ** -0.5
    results = []
    for name, lens, fp8 in [("decode_fp8", [(1, 700), (1, 33), (1, 4096), (1, 1500)], True),
                            ("mixed_bf16", [(1, 900), (300, 300), (64, 1064), (1, 77)], False)]:
        ks, vs = (0.37, 0.61) if fp8 else (1.0, 1.0)
        inp = make_paged_inputs(lens, Hq, Hkv, D, bs, seed=9, kv_dtype=torch.float8_e4m3fn if fp8 else None,
                                k_scale=ks, v_scale=vs)
        qs, kvs = shard_heads(rank, world, Hq, Hkv)
        mine = run(inp, qs, kvs, dev, fp8, ks, vs, scale)
        gathered = gather_head_shards(mine, world)
        if rank == 0:
            full = run(inp, slice(0, Hq), slice(0, Hkv), dev, fp8, ks, vs, scale)
            from oracle.paged_attention import unified_attention_ref
            ref = unified_attention_ref(inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"],
                                        inp["block_table"], scale, k_scale=ks, v_scale=vs, out_dtype=torch.float32)
            err = (gathered.float().cpu() - ref).abs()
            tol = (FP8_ATOL if fp8 else BF16_ATOL) + BF16_RTOL * ref.abs()
            # Per-head arithmetic is the same, but the split-KV partition (hence the fp32 summation
            # order of a split item) depends on how many heads a GPU holds: equal up to output rounding.
            d = float((gathered.float() - full.float()).abs().max())
            results.append(dict(case=name, world=world, sharded_equals_unsharded_bits=bool(torch.equal(gathered, full)),
                                sharded_vs_unsharded_max_abs=d, max_abs_err_vs_oracle=float(err.max()),
                                within_tolerance=bool((err <= tol).all()) and d <= 8e-3))
    if rank == 0:
        ok = all(r["within_tolerance"] for r in results)
        print(json.dumps(dict(verify_head_shards="ok" if ok else "FAILED", backend="nccl" if world > 1 else "none",
                              results=results)))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    sys.exit(0)


if __name__ == "__main__":
    main()

"""Generates tests/golden/pth_*.npz: per-(token, head) quantised KV caches (test infrastructure).

The algorithm of this mode is not in `/root/reference`: it lives in the vLLM the reference plugs into (third-party
dependency; pinned here: the installed vllm 0.22.0).  This script runs vLLM's own Triton kernels with Triton's CPU
interpreter:

    TRITON_INTERPRET=1 python oracle/make_golden_pth.py

  * `triton_reshape_and_cache_flash_per_token_head_quant` (vllm/v1/attention/ops/triton_reshape_and_cache_flash.py:247)
    -> `k_cache`, `v_cache` (stored as uint8 bit patterns), `k_scale_cache`, `v_scale_cache`
  * `unified_attention(..., kv_quant_mode=INT8/FP8_PER_TOKEN_HEAD, k_scale_cache=, v_scale_cache=)`
    (vllm/v1/attention/ops/triton_unified_attention.py:771) on the caches it just wrote -> `out_triton`

Known defects of the Triton 3.6 INTERPRETER (not of the compiled kernel, which uses the hardware conversion): its
fp32 -> e4m3 conversion drops the carry when rounding moves a value up into the next binade (124.3 -> 64 instead of
128), rounds exact ties away from zero (84 -> 88, not 80) and truncates subnormal results.  The fp8 cache fixtures
therefore differ from round-to-nearest-even in under 2 % of the codes; tests/test_oracle_per_token_head.py accepts a
mismatch only at those three kinds of place (the GPU suite compares with the compiled Triton kernel, which has no such defect).  The int8
fixtures are exact.
"""
from __future__ import annotations

import os
import sys

os.environ.setdefault("TRITON_INTERPRET", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

# name, (q_len, kv_len) per sequence, Hq, Hkv, D, block_size, num_blocks, kv dtype, window, softcap
CASES = [
    ("int8_mixed", [(1, 67), (5, 18), (33, 97), (16, 16)], 8, 2, 64, 16, 24, "int8", 0, 0.0),
    ("int8_decode_window_softcap", [(1, 130), (1, 40), (1, 9), (1, 1)], 8, 2, 64, 16, 24, "int8", 24, 30.0),
    ("fp8_mixed", [(1, 50), (7, 39), (20, 20)], 4, 2, 128, 16, 16, "fp8", 0, 0.0),
    ("fp8_decode", [(1, 200), (1, 64), (1, 17)], 16, 4, 128, 16, 32, "fp8", 0, 0.0),
]


def main():
    from vllm.v1.attention.ops.triton_reshape_and_cache_flash import (
        triton_reshape_and_cache_flash_per_token_head_quant as write_quant,
    )
    from vllm.v1.attention.ops.triton_unified_attention import unified_attention
    from vllm.v1.kv_cache_interface import KVQuantMode

    os.makedirs(GOLDEN_DIR, exist_ok=True)
    for name, lens, Hq, Hkv, D, bs, NB, kvd, window, softcap in CASES:
        g = torch.Generator().manual_seed(sum(ord(c) for c in name))
        kv_dtype = torch.int8 if kvd == "int8" else torch.float8_e4m3fn
        mode = KVQuantMode.INT8_PER_TOKEN_HEAD if kvd == "int8" else KVQuantMode.FP8_PER_TOKEN_HEAD
        q_lens = [a for a, _ in lens]
        kv_lens = [b for _, b in lens]
        T = sum(q_lens)
        q = (torch.rand(T, Hq, D, generator=g) * 2 - 1).half()
        # unique pages per sequence (a cache that is WRITTEN cannot share pages between sequences)
        pages = [(kl + bs - 1) // bs for kl in kv_lens]
        assert sum(pages) <= NB
        perm = torch.randperm(NB, generator=g)[: sum(pages)].to(torch.int32)
        bt = torch.zeros(len(lens), max(pages), dtype=torch.int32)
        slots = []
        o = 0
        for i, n in enumerate(pages):
            bt[i, :n] = perm[o : o + n]
            o += n
            for t in range(kv_lens[i]):
                slots.append(int(bt[i, t // bs]) * bs + t % bs)
        n_kv = len(slots)
        # rows with very different magnitudes, so that a per-tensor scale would not do
        mag = torch.exp(torch.rand(n_kv, Hkv, 1, generator=g) * 6 - 3)
        key = ((torch.rand(n_kv, Hkv, D, generator=g) * 2 - 1) * mag).half()
        value = ((torch.rand(n_kv, Hkv, D, generator=g) * 2 - 1) * mag).half()
        slot_mapping = torch.tensor(slots, dtype=torch.int64)
        slot_mapping[3] = -1  # one padding slot (the kernel skips it, :171-173)
        kc = torch.zeros(NB, bs, Hkv, D, dtype=kv_dtype)
        vc = torch.zeros_like(kc)
        ks = torch.ones(NB, bs, Hkv, dtype=torch.float32)
        vs = torch.ones(NB, bs, Hkv, dtype=torch.float32)
        write_quant(key, value, kc, vc, ks, vs, slot_mapping)
        out = torch.zeros_like(q)
        cu = torch.tensor([0] + list(np.cumsum(q_lens)), dtype=torch.int32)
        sl = torch.tensor(kv_lens, dtype=torch.int32)
        scale = D ** -0.5
        unified_attention(
            q=q, k=kc, v=vc, out=out, cu_seqlens_q=cu, max_seqlen_q=max(q_lens), seqused_k=sl,
            max_seqlen_k=max(kv_lens), softmax_scale=scale, causal=True,
            window_size=(window - 1, 0) if window > 0 else (-1, -1), block_table=bt, softcap=softcap,
            q_descale=None, k_descale=None, v_descale=None, kv_quant_mode=mode, k_scale_cache=ks, v_scale_cache=vs,
        )
        path = os.path.join(GOLDEN_DIR, f"pth_{name}.npz")
        np.savez_compressed(
            path, q=q.numpy(), key=key.numpy(), value=value.numpy(), slot_mapping=slot_mapping.numpy(),
            k_cache=kc.view(torch.uint8).numpy(), v_cache=vc.view(torch.uint8).numpy(), k_scale_cache=ks.numpy(),
            v_scale_cache=vs.numpy(), block_table=bt.numpy(), cu_seqlens_q=cu.numpy(), seq_lens=sl.numpy(),
            scale=np.float32(scale), window=np.int32(window), softcap=np.float32(softcap), kv_dtype=kvd,
            out_triton=out.numpy(), source="vllm 0.22.0 Triton kernels, TRITON_INTERPRET=1",
        )
        print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()

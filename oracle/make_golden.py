"""Generates tests/golden/*.npz by running the REFERENCE itself (test infrastructure).

Run in the build container (where /root/reference is mounted):

    TRITON_INTERPRET=1 python oracle/make_golden.py

For every case it stores the seeded inputs and
  * `out_triton`  — the reference Triton kernels (2D path, or 3D + reduce_segments)
                    executed by Triton's CPU interpreter through the reference
                    launcher `unified_attention`
                    (ibm-triton-lib/ibm_triton_lib/kernels/triton_unified_attention.py:839),
  * `out_torch`   — the reference's torch function `ref_paged_attn`
                    (scripts/vllm_utils.py:433) where it supports the feature,
and for the cache-write case the result of the reference's
`ref_reshape_and_cache_flash` (scripts/vllm_utils.py:377).
The GPU box has no /root/reference; tests read only the committed .npz files.
"""
from __future__ import annotations

import os
import sys

os.environ.setdefault("TRITON_INTERPRET", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from oracle.reference_runner import load_reference_kernels, load_reference_torch_utils  # noqa: E402

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

# name, (q_len, kv_len) per sequence, Hq, Hkv, D, block_size, num_blocks, dtype, window, softcap, alibi, (BLOCK_M, BLOCK_N)
CASES = [
    ("decode3d_f32", [(1, 130), (1, 64), (1, 17), (1, 1)], 8, 2, 128, 16, 48, "float32", 0, 0.0, False, (16, 32)),
    ("decode3d_f16_gqa4", [(1, 512), (1, 300), (1, 33), (1, 257)], 16, 4, 128, 16, 96, "float16", 0, 0.0, False, (16, 32)),
    ("decode3d_f32_alibi", [(1, 77), (1, 16), (1, 5)], 8, 2, 64, 16, 24, "float32", 0, 0.0, True, (16, 32)),
    ("decode3d_f32_window", [(1, 130), (1, 40), (1, 9)], 8, 2, 64, 16, 24, "float32", 24, 0.0, False, (16, 32)),
    ("mixed2d_f32", [(1, 67), (5, 18), (33, 97), (16, 16)], 8, 2, 64, 16, 32, "float32", 0, 0.0, False, (16, 32)),
    ("mixed2d_f32_softcap_window", [(1, 67), (5, 18), (33, 97), (16, 16)], 8, 2, 64, 16, 32, "float32", 32, 50.0, False, (16, 32)),
    ("mixed2d_f32_alibi", [(3, 40), (20, 20), (1, 9)], 4, 4, 64, 16, 16, "float32", 0, 0.0, True, (16, 16)),
    ("mixed2d_f16_d128_bs32", [(1, 100), (40, 72), (64, 64)], 8, 2, 128, 32, 16, "float16", 0, 0.0, False, (32, 32)),
    ("prefill2d_f32_m64", [(70, 70), (129, 129)], 4, 2, 64, 16, 24, "float32", 0, 0.0, False, (64, 64)),
]


def build_inputs(name, lens, Hq, Hkv, D, bs, NB, dtype, seed=0):
    g = torch.Generator().manual_seed(seed + sum(ord(c) for c in name))
    td = getattr(torch, dtype)
    q_lens = [a for a, _ in lens]
    kv_lens = [b for _, b in lens]
    T = sum(q_lens)
    # values live on the fp16 grid so the fixtures can be stored as fp16 losslessly
    q = (torch.rand(T, Hq, D, generator=g) * 2 - 1).half().to(td)
    k_cache = (torch.rand(NB, bs, Hkv, D, generator=g) * 2 - 1).half().to(td)
    v_cache = (torch.rand(NB, bs, Hkv, D, generator=g) * 2 - 1).half().to(td)
    max_blocks = max((kl + bs - 1) // bs for kl in kv_lens)
    # with replacement, as the reference tests draw them (scripts/test.py:210-214)
    block_table = torch.randint(0, NB, (len(lens), max_blocks), generator=g, dtype=torch.int32)
    cu = torch.tensor([0] + list(np.cumsum(q_lens)), dtype=torch.int32)
    seq_lens = torch.tensor(kv_lens, dtype=torch.int32)
    return q, k_cache, v_cache, block_table, cu, seq_lens


def main():
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    ref_utils = load_reference_torch_utils()
    for (name, lens, Hq, Hkv, D, bs, NB, dtype, window, softcap, alibi, (bm, bn)) in CASES:
        mod = load_reference_kernels(bm, bn)
        q, kc, vc, bt, cu, sl = build_inputs(name, lens, Hq, Hkv, D, bs, NB, dtype)
        q_lens = [a for a, _ in lens]
        kv_lens = [b for _, b in lens]
        scale = D ** -0.5
        slopes = None
        if alibi:
            slopes = torch.tensor([2.0 ** (-(i + 1) * 8.0 / Hq) for i in range(Hq)], dtype=torch.float32)
        out = torch.zeros_like(q)
        one = torch.ones(1, dtype=torch.float32)
        mod.unified_attention(
            q=q, k=kc, v=vc, out=out, cu_seqlens_q=cu, max_seqlen_q=max(q_lens), seqused_k=sl,
            max_seqlen_k=max(kv_lens), avg_seqlen_q=int(np.mean(q_lens)), avg_seqlen_k=int(np.mean(kv_lens)),
            softmax_scale=scale, causal=True, window_size=(window - 1, 0) if window > 0 else (-1, -1),
            block_table=bt, softcap=softcap, q_descale=None,
            k_descale=one.expand(len(lens), Hkv), v_descale=one.expand(len(lens), Hkv), alibi_slopes=slopes,
        )
        save = dict(
            q=q.half().numpy(), k_cache=kc.half().numpy(), v_cache=vc.half().numpy(), dtype=np.array(dtype),
            block_table=bt.numpy(), cu_seqlens_q=cu.numpy(),
            seq_lens=sl.numpy(), scale=np.float32(scale), window=np.int32(window), softcap=np.float32(softcap),
            out_triton=out.numpy(), path=np.array("3d" if max(q_lens) == 1 else "2d"),
        )
        if slopes is not None:
            save["alibi_slopes"] = slopes.numpy()
        else:
            # the reference torch function has no ALiBi argument
            o2 = ref_utils.ref_paged_attn(
                query=q.clone().float(), key_cache=kc.float(), value_cache=vc.float(), query_lens=q_lens,
                kv_lens=kv_lens, block_tables=bt, scale=scale,
                sliding_window=window if window > 0 else None, soft_cap=softcap if softcap > 0 else None,
            )
            save["out_torch"] = o2.numpy()
        np.savez(os.path.join(GOLDEN_DIR, f"attn_{name}.npz"), **save)
        print(f"{name}: path={save['path']} out_triton max|x|={np.abs(save['out_triton']).max():.4f}", flush=True)

    # cache write through the reference's torch function
    g = torch.Generator().manual_seed(7)
    T, Hkv, D, bs, NB = 42, 8, 64, 16, 8
    key = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
    value = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
    slots = torch.randperm(NB * bs, generator=g)[:T].to(torch.int64)
    kc = torch.zeros(NB, bs, Hkv, D, dtype=torch.bfloat16)
    vc = torch.zeros(NB, bs, Hkv, D, dtype=torch.bfloat16)
    ref_utils.ref_reshape_and_cache_flash(key, value, kc, vc, slots, bs, T)
    np.savez(
        os.path.join(GOLDEN_DIR, "cache_write_bf16.npz"),
        key=key.view(torch.int16).numpy(), value=value.view(torch.int16).numpy(), slot_mapping=slots.numpy(),
        k_cache=kc.view(torch.int16).numpy(), v_cache=vc.view(torch.int16).numpy(), block_size=np.int32(bs),
    )
    print("cache_write_bf16 done")


if __name__ == "__main__":
    main()

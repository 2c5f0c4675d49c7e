"""Shared test helpers: golden loading, synthetic paged-KV inputs, tolerance check."""
from __future__ import annotations

import glob
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# BASELINE.json north_star tolerances
BF16_ATOL, BF16_RTOL = 2e-2, 1e-2
FP8_ATOL = 6e-2


def golden_attn_files():
    return sorted(glob.glob(os.path.join(GOLDEN_DIR, "attn_*.npz")))


def load_golden_attn(path, dtype=None):
    z = np.load(path)
    td = getattr(torch, str(z["dtype"])) if dtype is None else dtype
    d = dict(
        name=os.path.basename(path)[5:-4],
        q=torch.from_numpy(z["q"]).to(td),
        k_cache=torch.from_numpy(z["k_cache"]).to(td),
        v_cache=torch.from_numpy(z["v_cache"]).to(td),
        block_table=torch.from_numpy(z["block_table"]),
        cu_seqlens_q=torch.from_numpy(z["cu_seqlens_q"]),
        seq_lens=torch.from_numpy(z["seq_lens"]),
        scale=float(z["scale"]),
        window=int(z["window"]),
        softcap=float(z["softcap"]),
        out_triton=torch.from_numpy(z["out_triton"]).float(),
        out_torch=torch.from_numpy(z["out_torch"]).float() if "out_torch" in z.files else None,
        alibi_slopes=torch.from_numpy(z["alibi_slopes"]) if "alibi_slopes" in z.files else None,
        path=str(z["path"]),
        dtype=td,
    )
    return d


def assert_close(actual: torch.Tensor, expected: torch.Tensor, atol: float, rtol: float, what: str = ""):
    a, e = actual.float().cpu(), expected.float().cpu()
    assert a.shape == e.shape, f"{what}: shape {tuple(a.shape)} vs {tuple(e.shape)}"
    both_nan = torch.isnan(a) & torch.isnan(e)
    err = (a - e).abs()
    err = torch.where(both_nan, torch.zeros_like(err), err)
    bound = atol + rtol * e.abs()
    bad = ~(err <= bound) & ~both_nan
    assert not bad.any(), (
        f"{what}: {int(bad.sum())}/{bad.numel()} outside atol={atol} rtol={rtol}; "
        f"max abs err {float(torch.nan_to_num(err).max()):.3e}"
    )


def make_paged_inputs(
    q_kv_lens, num_q_heads, num_kv_heads, head_size, block_size, dtype=torch.bfloat16, num_blocks=None,
    seed=0, unique_pages=True, kv_dtype=None, k_scale=1.0, v_scale=1.0, interleaved=False, device="cpu",
    max_value=1.0,
):
    """Synthetic inputs in the reference's layout (scripts/benchmark.py:1053-1227):
    q [T,Hq,D]; caches [NB,bs,Hkv,D]; int32 block_tables; int32 seq_lens; cu_seqlens_q.
    `interleaved=True` builds the vLLM>=0.11 allocation [NB,2,bs,Hkv,D] and returns
    K/V as unbind(1) views (block stride = 2*bs*Hkv*D)."""
    g = torch.Generator().manual_seed(seed)
    q_lens = [a for a, _ in q_kv_lens]
    kv_lens = [b for _, b in q_kv_lens]
    T = sum(q_lens)
    blocks_per_seq = [(kl + block_size - 1) // block_size for kl in kv_lens]
    need = sum(blocks_per_seq)
    if num_blocks is None:
        num_blocks = need + 7
    q = ((torch.rand(T, num_q_heads, head_size, generator=g) * 2 - 1) * max_value).to(dtype)
    shape = (num_blocks, 2, block_size, num_kv_heads, head_size)
    kv = ((torch.rand(shape, generator=g) * 2 - 1) * max_value)
    if kv_dtype == torch.float8_e4m3fn:
        kv[:, 0] /= k_scale
        kv[:, 1] /= v_scale
        kv = kv.clamp(-448, 448).to(torch.float8_e4m3fn)
    else:
        kv = kv.to(dtype)
    if not interleaved:
        k_cache = kv[:, 0].contiguous()
        v_cache = kv[:, 1].contiguous()
    else:
        k_cache, v_cache = kv.unbind(1)
    max_blocks = max(blocks_per_seq) if blocks_per_seq else 1
    bt = torch.zeros(len(q_kv_lens), max(max_blocks, 1), dtype=torch.int32)
    if unique_pages:
        perm = torch.randperm(num_blocks, generator=g)[:need].to(torch.int32)
        o = 0
        for i, n in enumerate(blocks_per_seq):
            bt[i, :n] = perm[o : o + n]
            o += n
    else:
        bt = torch.randint(0, num_blocks, bt.shape, generator=g, dtype=torch.int32)
    cu = torch.tensor([0] + list(np.cumsum(q_lens)), dtype=torch.int32)
    sl = torch.tensor(kv_lens, dtype=torch.int32)
    out = dict(q=q, k_cache=k_cache, v_cache=v_cache, kv=kv, block_table=bt, cu_seqlens_q=cu, seq_lens=sl,
               q_lens=q_lens, kv_lens=kv_lens)
    if device != "cpu":
        for k, v in list(out.items()):
            if isinstance(v, torch.Tensor):
                out[k] = v.to(device)
        if interleaved:
            out["k_cache"], out["v_cache"] = out["kv"].unbind(1)
    return out

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


def device_attention_reference(q, k_cache, v_cache, cu_seqlens_q, seq_lens, block_table, softmax_scale, k_scale=1.0,
                               v_scale=1.0, sliding_window=0, softcap=0.0, rows=None):
    """fp32 torch reference evaluated ON THE DEVICE the tensors live on: the same arithmetic, statement by statement,
    as oracle/paged_attention.py:unified_attention_ref (reference kernel_unified_attention_2d, triton_unified_attention.py
    :400-511: fp8 -> fp32 * scale -> query dtype, scale after the dot, softcap before the mask, causal/context + window
    mask, fp32 softmax, acc / L) - for BASELINE.json's full sizes, where the CPU oracle stays the arbiter on a sample
    of sequences (tests/test_gpu_full_size_numeric.py checks this function against it there).
    `rows`: iterable of sequence indices to evaluate (default all); other rows of the result are NaN."""
    T, Hq, D = q.shape
    Hkv, bs = k_cache.shape[2], k_cache.shape[1]
    g = Hq // Hkv
    cu = [int(x) for x in cu_seqlens_q.tolist()]
    sl = [int(x) for x in seq_lens.tolist()]
    out = torch.full((T, Hq, D), float("nan"), dtype=torch.float32, device=q.device)
    fp8 = k_cache.dtype == torch.float8_e4m3fn
    torch.backends.cuda.matmul.allow_tf32 = False  # a true fp32 reference
    torch.set_float32_matmul_precision("highest")

    def rows_of(cache, b, n, scale):
        idx = block_table[b, : (n + bs - 1) // bs].long()
        r = cache[idx].reshape(-1, Hkv, D)[:n]
        if fp8:
            return (r.to(torch.float32) * scale).to(q.dtype).to(torch.float32)
        return r.to(torch.float32)

    for b in (range(len(sl)) if rows is None else rows):
        q0, q1 = cu[b], cu[b + 1]
        ql, kl = q1 - q0, sl[b]
        if ql == 0:
            continue
        if kl == 0:
            out[q0:q1] = 0
            continue
        ctx = kl - ql
        k, v = rows_of(k_cache, b, kl, k_scale), rows_of(v_cache, b, kl, v_scale)  # [kv, Hkv, D]
        qb = q[q0:q1].to(torch.float32).view(ql, Hkv, g, D)
        s = softmax_scale * torch.einsum("qhgd,khd->hgqk", qb, k)
        if softcap > 0:
            s = softcap * torch.tanh(s / softcap)
        qpos = torch.arange(ql, device=q.device).view(1, 1, ql, 1)
        kpos = torch.arange(kl, device=q.device).view(1, 1, 1, kl)
        vis = kpos < ctx + qpos + 1
        if sliding_window > 0:
            vis = vis & ((ctx + qpos - kpos) < sliding_window)
        s = s.masked_fill(~vis, float("-inf"))
        m = s.max(dim=-1, keepdim=True).values
        m = torch.where(m > float("-inf"), m, torch.zeros_like(m))
        p = torch.exp(s - m)
        l = p.sum(dim=-1, keepdim=True)
        o = torch.einsum("hgqk,khd->qhgd", p, v) / l.permute(2, 0, 1, 3)
        out[q0:q1] = o.reshape(ql, Hq, D)
        del s, p, k, v
    return out

"""CPU oracle (test infrastructure, not product): paged-KV attention.

fp32 restatement of the reference attention path.  Each function cites the
reference lines it follows (paths relative to /root/reference):

* `unified_attention_ref`  — the arithmetic of `kernel_unified_attention_2d`
  (ibm-triton-lib/ibm_triton_lib/kernels/triton_unified_attention.py:275-523)
  evaluated densely per sequence instead of tile by tile (results are
  tile-size independent up to fp32 summation order).
* `paged_attn_torch_ref`   — `ref_paged_attn` (scripts/vllm_utils.py:433-488).
* `split_kv_partials_ref` / `reduce_segments_ref` — `kernel_unified_attention_3d`
  (:526-754) and `reduce_segments` (:757-836).

Parity status: pinned by tests/golden (see oracle/__init__.py).
"""
from __future__ import annotations

import math
from typing import Optional, Sequence

import torch

_FP8 = torch.float8_e4m3fn


def _dequant(cache_rows: torch.Tensor, scale: float, compute_dtype: torch.dtype) -> torch.Tensor:
    """K/V dequantisation as triton_unified_attention.py:434-455 does it:
    fp8 -> fp32, times the per-tensor scale, rounded to the query dtype."""
    if cache_rows.dtype == _FP8:
        return (cache_rows.to(torch.float32) * scale).to(compute_dtype).to(torch.float32)
    return cache_rows.to(torch.float32)


def gather_kv(cache: torch.Tensor, block_table_row: torch.Tensor, kv_len: int) -> torch.Tensor:
    """[num_blocks, block_size, Hkv, D] + one block-table row -> [kv_len, Hkv, D]
    (scripts/vllm_utils.py:456-462)."""
    block_size = cache.shape[1]
    n_blocks = (kv_len + block_size - 1) // block_size
    idx = block_table_row[:n_blocks].to(torch.long)
    rows = cache[idx].reshape(-1, cache.shape[2], cache.shape[3])
    return rows[:kv_len]


def _softcap(s: torch.Tensor, cap: float) -> torch.Tensor:
    # apply_softcap (:24-29) is cap*tanh(s/cap) written with exp(); tanh is the
    # same function without the overflow of the exp form.
    return cap * torch.tanh(s / cap)


def unified_attention_ref(
    q: torch.Tensor,  # [T, Hq, D]
    k_cache: torch.Tensor,  # [NB, bs, Hkv, D]
    v_cache: torch.Tensor,
    cu_seqlens_q: Sequence[int],  # [B+1]
    seq_lens: Sequence[int],  # [B]
    block_table: torch.Tensor,  # [B, MB] int
    softmax_scale: float,
    sliding_window: int = 0,  # 0 = off; kernel's SLIDING_WINDOW (= 1 + window_size[0])
    softcap: float = 0.0,
    alibi_slopes: Optional[torch.Tensor] = None,  # [Hq] fp32
    k_scale: float = 1.0,
    v_scale: float = 1.0,
    round_p_to: Optional[torch.dtype] = None,  # kernel casts P to V's dtype before PV (:508)
    out_dtype: Optional[torch.dtype] = None,
    sinks: Optional[torch.Tensor] = None,  # [Hq]: per-head logit that only joins the softmax denominator
    output_scale: Optional[float] = None,  # static per-tensor FP8 output quantisation; result is float8_e4m3fn
    local_chunk: int = 0,  # chunked local attention: a query only sees keys of its own chunk of positions
    qq_bias: Optional[torch.Tensor] = None,  # [q, n] fp32: bias between the query tokens of a step
    return_lse: bool = False,  # also return fp32 [T, Hq]: ln sum_j exp(s_ij) (+ exp(sink)); -inf where nothing is visible
    k_scale_cache: Optional[torch.Tensor] = None,  # [NB, bs, Hkv] fp32: per-(token, head) scales (fp8 or int8 cache)
    v_scale_cache: Optional[torch.Tensor] = None,
):
    """`k_scale_cache` / `v_scale_cache` (upstream KVQuantMode.FP8_PER_TOKEN_HEAD / INT8_PER_TOKEN_HEAD, vLLM 0.22
    vllm/v1/attention/ops/triton_unified_attention.py:39-55, 469-489, 505-508, 539-544): K and V are cast to the query
    dtype as stored, S = dot(Q, K) * (softmax_scale * k_scale[key]) and acc += (P * v_scale[key]).to(q dtype) @ V.
    `qq_bias` (upstream, vllm/v1/attention/ops/triton_unified_attention.py:526-528 with
    triton_attention_helpers.py:344-357): after masking and ALiBi, S[query i][key j] += qq_bias[i][j - ctx] for
    0 <= j - ctx < qq_bias.stride(0), the same matrix for every sequence and head.
    `local_chunk` states chunked local attention (iRoPE / Llama-4) directly as a mask: position p
    attends to keys in [floor(p / chunk) * chunk, p].  The reference gets the same result by cutting
    the batch into virtual sequences (`make_local_attention_virtual_batches`, used at
    ibm-triton-lib/ibm_triton_lib/backend/triton_attn.py:155-190, 423-444) and running the unchanged kernel.
    `output_scale` follows the upstream kernel epilogue (installed vLLM
    vllm/v1/attention/ops/triton_unified_attention.py:608-610, 722-724, 986): the fp32 result times
    the fp32 reciprocal of the scale, clamped to +-448, stored as e4m3 (round to nearest even).
    `sinks` is an upstream addition to the same kernel (vLLM's
    vllm/v1/attention/ops/triton_attention_helpers.py `init_softmax_M`: the running max starts at
    the head's sink value and the running sum at 1, i.e. one extra key with logit `sink`, unscaled,
    and a zero value row)."""
    T, Hq, D = q.shape
    Hkv = k_cache.shape[2]
    g = Hq // Hkv
    cu = [int(x) for x in cu_seqlens_q]
    sl = [int(x) for x in seq_lens]
    out = torch.zeros(T, Hq, D, dtype=torch.float32)
    lse_out = torch.full((T, Hq), float("-inf"), dtype=torch.float32)
    cdt = q.dtype
    for b in range(len(sl)):
        q0, q1 = cu[b], cu[b + 1]
        q_len, kv_len = q1 - q0, sl[b]
        if q_len == 0 or kv_len == 0:
            continue  # nothing visible: output stays 0 (reduce_segments :828)
        ctx = kv_len - q_len  # :374
        qb = q[q0:q1].to(torch.float32)  # [q, Hq, D]
        per_token_head = k_scale_cache is not None
        if per_token_head:
            k = gather_kv(k_cache, block_table[b], kv_len).to(torch.float32)  # plain cast (_cast_kv_tile)
            v = gather_kv(v_cache, block_table[b], kv_len).to(torch.float32)
            ks = gather_kv(k_scale_cache.unsqueeze(-1), block_table[b], kv_len)[..., 0].to(torch.float32)  # [kv, Hkv]
            vs = gather_kv(v_scale_cache.unsqueeze(-1), block_table[b], kv_len)[..., 0].to(torch.float32)
            ks = ks.repeat_interleave(g, dim=1).T.unsqueeze(1)  # [Hq, 1, kv]
            vs = vs.repeat_interleave(g, dim=1).T.unsqueeze(1)
        else:
            k = _dequant(gather_kv(k_cache, block_table[b], kv_len), k_scale, cdt)
            v = _dequant(gather_kv(v_cache, block_table[b], kv_len), v_scale, cdt)
        k = k.repeat_interleave(g, dim=1)  # [kv, Hq, D]
        v = v.repeat_interleave(g, dim=1)
        if per_token_head:
            s = torch.einsum("qhd,khd->hqk", qb, k) * (torch.tensor(softmax_scale, dtype=torch.float32) * ks)
        else:
            s = softmax_scale * torch.einsum("qhd,khd->hqk", qb, k)  # :465 (scale after the dot)
        if softcap > 0:
            s = _softcap(s, softcap)  # :467-468, before masking
        qpos = torch.arange(q_len).view(1, q_len, 1)
        kpos = torch.arange(kv_len).view(1, 1, kv_len)
        visible = kpos < ctx + qpos + 1  # :460
        if sliding_window > 0:
            visible = visible & ((ctx + qpos - kpos) < sliding_window)  # :474-479
        if local_chunk > 0:
            visible = visible & (kpos >= ((ctx + qpos) // local_chunk) * local_chunk)
        s = torch.where(visible, s, torch.full_like(s, float("-inf")))
        if alibi_slopes is not None:
            s = s + alibi_slopes.to(torch.float32).view(Hq, 1, 1) * (kpos - ctx)  # :481-482
        if qq_bias is not None:
            n_rel = min(qq_bias.stride(0), q_len, qq_bias.shape[1])
            bias = torch.zeros(q_len, kv_len, dtype=torch.float32)
            bias[:, ctx:ctx + n_rel] = qq_bias[:q_len, :n_rel].to(torch.float32)
            s = s + bias.view(1, q_len, kv_len)
        m = s.max(dim=-1, keepdim=True).values
        if sinks is not None:
            m = torch.maximum(m, sinks.to(torch.float32).view(Hq, 1, 1))
        m = torch.where(m > float("-inf"), m, torch.zeros_like(m))  # :489
        p = torch.exp(s - m)
        l = p.sum(dim=-1, keepdim=True)
        if sinks is not None:
            l = l + torch.exp(sinks.to(torch.float32).view(Hq, 1, 1) - m)
        if per_token_head:
            p = p * vs
        if round_p_to is not None:
            p = p.to(round_p_to).to(torch.float32)
        o = torch.einsum("hqk,khd->qhd", p, v) / l.permute(1, 0, 2)  # :508-511
        out[q0:q1] = o
        # what FlashAttention calls softmax_lse (the quantity decode context parallelism combines partial results
        # with, vllm/vllm/attention/backends/abstract.py:130-153): ln of the softmax denominator in natural units
        lse_out[q0:q1] = torch.where(l > 0, m + torch.log(l), torch.full_like(l, float("-inf"))).squeeze(-1).T
    if output_scale is not None:
        inv = 1.0 / torch.tensor(float(output_scale), dtype=torch.float32)  # fp32 reciprocal, as `1 / output_scale`
        res = (out * inv).clamp(-448.0, 448.0).to(torch.float8_e4m3fn)
    else:
        res = out.to(out_dtype or q.dtype)
    return (res, lse_out) if return_lse else res


def paged_attn_torch_ref(
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    query_lens: Sequence[int],
    kv_lens: Sequence[int],
    block_tables: torch.Tensor,
    scale: float,
    sliding_window: Optional[int] = None,
    soft_cap: Optional[float] = None,
) -> torch.Tensor:
    """`ref_paged_attn` (scripts/vllm_utils.py:433-488) without its in-place
    scaling of `query`: triu causal mask, optional sliding
    window, tanh softcap before masking, softmax in fp32 cast to V's dtype."""
    outs = []
    start = 0
    for i, (ql, kl) in enumerate(zip(query_lens, kv_lens)):
        q = query[start : start + ql] * scale
        k = gather_kv(key_cache, block_tables[i], kl)
        v = gather_kv(value_cache, block_tables[i], kl)
        if q.shape[1] != k.shape[1]:
            rep = q.shape[1] // k.shape[1]
            k = k.repeat_interleave(rep, dim=1)
            v = v.repeat_interleave(rep, dim=1)
        attn = torch.einsum("qhd,khd->hqk", q, k).float()
        ones = torch.ones(ql, kl)
        mask = torch.triu(ones, diagonal=kl - ql + 1).bool()
        if sliding_window is not None:
            mask |= torch.triu(ones, diagonal=kl - (ql + sliding_window) + 1).bool().logical_not()
        if soft_cap is not None and soft_cap > 0:
            attn = soft_cap * torch.tanh(attn / soft_cap)
        attn.masked_fill_(mask, float("-inf"))
        attn = torch.softmax(attn, dim=-1).to(v.dtype)
        outs.append(torch.einsum("hqk,khd->qhd", attn, v))
        start += ql
    return torch.cat(outs, dim=0)


def split_kv_partials_ref(
    q: torch.Tensor,  # [B, Hq, D]  (decode: one token per sequence)
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    seq_lens: Sequence[int],
    block_table: torch.Tensor,
    softmax_scale: float,
    num_segments: int = 16,
):
    """Per-segment un-normalised partials of `kernel_unified_attention_3d`
    (:526-754) for q_len == 1: returns (acc [B,Hq,S,D], M [B,Hq,S], L [B,Hq,S]).
    `L` starts at 1.0 and `M` at -inf (:624-625); the first tile's
    alpha = exp(-inf) = 0 removes the 1.0.  Empty segments keep (0, -inf, 0)
    here; the reference leaves them unwritten and masks them in the reduce."""
    B, Hq, D = q.shape
    Hkv, bs = k_cache.shape[2], k_cache.shape[1]
    g = Hq // Hkv
    acc = torch.zeros(B, Hq, num_segments, D)
    M = torch.full((B, Hq, num_segments), float("-inf"))
    L = torch.zeros(B, Hq, num_segments)
    for b, kv_len in enumerate(int(x) for x in seq_lens):
        blocks_per_segment = -(-kv_len // (num_segments * bs))  # :592
        k = gather_kv(k_cache, block_table[b], kv_len).float().repeat_interleave(g, dim=1)
        v = gather_kv(v_cache, block_table[b], kv_len).float().repeat_interleave(g, dim=1)
        for sgm in range(num_segments):
            lo = sgm * blocks_per_segment * bs
            if lo >= kv_len:  # :594
                continue
            hi = min((sgm + 1) * blocks_per_segment * bs, kv_len)
            s = softmax_scale * torch.einsum("hd,khd->hk", q[b].float(), k[lo:hi])
            m = s.max(dim=-1).values
            p = torch.exp(s - m[:, None])
            acc[b, :, sgm] = torch.einsum("hk,khd->hd", p, v[lo:hi])
            M[b, :, sgm] = m
            L[b, :, sgm] = p.sum(-1)
    return acc, M, L


def reduce_segments_ref(acc: torch.Tensor, M: torch.Tensor, L: torch.Tensor) -> torch.Tensor:
    """`reduce_segments` (:757-836): rescale by exp(M_s - max_s M_s), sum, divide;
    0 where the total exp-sum is 0 (:828)."""
    overall = M.max(dim=-1, keepdim=True).values
    w = torch.exp(M - overall)
    w = torch.where(torch.isnan(w), torch.zeros_like(w), w)
    expsum = (L * w).sum(-1)
    num = (acc * w[..., None]).sum(-2)
    return torch.where(expsum[..., None] == 0, torch.zeros_like(num), num / expsum[..., None])


def attention_flops(query_lens: Sequence[int], seq_lens: Sequence[int], num_q_heads: int, head_size: int) -> int:
    """Algorithmic FLOPs of the causal path, SURVEY.md §8(d):
    sum_b 4*Hq*D*(q_b*ctx_b + q_b(q_b+1)/2)."""
    total = 0
    for ql, sl in zip(query_lens, seq_lens):
        ctx = sl - ql
        total += 4 * num_q_heads * head_size * (ql * ctx + ql * (ql + 1) // 2)
    return total


def decode_bytes(seq_lens: Sequence[int], num_q_heads: int, num_kv_heads: int, head_size: int,
                 block_size: int, kv_elem_bytes: int, q_elem_bytes: int = 2, num_tokens: Optional[int] = None) -> int:
    """Algorithmic bytes of one decode call, SURVEY.md §8(d)."""
    B = len(seq_lens)
    T = B if num_tokens is None else num_tokens
    kv = sum(int(s) for s in seq_lens) * num_kv_heads * head_size * 2 * kv_elem_bytes
    qo = T * num_q_heads * head_size * (q_elem_bytes + 2)
    bt = sum(math.ceil(int(s) / block_size) for s in seq_lens) * 4
    return kv + qo + bt + (2 * B + 1) * 4

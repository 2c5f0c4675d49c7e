"""Seeded random sweep of the whole argument space of `unified_attention` against the CPU oracle.

The hand-picked cases elsewhere pin each feature; this walks the combinations nobody wrote down: head size x
GQA group x block size x dtype x FP8 cache x sliding window x softcap x ALiBi x sinks x FP8 output, on batches
that mix empty, single-token, short-extend and prompt sequences, through the library's own dispatch (so every
kernel and every routing rule is hit).  Tolerances are BASELINE.json's (bf16 2e-2 / 1e-2, FP8 KV 6e-2)."""
import random

import pytest
import torch

from helpers import BF16_ATOL, BF16_RTOL, FP8_ATOL, assert_close, make_paged_inputs
from oracle.paged_attention import unified_attention_ref

pytestmark = pytest.mark.gpu

N_CASES = 256


def _case(seed):
    r = random.Random(1000 + seed)
    D = r.choice([64, 128, 128, 128, 256])
    Hkv = r.choice([1, 2, 4, 8])
    g = r.choice([1, 2, 3, 4, 4, 5, 6, 8, 8] + ([12, 16] if D <= 128 else []))
    bs = r.choice([16, 16, 32, 64, 128])
    fp8 = r.random() < 0.35
    dtype = r.choice([torch.bfloat16, torch.float16])
    kind = r.choice(["decode", "decode", "mixed", "mixed", "prefill"])
    n = r.randint(1, 9)
    lens = []
    for _ in range(n):
        if kind == "decode":
            lens.append((1, r.choice([1, 2, bs, bs + 1, r.randint(1, 2500)])))
        elif kind == "prefill":
            q = r.randint(1, 400)
            lens.append((q, q + r.choice([0, 0, r.randint(1, 900)])))
        else:
            q = r.choice([1, 1, r.randint(2, 20), r.randint(21, 300)])
            lens.append((q, q + r.randint(0, 1200)))
    if r.random() < 0.15:  # a sequence with nothing cached and nothing visible
        lens[r.randrange(len(lens))] = (1, 0)
    return dict(
        D=D, Hkv=Hkv, Hq=Hkv * g, bs=bs, fp8=fp8, dtype=dtype, lens=lens,
        window=r.choice([0, 0, 0, r.randint(1, 300)]), softcap=r.choice([0.0, 0.0, 0.0, r.uniform(5.0, 50.0)]),
        alibi=r.random() < 0.2, sinks=r.random() < 0.25, out_fp8=r.random() < 0.2,
        scales=(r.choice([1.0, 0.5, 0.37]), r.choice([1.0, 0.5, 0.61])) if fp8 else (1.0, 1.0),
        interleaved=r.random() < 0.5, qmax=r.choice([1.0, 1.0, 3.0]),
    )


@pytest.mark.parametrize("seed", range(N_CASES))
def test_random_configuration_matches_oracle(seed):
    from vllm_triton_backend_b200 import _cabi, ops

    c = _case(seed)
    dev = "cuda"
    Hq, Hkv, D, bs = c["Hq"], c["Hkv"], c["D"], c["bs"]
    ks, vs = c["scales"]
    lens = c["lens"]
    # make_paged_inputs needs at least one page per sequence; an empty sequence keeps its (unused) page
    alloc = [(a, max(b, 1)) for a, b in lens]
    inp = make_paged_inputs(alloc, Hq, Hkv, D, bs, dtype=c["dtype"], seed=seed, max_value=c["qmax"],
                            kv_dtype=torch.float8_e4m3fn if c["fp8"] else None, k_scale=ks, v_scale=vs,
                            interleaved=c["interleaved"])
    inp["seq_lens"] = torch.tensor([b for _, b in lens], dtype=torch.int32)
    scale = D ** -0.5
    slopes = torch.tensor([2.0 ** (-(i + 1) * 8.0 / Hq) for i in range(Hq)], dtype=torch.float32) if c["alibi"] else None
    sinks = torch.linspace(-2.0, 4.0, Hq, dtype=torch.float32) if c["sinks"] else None
    out_scale = 0.02 if c["out_fp8"] else None
    ref = unified_attention_ref(
        inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"], inp["block_table"], scale,
        sliding_window=c["window"], softcap=c["softcap"], alibi_slopes=slopes, k_scale=ks, v_scale=vs, sinks=sinks,
        out_dtype=torch.float32)
    q = inp["q"].to(dev)
    kv = inp["kv"].to(dev)
    kc, vc = (kv.unbind(1) if c["interleaved"] else (kv[:, 0].contiguous(), kv[:, 1].contiguous()))
    if out_scale is None:
        out = torch.full_like(q, float("nan"))
    else:
        out = torch.full(q.shape, 0x7F, dtype=torch.uint8, device=dev).view(torch.float8_e4m3fn)
    ops.unified_attention(
        q=q, k=kc, v=vc, out=out, cu_seqlens_q=inp["cu_seqlens_q"].to(dev), max_seqlen_q=max(a for a, _ in lens),
        seqused_k=inp["seq_lens"].to(dev), max_seqlen_k=max(max(b for _, b in lens), 1), softmax_scale=scale, causal=True,
        window_size=(c["window"] - 1, 0) if c["window"] > 0 else (-1, -1), block_table=inp["block_table"].to(dev),
        softcap=c["softcap"], q_descale=None, k_descale=torch.tensor([ks], dtype=torch.float32, device=dev),
        v_descale=torch.tensor([vs], dtype=torch.float32, device=dev), alibi_slopes=None if slopes is None else slopes.to(dev),
        sinks=None if sinks is None else sinks.to(dev),
        output_scale=None if out_scale is None else torch.tensor([out_scale], dtype=torch.float32, device=dev))
    torch.cuda.synchronize()
    what = f"seed {seed} ({_cabi.last_kernel()}): " + ", ".join(f"{k}={v}" for k, v in c.items() if k != "lens") + f", lens={lens}"
    atol = FP8_ATOL if c["fp8"] else BF16_ATOL
    if out_scale is None:
        assert_close(out.cpu(), ref, atol=atol, rtol=BF16_RTOL, what=what)
    else:  # dequantised e4m3: one code step (2^-3 relative) on top of the unquantised tolerance
        got = out.cpu().float() * out_scale
        assert not torch.isnan(got).any(), what
        assert_close(got, ref.clamp(-448 * out_scale, 448 * out_scale), atol=atol + out_scale * 2.0 ** -6, rtol=BF16_RTOL + 0.125, what=what)

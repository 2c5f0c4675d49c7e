"""Numeric parity at BASELINE.json's FULL sizes (the property tests of test_gpu_full_size_properties.py check the
index path; these check the softmax arithmetic - long split-KV merges, 8k-token lazy rescale - at the sizes the
metric is quoted on):

  C2  exactly: 64 sequences x ctx 4096, bf16, 32q/8kv - every row against the CPU oracle
  C4  64 sequences x ctx 32768, FP8-e4m3 KV (scales 0.5 and 0.37), all 8 KV heads and the 1-KV-head shard one of 8
      GPUs holds - every row against the fp32 torch reference evaluated on the device (same statements as the
      oracle; helpers.device_attention_reference), the CPU oracle as arbiter on sampled sequences
  C3  SURVEY.md 8(d)'s chunked-prefill batch (32 decode rows + a fresh 2048-token prompt + two chunked continuations
      + 4 prefix-cache hits), bf16 and FP8 cache - every row against the device reference, sampled sequences
      against the CPU oracle

Tolerances are BASELINE.json's: bf16 max-abs 2e-2 + rtol 1e-2, FP8 KV 6e-2.  Reference arithmetic:
/root/reference/scripts/vllm_utils.py:433-488, /root/reference/scripts/benchmark.py:987-1347 (the batch)."""
import random

import pytest
import torch

from helpers import BF16_ATOL, BF16_RTOL, FP8_ATOL, assert_close, device_attention_reference
from oracle.paged_attention import unified_attention_ref

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _fill(t, g, fp8_scale=None):
    flat = t.view(-1)
    chunk = 1 << 26
    for o in range(0, flat.numel(), chunk):
        n = min(chunk, flat.numel() - o)
        x = torch.rand(n, device=t.device, generator=g) * 2 - 1
        if fp8_scale is not None:
            flat[o:o + n] = (x / fp8_scale).to(torch.float8_e4m3fn).view(torch.uint8)
        else:
            flat[o:o + n] = x.to(t.dtype)


def _paged(lens, Hq, Hkv, D, bs, fp8, k_scale, v_scale, seed):
    """Caches resident on the device (too big to build on the host at C4 sizes), unique random pages."""
    g = torch.Generator(device=DEV).manual_seed(seed)
    pages = [(kl + bs - 1) // bs for _, kl in lens]
    NB = sum(pages) + 64
    if fp8:
        kc = torch.empty(NB, bs, Hkv, D, dtype=torch.uint8, device=DEV)
        vc = torch.empty_like(kc)
        _fill(kc, g, k_scale), _fill(vc, g, v_scale)
        kc, vc = kc.view(torch.float8_e4m3fn), vc.view(torch.float8_e4m3fn)
    else:
        kc = torch.empty(NB, bs, Hkv, D, dtype=torch.bfloat16, device=DEV)
        vc = torch.empty_like(kc)
        _fill(kc, g), _fill(vc, g)
    perm = torch.randperm(NB, device=DEV, generator=g).to(torch.int32)
    bt = torch.zeros(len(lens), max(pages), dtype=torch.int32, device=DEV)
    o = 0
    for i, n in enumerate(pages):
        bt[i, :n] = perm[o:o + n]
        o += n
    T = sum(ql for ql, _ in lens)
    q = (torch.rand(T, Hq, D, device=DEV, generator=g) * 2 - 1).to(torch.bfloat16)
    cu = torch.tensor([0] + list(torch.tensor([ql for ql, _ in lens]).cumsum(0)), dtype=torch.int32, device=DEV)
    sl = torch.tensor([kl for _, kl in lens], dtype=torch.int32, device=DEV)
    return dict(q=q, kc=kc, vc=vc, bt=bt, cu=cu, sl=sl, lens=lens)


def _ours(p, k_scale=None, v_scale=None):
    from vllm_triton_backend_b200 import _cabi, ops

    out = torch.full_like(p["q"], float("nan"))
    ks = None if k_scale is None else torch.tensor([k_scale], dtype=torch.float32, device=DEV)
    vs = None if v_scale is None else torch.tensor([v_scale], dtype=torch.float32, device=DEV)
    D = p["q"].shape[-1]
    ops.unified_attention(q=p["q"], k=p["kc"], v=p["vc"], out=out, cu_seqlens_q=p["cu"],
                          max_seqlen_q=max(a for a, _ in p["lens"]), seqused_k=p["sl"], max_seqlen_k=max(b for _, b in p["lens"]),
                          softmax_scale=D ** -0.5, causal=True, window_size=(-1, -1), block_table=p["bt"], softcap=0,
                          q_descale=None, k_descale=ks, v_descale=vs)
    torch.cuda.synchronize()
    return out, _cabi.last_kernel()


def _oracle_rows(p, seqs, k_scale=1.0, v_scale=1.0):
    """CPU oracle on a sample of sequences: their pages are gathered into a small cache of their own."""
    bs = p["kc"].shape[1]
    D = p["q"].shape[-1]
    res = {}
    cu = p["cu"].tolist()
    for b in seqs:
        ql, kl = p["lens"][b]
        n = (kl + bs - 1) // bs
        idx = p["bt"][b, :n].long()
        kc, vc = p["kc"][idx].cpu(), p["vc"][idx].cpu()
        bt = torch.arange(n, dtype=torch.int32).view(1, n)
        res[b] = unified_attention_ref(p["q"][cu[b]:cu[b + 1]].cpu(), kc, vc, [0, ql], [kl], bt, D ** -0.5, k_scale=k_scale,
                                       v_scale=v_scale, out_dtype=torch.float32)
    return res


def _check(p, out, atol, k_scale=1.0, v_scale=1.0, oracle_seqs=(), what=""):
    ref = device_attention_reference(p["q"], p["kc"], p["vc"], p["cu"], p["sl"], p["bt"], p["q"].shape[-1] ** -0.5,
                                     k_scale=k_scale, v_scale=v_scale)
    assert not torch.isnan(out).any() and not torch.isnan(ref).any()
    assert_close(out, ref, atol=atol, rtol=BF16_RTOL, what=f"{what}: every row vs the device fp32 reference")
    cu = p["cu"].tolist()
    for b, exp in _oracle_rows(p, oracle_seqs, k_scale, v_scale).items():
        rows = slice(cu[b], cu[b + 1])
        # the device reference restates the oracle: fp32 summation order is all that may differ
        assert_close(ref[rows], exp, atol=2e-4, rtol=1e-4, what=f"{what}: device reference vs CPU oracle, sequence {b}")
        assert_close(out[rows], exp, atol=atol, rtol=BF16_RTOL, what=f"{what}: sequence {b} vs the CPU oracle")


def test_c2_every_row_against_the_cpu_oracle():
    lens = [(1, 4096)] * 64
    p = _paged(lens, 32, 8, 128, 16, False, 1.0, 1.0, seed=2)
    out, kern = _ours(p)
    assert kern.startswith("attn_decode_split_kv")
    _check(p, out, BF16_ATOL, oracle_seqs=range(64), what="C2")


@pytest.mark.parametrize("Hq,Hkv", [(64, 8), (8, 1)], ids=["8_kv_heads", "1_kv_head_shard"])
@pytest.mark.parametrize("scale", [0.5, 0.37])
def test_c4_fp8_decode_at_32k(Hq, Hkv, scale):
    lens = [(1, 32768)] * 64
    p = _paged(lens, Hq, Hkv, 128, 16, True, scale, scale, seed=4)
    out, kern = _ours(p, scale, scale)
    assert kern.startswith("attn_decode_split_kv")
    _check(p, out, FP8_ATOL, scale, scale, oracle_seqs=(0, 37) if Hkv == 8 else (0, 21, 42, 63), what=f"C4 {Hq}q/{Hkv}kv scale {scale}")


def _c3_batch(seed=0):
    rnd = random.Random(seed)
    dec = [(1, rnd.randint(128, 8192)) for _ in range(32)]
    big = [(2048, 2048), (2048, 6144), (1024, 7168)]  # fresh prompt, two chunked continuations (ctx 4096, 6144)
    hits = [(rnd.randint(128, 1024), 0) for _ in range(4)]
    hits = [(q, q + (1024 if i % 2 == 0 else 2048)) for i, (q, _) in enumerate(hits)]  # prefix-cache hits
    seqs, rest = [], big + hits
    while dec or rest:  # alternate, as the reference harness does (scripts/benchmark.py:1101-1112)
        if rest:
            seqs.append(rest.pop(0))
        if dec:
            seqs.append(dec.pop(0))
    return seqs


@pytest.mark.parametrize("fp8", [False, True], ids=["bf16", "fp8_kv"])
def test_c3_chunked_prefill_mixed_batch(fp8):
    from vllm_triton_backend_b200 import _cabi

    lens = _c3_batch()
    sc = 0.37 if fp8 else 1.0
    p = _paged(lens, 32, 8, 128, 16, fp8, sc, sc, seed=6)
    _cabi.reset_launch_count()
    out, _ = _ours(p, sc if fp8 else None, sc if fp8 else None)
    assert _cabi.launch_count() >= 3, "prompts, short extends and decode rows take different kernels"
    # sampled for the CPU oracle: the fresh 2048-token prompt, a prefix hit, three decode rows
    hit = next(i for i, (q, kv) in enumerate(lens) if 1 < q < 2048 and kv - q in (1024, 2048))
    dec = [i for i, (q, _) in enumerate(lens) if q == 1][:3]
    _check(p, out, FP8_ATOL if fp8 else BF16_ATOL, sc, sc, oracle_seqs=[0, hit] + dec, what="C3 " + ("fp8" if fp8 else "bf16"))

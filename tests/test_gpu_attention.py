"""K2/K3 parity on the GPU against the CPU oracle and the reference golden
vectors (tolerances from BASELINE.json: bf16 max-abs 2e-2 / rtol 1e-2, FP8 KV 6e-2).
All calls go through the C ABI (ops.unified_attention -> libb200attn.so)."""
import pytest
import torch

from helpers import (BF16_ATOL, BF16_RTOL, FP8_ATOL, assert_close, golden_attn_files, load_golden_attn,
                     make_paged_inputs)
from oracle.paged_attention import unified_attention_ref

pytestmark = pytest.mark.gpu

GENERIC, PREFILL, DECODE_OR_AUTO, PREFILL_TC = 1, 2, None, 4


def _run(inp, scale, path, window=0, softcap=0.0, alibi=None, k_scale=None, v_scale=None, dtype=None, sinks=None):
    from vllm_triton_backend_b200 import _cabi, ops

    dev = "cuda"
    q = inp["q"].to(dev) if dtype is None else inp["q"].to(dev, dtype)
    kc, vc = inp["k_cache"], inp["v_cache"]
    if dtype is not None and kc.dtype != torch.float8_e4m3fn:
        kc, vc = kc.to(dtype), vc.to(dtype)
    kc, vc = kc.to(dev), vc.to(dev)
    out = torch.full_like(q, float("nan"))
    q_lens = (inp["cu_seqlens_q"][1:] - inp["cu_seqlens_q"][:-1])
    max_q = int(q_lens.max())
    force = path
    if path == PREFILL and max_q > 1:
        force = 22
    ops.unified_attention(
        q=q, k=kc, v=vc, out=out, cu_seqlens_q=inp["cu_seqlens_q"].to(dev), max_seqlen_q=max_q,
        seqused_k=inp["seq_lens"].to(dev), max_seqlen_k=int(inp["seq_lens"].max()), softmax_scale=scale, causal=True,
        window_size=(window - 1, 0) if window > 0 else (-1, -1), block_table=inp["block_table"].to(dev),
        softcap=softcap, q_descale=None,
        k_descale=None if k_scale is None else torch.tensor([k_scale], dtype=torch.float32, device=dev),
        v_descale=None if v_scale is None else torch.tensor([v_scale], dtype=torch.float32, device=dev),
        alibi_slopes=None if alibi is None else alibi.to(dev), force_selection=force,
        sinks=None if sinks is None else sinks.to(dev),
    )
    torch.cuda.synchronize()
    return out.cpu(), _cabi.last_kernel()


def _oracle(inp, scale, window=0, softcap=0.0, alibi=None, k_scale=1.0, v_scale=1.0, dtype=None, sinks=None):
    q = inp["q"] if dtype is None else inp["q"].to(dtype)
    kc, vc = inp["k_cache"], inp["v_cache"]
    if dtype is not None and kc.dtype != torch.float8_e4m3fn:
        kc, vc = kc.to(dtype), vc.to(dtype)
    return unified_attention_ref(
        q, kc, vc, inp["cu_seqlens_q"], inp["seq_lens"], inp["block_table"], scale, sliding_window=window,
        softcap=softcap, alibi_slopes=alibi, k_scale=k_scale, v_scale=v_scale, out_dtype=torch.float32, sinks=sinks,
    )


# ---------------------------------------------------------------------------------
# golden vectors produced by the reference Triton kernels
# ---------------------------------------------------------------------------------
@pytest.mark.parametrize("path", golden_attn_files(), ids=lambda p: p.split("attn_")[-1][:-4])
def test_golden_native_dtype_generic_kernel(path):
    """fp32/fp16 inputs as the reference ran them; SIMT kernel; tight tolerance."""
    c = load_golden_attn(path)
    out, kern = _run(c, c["scale"], GENERIC, window=c["window"], softcap=c["softcap"], alibi=c["alibi_slopes"])
    assert kern == "attn_generic"
    tight = c["dtype"] == torch.float32
    assert_close(out, c["out_triton"], atol=1e-4 if tight else 2e-3, rtol=1e-4 if tight else 2e-3, what=c["name"])


@pytest.mark.parametrize("path", golden_attn_files(), ids=lambda p: p.split("attn_")[-1][:-4])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_golden_tensor_core_kernels(path, dtype):
    """Same inputs cast to bf16/fp16 through the kernel the dispatcher picks."""
    c = load_golden_attn(path)
    out, kern = _run(c, c["scale"], DECODE_OR_AUTO, window=c["window"], softcap=c["softcap"], alibi=c["alibi_slopes"],
                     dtype=dtype)
    assert kern.startswith(("attn_decode_split_kv", "attn_prefill_mma", "attn_prefill_tc")), kern
    # golden inputs live on the fp16 grid; bf16 casts them, so compare with the oracle on the cast inputs too
    assert_close(out, _oracle(c, c["scale"], c["window"], c["softcap"], c["alibi_slopes"], dtype=dtype),
                 atol=BF16_ATOL, rtol=BF16_RTOL, what=c["name"] + " vs oracle")
    if dtype == torch.float16:
        assert_close(out, c["out_triton"], atol=BF16_ATOL, rtol=BF16_RTOL, what=c["name"] + " vs reference kernels")


# ---------------------------------------------------------------------------------
# decode (3D path)
# ---------------------------------------------------------------------------------
DECODE_SHAPES = [
    # (kv lens, Hq, Hkv, D, block_size)
    ([512, 512, 512, 512], 32, 8, 128, 16),          # C1
    ([1328, 18, 463, 1, 16, 17, 2011], 8, 2, 128, 16),
    ([523, 37, 2011], 4, 4, 128, 16),                # MHA (group 1)
    ([300, 4096, 129], 64, 8, 128, 16),              # 70B-shaped, group 8
    ([700, 33], 16, 1, 128, 32),                     # MQA-16 -> two n-tiles
    ([257, 1000], 8, 2, 64, 16),
    ([257, 1000], 8, 2, 256, 16),
    ([90, 3000, 5], 32, 8, 128, 128),
]


@pytest.mark.parametrize("lens,Hq,Hkv,D,bs", DECODE_SHAPES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("interleaved", [False, True])
def test_decode_matches_oracle(lens, Hq, Hkv, D, bs, dtype, interleaved):
    inp = make_paged_inputs([(1, n) for n in lens], Hq, Hkv, D, bs, dtype=dtype, seed=len(lens) + D, interleaved=interleaved)
    scale = D ** -0.5
    out, kern = _run(inp, scale, DECODE_OR_AUTO)
    assert kern.startswith("attn_decode_split_kv"), kern
    assert_close(out, _oracle(inp, scale), atol=BF16_ATOL, rtol=BF16_RTOL, what="decode")
    assert not torch.isnan(out).any()


@pytest.mark.parametrize("hpc", [1, 2, 4, 8])
@pytest.mark.parametrize("tma_split", [0, 1])
def test_decode_split_and_head_grouping_invariance(hpc, tma_split):
    """Result must not depend on how KV is cut across CTAs / warps / TMA boxes."""
    from vllm_triton_backend_b200 import _cabi

    lens = [4096, 100, 1777, 16, 3000, 15, 2048, 640]
    inp = make_paged_inputs([(1, n) for n in lens], 32, 8, 128, 16, dtype=torch.bfloat16, seed=3)
    scale = 128 ** -0.5
    try:
        _cabi.set_tuning(hpc, 0, 0, tma_split)
        out, _ = _run(inp, scale, DECODE_OR_AUTO)
    finally:
        _cabi.set_tuning(0, 0, 0, 0)
    assert_close(out, _oracle(inp, scale), atol=BF16_ATOL, rtol=BF16_RTOL, what=f"hpc={hpc}")


@pytest.mark.parametrize("feature", ["window", "softcap", "alibi", "all"])
@pytest.mark.parametrize("D", [64, 128])
def test_decode_features(feature, D):
    lens = [1328, 18, 463, 200, 64, 65]
    Hq, Hkv = 8, 2
    inp = make_paged_inputs([(1, n) for n in lens], Hq, Hkv, D, 16, dtype=torch.bfloat16, seed=11)
    window = 128 if feature in ("window", "all") else 0
    softcap = 50.0 if feature in ("softcap", "all") else 0.0
    alibi = torch.tensor([2.0 ** (-(i + 1)) for i in range(Hq)]) if feature in ("alibi", "all") else None
    scale = D ** -0.5
    out, kern = _run(inp, scale, DECODE_OR_AUTO, window=window, softcap=softcap, alibi=alibi)
    assert kern.startswith("attn_decode_split_kv")
    assert_close(out, _oracle(inp, scale, window, softcap, alibi), atol=BF16_ATOL, rtol=BF16_RTOL, what=feature)


@pytest.mark.parametrize("scales", [(0.5, 0.5), (0.37, 0.21), (1.0, 1.0)])
@pytest.mark.parametrize("Hq,Hkv", [(64, 8), (8, 1), (32, 8)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_decode_fp8_kv(scales, Hq, Hkv, dtype):
    lens = [2049, 333, 16, 1]
    inp = make_paged_inputs([(1, n) for n in lens], Hq, Hkv, 128, 16, dtype=dtype, seed=5,
                            kv_dtype=torch.float8_e4m3fn, k_scale=scales[0], v_scale=scales[1])
    scale = 128 ** -0.5
    out, kern = _run(inp, scale, DECODE_OR_AUTO, k_scale=scales[0], v_scale=scales[1])
    assert kern.startswith("attn_decode_split_kv")
    assert_close(out, _oracle(inp, scale, k_scale=scales[0], v_scale=scales[1]), atol=FP8_ATOL, rtol=BF16_RTOL, what="fp8 decode")


def test_decode_ignores_garbage_beyond_seq_len():
    """Slots past seq_len in the last page (and unused pages) may hold NaN/Inf."""
    lens = [17, 100, 1]
    inp = make_paged_inputs([(1, n) for n in lens], 8, 2, 128, 16, dtype=torch.bfloat16, seed=9)
    ref = _oracle(inp, 0.1)
    bt = inp["block_table"]
    for b, n in enumerate(lens):
        page = int(bt[b, (n - 1) // 16])
        inp["k_cache"][page, n % 16 if n % 16 else 16:] = float("nan")
        inp["v_cache"][page, n % 16 if n % 16 else 16:] = float("inf")
    out, _ = _run(inp, 0.1, DECODE_OR_AUTO)
    assert_close(out, ref, atol=BF16_ATOL, rtol=BF16_RTOL, what="garbage tail")


def test_decode_empty_and_single_token_sequences():
    inp = make_paged_inputs([(1, 1), (1, 0), (1, 2), (1, 16)], 8, 2, 128, 16, dtype=torch.bfloat16, seed=2)
    out, _ = _run(inp, 0.3, DECODE_OR_AUTO)
    ref = _oracle(inp, 0.3)
    ref[1] = 0  # no visible key: reduce_segments returns 0 (reference :828)
    assert_close(out, ref, atol=BF16_ATOL, rtol=BF16_RTOL, what="tiny")


# ---------------------------------------------------------------------------------
# prefill / prefix / chunked / mixed (2D path)
# ---------------------------------------------------------------------------------
MIXED = [
    # (q, kv) lens as in vllm/tests/kernels/attention/test_triton_unified_attention.py:85
    ([(1, 1328), (5, 18), (129, 463)], 4, 4, 128, 16),
    ([(1, 1328), (5, 18), (129, 463)], 8, 2, 128, 16),
    ([(1, 523), (1, 37), (1, 2011)], 8, 2, 256, 16),
    ([(1, 1328), (5, 18), (129, 463)], 8, 2, 256, 16),
    ([(200, 200), (64, 64), (1, 1), (77, 300)], 32, 8, 128, 16),   # fresh prompts + prefix hit
    ([(512, 1536), (33, 33), (1, 700), (1, 9)], 32, 8, 128, 32),   # chunked continuation + decodes
    ([(100, 100), (3, 50)], 40, 8, 64, 16),                        # group 5 (not a power of two)
    ([(70, 70), (129, 129)], 16, 1, 128, 16),                      # MQA 16
]


@pytest.mark.parametrize("lens,Hq,Hkv,D,bs", MIXED)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_prefill_mixed_matches_oracle(lens, Hq, Hkv, D, bs, dtype):
    inp = make_paged_inputs(lens, Hq, Hkv, D, bs, dtype=dtype, seed=Hq + D, interleaved=True)
    scale = D ** -0.5
    out, kern = _run(inp, scale, PREFILL)
    assert kern == "attn_prefill_mma", kern
    assert_close(out, _oracle(inp, scale), atol=BF16_ATOL, rtol=BF16_RTOL, what="prefill")


@pytest.mark.parametrize("feature", ["window", "softcap", "alibi", "all"])
def test_prefill_features(feature):
    lens = [(1, 1328), (5, 18), (129, 463), (300, 300)]
    Hq, Hkv, D = 8, 2, 128
    inp = make_paged_inputs(lens, Hq, Hkv, D, 16, dtype=torch.bfloat16, seed=21)
    window = 64 if feature in ("window", "all") else 0
    softcap = 50.0 if feature in ("softcap", "all") else 0.0
    alibi = torch.tensor([2.0 ** (-(i + 1)) for i in range(Hq)]) if feature in ("alibi", "all") else None
    scale = D ** -0.5
    out, _ = _run(inp, scale, PREFILL, window=window, softcap=softcap, alibi=alibi)
    assert_close(out, _oracle(inp, scale, window, softcap, alibi), atol=BF16_ATOL, rtol=BF16_RTOL, what=feature)


@pytest.mark.parametrize("scales", [(0.5, 0.5), (0.37, 0.21)])
def test_prefill_fp8_kv(scales):
    lens = [(1, 523), (64, 200), (130, 130)]
    inp = make_paged_inputs(lens, 16, 2, 128, 16, dtype=torch.bfloat16, seed=6, kv_dtype=torch.float8_e4m3fn,
                            k_scale=scales[0], v_scale=scales[1])
    scale = 128 ** -0.5
    out, _ = _run(inp, scale, PREFILL, k_scale=scales[0], v_scale=scales[1])
    assert_close(out, _oracle(inp, scale, k_scale=scales[0], v_scale=scales[1]), atol=FP8_ATOL, rtol=BF16_RTOL, what="fp8 prefill")


def test_decode_batch_through_prefill_kernel_matches_decode_kernel():
    """force_selection=2 sends a decode-only batch down the 2D kernel (reference :884)."""
    lens = [1328, 18, 463, 2011]
    inp = make_paged_inputs([(1, n) for n in lens], 32, 8, 128, 16, dtype=torch.bfloat16, seed=8)
    a, ka = _run(inp, 0.088, DECODE_OR_AUTO)
    b, kb = _run(inp, 0.088, PREFILL)
    assert ka.startswith("attn_decode_split_kv") and kb == "attn_prefill_mma"
    assert_close(a, b, atol=1e-2, rtol=1e-2, what="2D vs 3D")


# ---------------------------------------------------------------------------------
# tcgen05 / TMEM prefill kernel (D=128, bf16/f16 KV)
# ---------------------------------------------------------------------------------
TC_CASES = [
    # (q, kv) lens, Hq, Hkv, block_size
    ([(128, 128)], 4, 1, 16),                                      # one full tile pair, group 4
    ([(64, 64)], 4, 1, 16),                                        # second subtile empty
    ([(1, 1328), (5, 18), (129, 463)], 4, 4, 16),                  # upstream shapes, MHA
    ([(1, 1328), (5, 18), (129, 463)], 8, 2, 16),
    ([(200, 200), (64, 64), (1, 1), (77, 300)], 32, 8, 16),        # fresh prompts + prefix hit
    ([(512, 1536), (33, 33), (1, 700), (1, 9)], 32, 8, 32),        # chunked continuation + decodes
    ([(300, 1000), (257, 257)], 16, 2, 64),                        # group 8, block 64
    ([(130, 900)], 8, 8, 128),                                     # block == tile
    ([(100, 700), (31, 31)], 16, 1, 256),                          # MQA 16, block > tile
    ([(1000, 1000)], 64, 8, 16),
]


@pytest.mark.parametrize("lens,Hq,Hkv,bs", TC_CASES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("interleaved", [False, True])
@pytest.mark.parametrize("D", [128, 64, 256])
def test_prefill_tc_matches_oracle(lens, Hq, Hkv, bs, dtype, interleaved, D):
    """Every sequence forced through the tcgen05 kernels (force_selection=4), head sizes 128, 64 and 256."""
    inp = make_paged_inputs(lens, Hq, Hkv, D, bs, dtype=dtype, seed=Hq + bs, interleaved=interleaved)
    scale = D ** -0.5
    out, kern = _run(inp, scale, PREFILL_TC)
    assert kern == ("attn_prefill_tc256" if D == 256 else "attn_prefill_tc"), kern
    assert_close(out, _oracle(inp, scale), atol=BF16_ATOL, rtol=BF16_RTOL, what="prefill_tc")
    assert not torch.isnan(out).any()


@pytest.mark.parametrize("Hq,Hkv", [(40, 8), (6, 2), (12, 4), (7, 1), (48, 4), (36, 3), (10, 1)])
@pytest.mark.parametrize("D", [128, 64, 256])
@pytest.mark.parametrize("forced", [True, False])
def test_prefill_tc_groups_that_are_not_powers_of_two(Hq, Hkv, D, forced):
    """GQA groups of 5, 3, 7, 12, 10: the Q tile pads the group to the next power of two (the padding rows hold the
    next kv head's queries, or nothing past the last head, and are never stored).  Forced and through AUTO, where
    these shapes must now reach the tcgen05 kernel (reference head / group coverage: one kernel serves them all,
    ibm-triton-lib/ibm_triton_lib/kernels/triton_unified_attention.py:275-523)."""
    lens = [(300, 300), (64, 1064), (1, 555), (5, 40), (513, 2049), (129, 129)]
    inp = make_paged_inputs(lens, Hq, Hkv, D, 16, dtype=torch.bfloat16, seed=Hq * 7 + D)
    scale = D ** -0.5
    out, kern = _run(inp, scale, PREFILL_TC if forced else DECODE_OR_AUTO)
    # (AUTO launches the tcgen05 kernel last when the single-token row was forked onto the second stream; the split-KV
    #  kernel does not take groups above 8 at head size 256, there the row goes to the mma.sync launch behind it)
    if forced or not (D == 256 and Hq // Hkv > 8):
        assert kern.startswith("attn_prefill_tc"), kern
    assert_close(out, _oracle(inp, scale), atol=BF16_ATOL, rtol=BF16_RTOL, what=f"prefill_tc g={Hq // Hkv} D={D}")
    assert not torch.isnan(out).any()


@pytest.mark.parametrize("window", [64, 128, 300])
@pytest.mark.parametrize("D", [128, 64, 256])
def test_prefill_tc_sliding_window(window, D):
    lens = [(1, 1328), (129, 463), (300, 300), (700, 1500)]
    inp = make_paged_inputs(lens, 8, 2, D, 16, dtype=torch.bfloat16, seed=23)
    scale = D ** -0.5
    out, kern = _run(inp, scale, PREFILL_TC, window=window)
    assert kern.startswith("attn_prefill_tc")
    assert_close(out, _oracle(inp, scale, window), atol=BF16_ATOL, rtol=BF16_RTOL, what=f"window {window}")


@pytest.mark.parametrize("softcap,sinks,window,alibi", [
    (30.0, False, 0, False), (50.0, False, 128, False), (0.0, True, 0, False), (20.0, True, 300, False),
    (0.0, False, 0, True), (30.0, False, 200, True),
])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("D,Hkv", [(128, 2), (64, 2), (64, 1), (128, 8), (256, 2), (256, 8)])
def test_prefill_tc_softcap_alibi_and_sinks(softcap, sinks, window, alibi, dtype, D, Hkv):
    """Softcap (tanh of S before masking), ALiBi (slope * (key position - context length)) and attention
    sinks (epilogue only) inside the tcgen05 kernel; large logits so that the cap actually bends them."""
    lens = [(1, 1328), (129, 463), (300, 300), (700, 1500), (5, 40)]
    Hq = 8 if Hkv < 8 else 40  # (40q / 8kv: a padded group of 5)
    inp = make_paged_inputs(lens, Hq, Hkv, D, 16, dtype=dtype, seed=37)
    inp["q"] = inp["q"] * 3
    sink_t = torch.linspace(-2.0, 6.0, Hq, dtype=torch.float32) if sinks else None
    slopes = torch.tensor([2.0 ** (-(i + 1) * 8.0 / Hq) for i in range(Hq)], dtype=torch.float32) if alibi else None
    scale = 0.4
    out, kern = _run(inp, scale, PREFILL_TC, window=window, softcap=softcap, sinks=sink_t, alibi=slopes)
    assert kern.startswith("attn_prefill_tc")
    ref = _oracle(inp, scale, window=window, softcap=softcap, sinks=sink_t, alibi=slopes)
    assert_close(out, ref, atol=BF16_ATOL, rtol=BF16_RTOL, what=f"prefill_tc softcap={softcap} sinks={sinks} alibi={alibi}")


@pytest.mark.parametrize("D", [128, 64, 256])
def test_prefill_tc_large_logits_exercise_the_lazy_rescale(D):
    """Row maxima that keep growing along the KV axis force O to be rescaled (head size 256: by whichever warpgroup
    owns the tile, with the maximum handed from one to the other)."""
    lens = [(256, 1024)]
    inp = make_paged_inputs(lens, 4, 1, D, 16, dtype=torch.bfloat16, seed=31)
    # keys grow with position so later tiles dominate: max logit climbs by >> 2^8 per tile
    bt = inp["block_table"][0]
    for pos in range(1024):
        inp["k_cache"][int(bt[pos // 16]), pos % 16] *= (1.0 + pos / 64.0)
    q = inp["q"]
    inp["q"] = (q.float().abs() * 2).to(q.dtype)
    kc = inp["k_cache"]
    inp["k_cache"] = kc.float().abs().to(kc.dtype)
    scale = 1.0 if D <= 128 else 0.5
    out, kern = _run(inp, scale, PREFILL_TC)
    assert kern.startswith("attn_prefill_tc")
    assert_close(out, _oracle(inp, scale), atol=BF16_ATOL, rtol=BF16_RTOL, what="lazy rescale")


def test_prefill_tc_ignores_garbage_beyond_seq_len():
    lens = [(100, 117), (40, 40), (200, 700), (64, 905)]
    inp = make_paged_inputs(lens, 8, 2, 128, 16, dtype=torch.bfloat16, seed=9)
    ref = _oracle(inp, 0.1)
    bt = inp["block_table"]
    for b, (_, n) in enumerate(lens):
        if n % 16:
            page = int(bt[b, (n - 1) // 16])
            inp["k_cache"][page, n % 16:] = float("nan")
            inp["v_cache"][page, n % 16:] = float("inf")
    out, _ = _run(inp, 0.1, PREFILL_TC)
    assert_close(out, ref, atol=BF16_ATOL, rtol=BF16_RTOL, what="garbage tail")


def test_mixed_batch_is_routed_across_three_kernels():
    """AUTO on a mixed batch: tcgen05 for long queries, mma.sync for short extends,
    split-KV decode for single-token rows; together they must cover every row."""
    from vllm_triton_backend_b200 import _cabi

    lens = [(1, 1328), (300, 300), (5, 18), (1, 77), (129, 463), (12, 900), (1, 1)]
    inp = make_paged_inputs(lens, 32, 8, 128, 16, dtype=torch.bfloat16, seed=41)
    scale = 128 ** -0.5
    _cabi.reset_launch_count()
    out, kern = _run(inp, scale, DECODE_OR_AUTO)
    # split-KV plan + decode and the mma.sync launch (forked onto a second stream, beside the prompts), then tcgen05
    assert _cabi.launch_count() == 4 and kern.startswith("attn_prefill_tc")
    assert not torch.isnan(out).any()
    assert_close(out, _oracle(inp, scale), atol=BF16_ATOL, rtol=BF16_RTOL, what="mixed routing")


def test_workspace_counters_are_left_clean():
    """The split-KV ticket counters must be zero again after a call (self-cleaning)."""
    from vllm_triton_backend_b200 import ops

    lens = [4096, 100, 1777, 16, 3000]
    inp = make_paged_inputs([(1, n) for n in lens], 32, 8, 128, 16, dtype=torch.bfloat16, seed=3, device="cuda")
    ws = torch.zeros(ops.workspace_bytes(5, 5, 32, 8, 128), dtype=torch.uint8, device="cuda")
    out = torch.empty_like(inp["q"])
    for _ in range(3):
        ops.unified_attention(q=inp["q"], k=inp["k_cache"], v=inp["v_cache"], out=out, cu_seqlens_q=inp["cu_seqlens_q"],
                              max_seqlen_q=1, seqused_k=inp["seq_lens"], max_seqlen_k=4096, softmax_scale=0.088,
                              causal=True, window_size=(-1, -1), block_table=inp["block_table"], softcap=0,
                              q_descale=None, k_descale=None, v_descale=None, workspace=ws)
    torch.cuda.synchronize()
    assert int(ws[: 5 * 8 * 4].view(torch.int32).abs().sum()) == 0
    assert int(ws[: 1 << 20].view(torch.int32).abs().sum()) == 0  # the whole counter region (include/b200attn.h)


@pytest.mark.parametrize("B,fp8,mixed", [(257, False, False), (300, True, False), (1000, False, False), (300, False, True)])
def test_more_than_256_sequences_take_the_shared_memory_partition(B, fp8, mixed):
    """Up to 256 sequences every warp derives the work partition in registers; beyond that warp 0 does and
    hands it over through shared memory.  Ragged lengths, empty sequences, and (mixed) prompts in between."""
    import random

    r = random.Random(B)
    lens = [(1, r.choice([0, 1, 16, 17, r.randint(1, 900), r.randint(1, 3000)])) for _ in range(B)]
    if mixed:
        for i in range(0, B, 37):
            q = r.randint(2, 200)
            lens[i] = (q, q + r.randint(0, 500))
    alloc = [(a, max(b, 1)) for a, b in lens]
    ks, vs = (0.37, 0.61) if fp8 else (1.0, 1.0)
    inp = make_paged_inputs(alloc, 16, 4, 128, 16, dtype=torch.bfloat16, seed=B, kv_dtype=torch.float8_e4m3fn if fp8 else None,
                            k_scale=ks, v_scale=vs)
    inp["seq_lens"] = torch.tensor([b for _, b in lens], dtype=torch.int32)
    scale = 128 ** -0.5
    out, kern = _run(inp, scale, DECODE_OR_AUTO, k_scale=ks, v_scale=vs)
    assert kern.startswith("attn_decode_split_kv")
    assert not torch.isnan(out).any()
    assert_close(out, _oracle(inp, scale, k_scale=ks, v_scale=vs), atol=FP8_ATOL if fp8 else BF16_ATOL, rtol=BF16_RTOL,
                 what=f"{B} sequences")


def test_one_workspace_serves_batches_of_changing_size():
    """Continuous batching: the same workspace sees a small batch, then a larger one, then a small one again.
    The ticket counters of the larger batch must not land where the smaller one left partial results
    (they did once: the counter region was sized by the call's own batch)."""
    from vllm_triton_backend_b200 import ops

    Hq, Hkv, D, bs = 32, 8, 128, 16
    big = ops.workspace_bytes(256, 256, Hq, Hkv, D)
    ws = torch.zeros(big, dtype=torch.uint8, device="cuda")
    scale = D ** -0.5
    for step, B in enumerate([2, 96, 3, 256, 17]):
        lens = [(1, 3000 + 37 * i if i % 3 == 0 else 40 + 11 * i) for i in range(B)]  # long (split over CTAs) and short
        inp = make_paged_inputs(lens, Hq, Hkv, D, bs, dtype=torch.bfloat16, seed=50 + step)
        out = torch.full_like(inp["q"], float("nan")).cuda()
        ops.unified_attention(q=inp["q"].cuda(), k=inp["k_cache"].cuda(), v=inp["v_cache"].cuda(), out=out,
                              cu_seqlens_q=inp["cu_seqlens_q"].cuda(), max_seqlen_q=1, seqused_k=inp["seq_lens"].cuda(),
                              max_seqlen_k=max(b for _, b in lens), softmax_scale=scale, causal=True, window_size=(-1, -1),
                              block_table=inp["block_table"].cuda(), softcap=0, q_descale=None, k_descale=None,
                              v_descale=None, workspace=ws)
        torch.cuda.synchronize()
        assert_close(out.cpu(), _oracle(inp, scale), atol=BF16_ATOL, rtol=BF16_RTOL, what=f"step {step}, batch {B}")
        assert int(ws[: 1 << 20].view(torch.int32).abs().sum()) == 0


# ---------------------------------------------------------------------------------
# shapes only the SIMT kernel takes
# ---------------------------------------------------------------------------------
@pytest.mark.parametrize("D", [32, 80, 96, 120])
def test_odd_head_sizes_run_on_the_generic_kernel(D):
    lens = [(1, 100), (7, 40), (33, 33)]
    inp = make_paged_inputs(lens, 8, 2, D, 16, dtype=torch.bfloat16, seed=D)
    out, kern = _run(inp, D ** -0.5, DECODE_OR_AUTO)
    assert kern == "attn_generic"
    assert_close(out, _oracle(inp, D ** -0.5), atol=BF16_ATOL, rtol=BF16_RTOL, what=f"D={D}")


def test_reference_argument_checks():
    from vllm_triton_backend_b200 import ops

    inp = make_paged_inputs([(1, 10)], 4, 4, 64, 16, device="cuda")
    kw = dict(q=inp["q"], k=inp["k_cache"], v=inp["v_cache"], out=torch.empty_like(inp["q"]),
              cu_seqlens_q=inp["cu_seqlens_q"], max_seqlen_q=1, seqused_k=inp["seq_lens"], max_seqlen_k=10,
              softmax_scale=1.0, window_size=(-1, -1), block_table=inp["block_table"], softcap=0,
              k_descale=None, v_descale=None)
    with pytest.raises(AssertionError, match="causal"):
        ops.unified_attention(causal=False, q_descale=None, **kw)
    with pytest.raises(AssertionError, match="Q scales"):
        ops.unified_attention(causal=True, q_descale=torch.ones(1, device="cuda"), **kw)


# ---------------------------------------------------------------------------------
# attention sinks
# ---------------------------------------------------------------------------------
@pytest.mark.parametrize("path", [GENERIC, PREFILL, DECODE_OR_AUTO])
@pytest.mark.parametrize("fp8", [False, True])
def test_attention_sinks_match_oracle(path, fp8):
    Hq, Hkv, D, bs = 16, 2, 128, 16
    lens = [(1, 700), (1, 33), (1, 4096), (1, 1)] if path is DECODE_OR_AUTO else [(1, 300), (70, 470), (129, 129), (1, 5)]
    ks, vs = (0.37, 0.61) if fp8 else (1.0, 1.0)
    inp = make_paged_inputs(lens, Hq, Hkv, D, bs, seed=23, kv_dtype=torch.float8_e4m3fn if fp8 else None, k_scale=ks,
                            v_scale=vs)
    sinks = torch.linspace(-3.0, 4.0, Hq, dtype=torch.float32)
    scale = D ** -0.5
    out, kern = _run(inp, scale, path, k_scale=ks if fp8 else None, v_scale=vs if fp8 else None, sinks=sinks, window=256)
    ref = _oracle(inp, scale, window=256, k_scale=ks, v_scale=vs, sinks=sinks)
    assert_close(out, ref, atol=FP8_ATOL if fp8 else BF16_ATOL, rtol=BF16_RTOL, what=f"sinks ({kern})")
    # a sink can only shrink the output: compare norms with the sink-free run
    out0, _ = _run(inp, scale, path, k_scale=ks if fp8 else None, v_scale=vs if fp8 else None, window=256)
    assert (out.float().norm(dim=-1) <= out0.float().norm(dim=-1) + 1e-2).all()


# ---------------------------------------------------------------------------------
# HND cache layout: same [NB, bs, Hkv, D] views, memory ordered [NB, Hkv, bs, D]
# ---------------------------------------------------------------------------------
def _as_hnd(t):
    """[NB, bs, H, D] view over memory laid out [NB, H, bs, D] (vLLM's VLLM_KV_CACHE_LAYOUT=HND)."""
    return t.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)


@pytest.mark.parametrize("lens,fp8", [
    ([(1, 700), (1, 33), (1, 4096), (1, 1)], False),            # decode (TMA boxes over HND strides)
    ([(1, 700), (1, 33), (1, 4096), (1, 1)], True),
    ([(300, 300), (64, 1064), (1, 555), (5, 40)], False),        # mixed: tcgen05 + mma.sync + decode
    ([(300, 300), (64, 1064), (1, 555)], True),
])
def test_hnd_cache_layout(lens, fp8):
    from vllm_triton_backend_b200 import ops

    Hq, Hkv, D, bs = 32, 8, 128, 16
    ks, vs = (0.37, 0.61) if fp8 else (1.0, 1.0)
    inp = make_paged_inputs(lens, Hq, Hkv, D, bs, seed=31, kv_dtype=torch.float8_e4m3fn if fp8 else None, k_scale=ks, v_scale=vs)
    scale = D ** -0.5
    ref = _oracle(inp, scale, k_scale=ks, v_scale=vs)
    hnd = dict(inp)
    kc8, vc8 = (inp["k_cache"].view(torch.uint8), inp["v_cache"].view(torch.uint8)) if fp8 else (inp["k_cache"], inp["v_cache"])
    kh, vh = _as_hnd(kc8).to("cuda"), _as_hnd(vc8).to("cuda")
    assert kh.stride(1) == D and kh.stride(2) == bs * D  # really HND in device memory
    if fp8:
        kh, vh = kh.view(torch.float8_e4m3fn), vh.view(torch.float8_e4m3fn)
    hnd["k_cache"], hnd["v_cache"] = kh, vh
    out, kern = _run(hnd, scale, DECODE_OR_AUTO, k_scale=ks if fp8 else None, v_scale=vs if fp8 else None)
    assert_close(out, ref, atol=FP8_ATOL if fp8 else BF16_ATOL, rtol=BF16_RTOL, what=f"HND ({kern})")
    # cache write into an HND cache: same bytes as into an NHD one
    T = 9
    g = torch.Generator().manual_seed(1)
    key = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).bfloat16().cuda()
    val = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).bfloat16().cuda()
    slots = torch.randperm(kh.shape[0] * bs, generator=g)[:T].to(torch.int64).cuda()
    nhd_k, nhd_v = kh.contiguous().clone(), vh.contiguous().clone()
    kvd = "fp8" if fp8 else "auto"
    ops.reshape_and_cache_flash(key, val, nhd_k, nhd_v, slots, kvd, ks, vs)
    ops.reshape_and_cache_flash(key, val, kh, vh, slots, kvd, ks, vs)
    torch.cuda.synchronize()
    as_b = lambda t: t.view(torch.uint8) if fp8 else t
    assert torch.equal(as_b(kh), as_b(nhd_k)) and torch.equal(as_b(vh), as_b(nhd_v))


# ---------------------------------------------------------------------------------
# FP8 prefill through the staging pool + tcgen05 kernel (csrc/kv_stage.cu)
# ---------------------------------------------------------------------------------
def _fp8_prefill(lens, ws_tokens, scales=(0.37, 0.61), Hq=32, Hkv=8, tune=0):
    from vllm_triton_backend_b200 import _cabi, ops

    D, bs = 128, 16
    inp = make_paged_inputs(lens, Hq, Hkv, D, bs, seed=41, kv_dtype=torch.float8_e4m3fn, k_scale=scales[0], v_scale=scales[1])
    dev = "cuda"
    q = inp["q"].to(dev)
    out = torch.full_like(q, float("nan"))
    T, B = q.shape[0], len(lens)
    nbytes = ops.workspace_bytes(T, B, Hq, Hkv, D) + (ops.fp8_staging_bytes(B, ws_tokens, bs, Hkv, D) if ws_tokens else 0)
    ws = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
    sc = [torch.tensor([s], dtype=torch.float32, device=dev) for s in scales]
    _cabi.set_tuning(0, 0, 0, tune)
    _cabi.reset_launch_count()
    try:
        ops.unified_attention(
            q=q, k=inp["k_cache"].to(dev), v=inp["v_cache"].to(dev), out=out, cu_seqlens_q=inp["cu_seqlens_q"].to(dev),
            max_seqlen_q=max(a for a, _ in lens), seqused_k=inp["seq_lens"].to(dev), max_seqlen_k=max(b for _, b in lens),
            softmax_scale=D ** -0.5, causal=True, window_size=(-1, -1), block_table=inp["block_table"].to(dev), softcap=0,
            q_descale=None, k_descale=sc[0], v_descale=sc[1], workspace=ws)
        torch.cuda.synchronize()
    finally:
        _cabi.set_tuning(0, 0, 0, 0)
    ref = _oracle(inp, D ** -0.5, k_scale=scales[0], v_scale=scales[1])
    return out.cpu(), ref, _cabi.launch_count()


def test_fp8_prefill_is_staged_for_the_tcgen05_kernel():
    lens = [(300, 300), (1, 555), (64, 1064), (5, 40), (513, 2049)]
    out, ref, launches = _fp8_prefill(lens, ws_tokens=4096)
    assert launches == 6  # stage plan, dequant, tcgen05, mma.sync (short extend), split-KV plan, decode row
    assert_close(out, ref, atol=FP8_ATOL, rtol=BF16_RTOL, what="fp8 prefill, staged")
    # the staged values are exactly what the reference forms on the fly, so bf16-level agreement holds too
    assert_close(out, ref, atol=BF16_ATOL, rtol=BF16_RTOL, what="fp8 prefill, staged (bf16 tolerance)")


def test_fp8_prefill_partial_staging_falls_back_per_sequence():
    lens = [(300, 300), (64, 1064), (513, 2049), (200, 900)]
    full, ref, _ = _fp8_prefill(lens, ws_tokens=8192)
    part, _, launches = _fp8_prefill(lens, ws_tokens=1500)   # room for the first two sequences only
    none, _, n0 = _fp8_prefill(lens, ws_tokens=0)             # no pool: mma.sync kernel for all
    off, _, n1 = _fp8_prefill(lens, ws_tokens=8192, tune=256)  # switched off by the tuning word
    assert launches == 6 and n0 == n1 == 3  # stage plan, dequant, tcgen05, mma.sync, split-KV plan, decode  |  mma.sync, plan, decode
    for name, o in (("full", full), ("partial", part), ("no pool", none), ("off", off)):
        assert_close(o, ref, atol=FP8_ATOL, rtol=BF16_RTOL, what=f"fp8 prefill ({name})")
        assert not torch.isnan(o).any()


def test_fp8_prefill_staging_in_a_cuda_graph():
    from vllm_triton_backend_b200 import ops

    Hq, Hkv, D, bs = 32, 8, 128, 16
    lens = [(256, 256), (128, 1152)]
    inp = make_paged_inputs(lens, Hq, Hkv, D, bs, seed=43, kv_dtype=torch.float8_e4m3fn, k_scale=0.5, v_scale=0.25)
    dev = "cuda"
    q, out = inp["q"].to(dev), torch.zeros_like(inp["q"]).to(dev)
    kc, vc = inp["k_cache"].to(dev), inp["v_cache"].to(dev)
    cu, bt = inp["cu_seqlens_q"].to(dev), inp["block_table"].to(dev)
    sl = torch.tensor([256, 128], dtype=torch.int32, device=dev)  # capture with context-free lengths
    ws = torch.zeros(ops.workspace_bytes(384, 2, Hq, Hkv, D) + ops.fp8_staging_bytes(2, 2048, bs, Hkv, D), dtype=torch.uint8, device=dev)
    sc = torch.tensor([0.5], device=dev), torch.tensor([0.25], device=dev)

    def call():
        ops.unified_attention(q=q, k=kc, v=vc, out=out, cu_seqlens_q=cu, max_seqlen_q=256, seqused_k=sl, max_seqlen_k=1152,
                              softmax_scale=D ** -0.5, causal=True, window_size=(-1, -1), block_table=bt, softcap=0,
                              q_descale=None, k_descale=sc[0], v_descale=sc[1], workspace=ws)
    call()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        call()
    sl.copy_(inp["seq_lens"].to(dev))
    out.fill_(float("nan"))
    g.replay()
    torch.cuda.synchronize()
    assert_close(out.cpu(), _oracle(inp, D ** -0.5, k_scale=0.5, v_scale=0.25), atol=FP8_ATOL, rtol=BF16_RTOL, what="staged, graph")


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("D", [64, 128])
@pytest.mark.parametrize("force,kernel", [(None, "attn_decode_split_kv"), (22, "attn_prefill_mma"), (4, "attn_prefill_tc"), (1, "attn_generic")])
def test_qq_bias_between_the_query_tokens_of_a_step(dtype, D, force, kernel):
    """Upstream `qq_bias` (tree / speculative-decoding masks): a [q, q] bias on the logits between a step's own
    query tokens, the same for every sequence and head; -inf entries cut a query off from a sibling branch.
    Applied inside every kernel (the reference applies it inside its one kernel,
    vllm/attention/ops/triton_unified_attention.py:66-79, 176-177, 301-307): AUTO sends the 8-token rows to the
    mma.sync kernel and the single-token row to the split-KV kernel."""
    from vllm_triton_backend_b200 import _cabi, ops

    lens = [(8, 700), (8, 8), (8, 33), (5, 1200), (1, 90)]
    Hq, Hkv, bs = 8, 2, 16
    inp = make_paged_inputs(lens, Hq, Hkv, D, bs, dtype=dtype, seed=61)
    g = torch.Generator().manual_seed(7)
    bias = torch.randn(8, 8, generator=g)
    bias[torch.rand(8, 8, generator=g) < 0.3] = float("-inf")
    bias.fill_diagonal_(0.0)  # a token always sees itself
    bias[0, 0] = 0.75         # (and the single-token row must pick up [0][0])
    scale = D ** -0.5
    dev = "cuda"
    out = torch.full_like(inp["q"], float("nan")).to(dev)
    ops.unified_attention(
        q=inp["q"].to(dev), k=inp["k_cache"].to(dev), v=inp["v_cache"].to(dev), out=out, cu_seqlens_q=inp["cu_seqlens_q"].to(dev),
        max_seqlen_q=8, seqused_k=inp["seq_lens"].to(dev), max_seqlen_k=1200, softmax_scale=scale, causal=True,
        window_size=(-1, -1), block_table=inp["block_table"].to(dev), softcap=0, q_descale=None, k_descale=None,
        v_descale=None, qq_bias=bias.to(dev), force_selection=force)
    torch.cuda.synchronize()
    assert _cabi.last_kernel() == kernel
    ref = unified_attention_ref(inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"],
                                inp["block_table"], scale, out_dtype=torch.float32, qq_bias=bias)
    assert_close(out.cpu(), ref, atol=BF16_ATOL, rtol=BF16_RTOL, what=f"qq_bias via {kernel}")


def test_qq_bias_with_long_trees_and_other_features():
    """qq_bias together with a sliding window, softcap and FP8 KV, on queries long enough for the tcgen05 kernel."""
    from vllm_triton_backend_b200 import _cabi, ops

    lens = [(40, 700), (33, 33), (1, 90), (64, 2000), (7, 300)]
    Hq, Hkv, D, bs = 16, 4, 128, 16
    dev = "cuda"
    for fp8 in (False, True):
        inp = make_paged_inputs(lens, Hq, Hkv, D, bs, dtype=torch.bfloat16, seed=62, kv_dtype=torch.float8_e4m3fn if fp8 else None,
                                k_scale=0.5, v_scale=0.5)
        g = torch.Generator().manual_seed(8)
        bias = torch.randn(64, 64, generator=g)
        bias[torch.rand(64, 64, generator=g) < 0.3] = float("-inf")
        bias.fill_diagonal_(0.0)
        sc = torch.tensor([0.5], dtype=torch.float32, device=dev) if fp8 else None
        out = torch.full_like(inp["q"], float("nan")).to(dev)
        ops.unified_attention(
            q=inp["q"].to(dev), k=inp["k_cache"].to(dev), v=inp["v_cache"].to(dev), out=out, cu_seqlens_q=inp["cu_seqlens_q"].to(dev),
            max_seqlen_q=64, seqused_k=inp["seq_lens"].to(dev), max_seqlen_k=2000, softmax_scale=0.2, causal=True,
            window_size=(511, 0), block_table=inp["block_table"].to(dev), softcap=30.0, q_descale=None, k_descale=sc,
            v_descale=sc, qq_bias=bias.to(dev))
        torch.cuda.synchronize()
        ref = unified_attention_ref(inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"],
                                    inp["block_table"], 0.2, sliding_window=512, softcap=30.0, k_scale=0.5 if fp8 else 1.0,
                                    v_scale=0.5 if fp8 else 1.0, out_dtype=torch.float32, qq_bias=bias)
        assert_close(out.cpu(), ref, atol=FP8_ATOL if fp8 else BF16_ATOL, rtol=BF16_RTOL, what=f"qq_bias + features fp8={fp8}")

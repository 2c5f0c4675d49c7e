"""Fused static per-tensor FP8 quantisation of the attention output (SURVEY.md 8f-2;
`output_scale` of AttentionImpl.forward / upstream `unified_attention`): every kernel of the path
writes e4m3 from its fp32 epilogue value, `out = e4m3_rn(clamp(o * (1 / output_scale), +-448))`
(installed vLLM vllm/v1/attention/ops/triton_unified_attention.py:608-610, 722-724, 986).

Checked against the oracle (oracle/paged_attention.py, `output_scale=`) and the upstream Triton
kernel on the same device tensors.  The fp32 value under the rounding differs by bf16-level noise
between implementations, so (i) the fused output must equal, byte for byte up to a neighbouring code on
rounding boundaries, the same kernel's 16-bit output quantised on the host, and (ii) against the
oracle / Triton the dequantised values agree to the unquantised tolerance plus one e4m3 step."""
import pytest
import torch

import triton_reference
from helpers import make_paged_inputs
from oracle.paged_attention import unified_attention_ref

pytestmark = pytest.mark.gpu

GENERIC, PREFILL, AUTO, PREFILL_TC = 1, 22, None, 4


def _e4m3_step(x: torch.Tensor) -> torch.Tensor:
    """Spacing of e4m3 codes around |x| (3 mantissa bits; subnormal step 2^-9)."""
    a = x.abs().clamp_min(2.0 ** -6)
    return torch.exp2(torch.floor(torch.log2(a)) - 3)


def _assert_same_codes(ours_u8: torch.Tensor, ref_u8: torch.Tensor, what: str, min_equal=0.97):
    """Strict: same fp32 value under the rounding up to 16-bit noise => same byte, or the neighbouring code."""
    a = ours_u8.view(torch.float8_e4m3fn).float()
    b = ref_u8.view(torch.float8_e4m3fn).float()
    assert not torch.isnan(a).any(), f"{what}: NaN code in the output"
    equal = (ours_u8 == ref_u8) | ((a == 0) & (b == 0))
    frac = equal.float().mean().item()
    worst = ((a - b).abs() / _e4m3_step(torch.maximum(a.abs(), b.abs()))).max().item()
    assert worst <= 1.0 + 1e-6, f"{what}: codes differ by {worst:.2f} e4m3 steps"
    assert frac >= min_equal, f"{what}: only {frac:.4f} of the bytes are identical"


def _assert_close_codes(ours_u8: torch.Tensor, ref_u8: torch.Tensor, out_scale: float, what: str, atol=1e-2):
    """Against another implementation of the attention itself: its fp32 result differs from ours by the
    16-bit noise the unquantised parity tests allow (BASELINE.json: 2e-2 abs), so the dequantised
    outputs agree to that plus one e4m3 step."""
    a = ours_u8.view(torch.float8_e4m3fn).float()
    b = ref_u8.view(torch.float8_e4m3fn).float()
    assert not torch.isnan(a).any(), f"{what}: NaN code in the output"
    excess = ((a - b).abs() - _e4m3_step(torch.maximum(a.abs(), b.abs()))) * out_scale
    assert excess.max().item() <= atol, f"{what}: dequantised outputs differ by one step + {excess.max().item():.4f}"


def _host_quant(out16: torch.Tensor, out_scale: float) -> torch.Tensor:
    inv = 1.0 / torch.tensor(out_scale, dtype=torch.float32)
    return (out16.float() * inv).clamp(-448.0, 448.0).to(torch.float8_e4m3fn).view(torch.uint8)


def _run(lens, Hq, Hkv, D, bs, force, out_scale, dtype=torch.bfloat16, fp8_kv=False, sinks=False, window=0, seed=0):
    from vllm_triton_backend_b200 import _cabi, ops

    dev = "cuda"
    ks, vs = (0.37, 0.61) if fp8_kv else (1.0, 1.0)
    inp = make_paged_inputs(lens, Hq, Hkv, D, bs, dtype=dtype, seed=seed,
                            kv_dtype=torch.float8_e4m3fn if fp8_kv else None, k_scale=ks, v_scale=vs)
    scale = D ** -0.5
    max_q, max_k = max(a for a, _ in lens), max(b for _, b in lens)
    sink_t = torch.linspace(-2.0, 3.0, Hq, dtype=torch.float32) if sinks else None
    d = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in inp.items()}
    out = torch.full(inp["q"].shape, 0x7F, dtype=torch.uint8, device=dev).view(torch.float8_e4m3fn)  # NaN pattern
    os_t = torch.tensor([out_scale], dtype=torch.float32, device=dev)
    kw = dict(
        q=d["q"], k=d["k_cache"], v=d["v_cache"], cu_seqlens_q=d["cu_seqlens_q"], max_seqlen_q=max_q,
        seqused_k=d["seq_lens"], max_seqlen_k=max_k, softmax_scale=scale, causal=True,
        window_size=(window - 1, 0) if window > 0 else (-1, -1), block_table=d["block_table"], softcap=0.0, q_descale=None,
        k_descale=torch.tensor([ks], dtype=torch.float32, device=dev),
        v_descale=torch.tensor([vs], dtype=torch.float32, device=dev),
        sinks=None if sink_t is None else sink_t.to(dev),
    )
    ops.unified_attention(out=out, output_scale=os_t, force_selection=force, **kw)
    torch.cuda.synchronize()
    kern = _cabi.last_kernel()
    out16 = torch.full_like(d["q"], float("nan"))  # the same kernels without the fused quantisation
    ops.unified_attention(out=out16, force_selection=force, **kw)
    torch.cuda.synchronize()
    _assert_same_codes(out.cpu().view(torch.uint8), _host_quant(out16.cpu(), out_scale), f"{kern}: fused vs unfused quantisation")
    ref = unified_attention_ref(inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"],
                                inp["block_table"], scale, sliding_window=window, k_scale=ks, v_scale=vs, sinks=sink_t,
                                output_scale=out_scale)
    return out.cpu().view(torch.uint8), ref.view(torch.uint8), kern, kw, os_t


DECODE_LENS = [(1, 4096), (1, 33), (1, 1), (1, 700), (1, 2049), (1, 16)]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("fp8_kv", [False, True])
def test_decode_direct_and_merged_items_quantise_the_output(dtype, fp8_kv):
    # long sequences are split over CTAs (merge epilogue), short ones finish in one CTA (direct epilogue)
    ours, ref, kern, _, _ = _run(DECODE_LENS, 32, 8, 128, 16, AUTO, out_scale=0.02, dtype=dtype, fp8_kv=fp8_kv, seed=3)
    assert kern.startswith("attn_decode_split_kv")
    _assert_close_codes(ours, ref, 0.02, "decode fp8 output")


@pytest.mark.parametrize("force,expect", [(GENERIC, "attn_generic"), (PREFILL, "attn_prefill_mma"), (PREFILL_TC, "attn_prefill_tc")])
def test_each_prefill_kernel_quantises_the_output(force, expect):
    lens = [(300, 300), (64, 1064), (129, 463), (256, 256)]
    ours, ref, kern, _, _ = _run(lens, 16, 4, 128, 16, force, out_scale=0.015, seed=5)
    assert kern.startswith(expect), kern
    _assert_close_codes(ours, ref, 0.015, f"{kern} fp8 output")


def test_mixed_batch_auto_dispatch_with_fp8_cache_and_output():
    # fp8 cache + fp8 output: staging -> tcgen05, mma.sync for the short extend, split-KV decode rows
    lens = [(300, 300), (1, 555), (64, 1064), (5, 40), (513, 2049), (1, 1)]
    ours, ref, kern, _, _ = _run(lens, 32, 8, 128, 16, AUTO, out_scale=0.01, fp8_kv=True, seed=7)
    _assert_close_codes(ours, ref, 0.01, "mixed fp8 kv + fp8 output", atol=3e-2)


def test_sinks_window_and_small_heads():
    ours, ref, kern, _, _ = _run([(1, 900), (200, 700), (1, 40), (333, 333)], 16, 2, 64, 16, AUTO, out_scale=0.02,
                                 sinks=True, window=128, seed=9)
    _assert_close_codes(ours, ref, 0.02, "sinks + window, D=64")


def test_saturation_and_empty_sequences():
    from vllm_triton_backend_b200 import ops

    # a scale far too small: every value saturates to +-448 (0x7E / 0xFE), never NaN (0x7F)
    ours, ref, _, _, _ = _run([(1, 300), (40, 40)], 8, 2, 128, 16, AUTO, out_scale=1e-12, seed=11)
    assert set(ours.unique().tolist()) <= {0x7E, 0xFE}
    assert (ours == ref).float().mean().item() > 0.99  # only the sign is left; it differs where the value is ~0
    # a sequence with nothing visible gets 0x00 bytes (reference reduce_segments :828)
    dev = "cuda"
    inp = make_paged_inputs([(1, 100), (1, 1), (1, 50)], 8, 2, 128, 16, seed=13)
    sl = inp["seq_lens"].clone()
    sl[1] = 0
    out = torch.full(inp["q"].shape, 0x7F, dtype=torch.uint8, device=dev).view(torch.float8_e4m3fn)
    ops.unified_attention(
        q=inp["q"].to(dev), k=inp["k_cache"].to(dev), v=inp["v_cache"].to(dev), out=out,
        cu_seqlens_q=inp["cu_seqlens_q"].to(dev), max_seqlen_q=1, seqused_k=sl.to(dev), max_seqlen_k=100,
        softmax_scale=128 ** -0.5, causal=True, window_size=(-1, -1), block_table=inp["block_table"].to(dev), softcap=0,
        q_descale=None, k_descale=None, v_descale=None, output_scale=torch.tensor([0.02], device=dev))
    torch.cuda.synchronize()
    got = out.cpu().view(torch.uint8)
    assert (got[1] == 0).all() and (got[0] != 0x7F).all() and (got[2] != 0x7F).all()


@pytest.mark.parametrize("lens", [DECODE_LENS, [(300, 300), (64, 1064), (1, 555)]], ids=["decode", "mixed"])
def test_matches_the_triton_kernels_output_scale(lens):
    why = triton_reference.available()
    if why:
        pytest.skip(why)
    from vllm.v1.attention.ops.triton_unified_attention import unified_attention as triton_unified_attention

    ours, _, _, kw, os_t = _run(lens, 32, 8, 128, 16, AUTO, out_scale=0.02, seed=15)
    ref = torch.zeros(kw["q"].shape, dtype=torch.uint8, device="cuda").view(torch.float8_e4m3fn)
    tr = triton_reference.TritonReference(len(lens), 32, 128)
    max_q = kw["max_seqlen_q"]
    extra = {}
    if max_q == 1:
        extra = dict(seq_threshold_3D=tr.max_seqs, num_par_softmax_segments=tr.NUM_SEGMENTS,
                     softmax_segm_output=tr.segm_output, softmax_segm_max=tr.segm_max, softmax_segm_expsum=tr.segm_expsum)
    triton_unified_attention(
        q=kw["q"], k=kw["k"], v=kw["v"], out=ref, cu_seqlens_q=kw["cu_seqlens_q"], max_seqlen_q=max_q,
        seqused_k=kw["seqused_k"], max_seqlen_k=kw["max_seqlen_k"], softmax_scale=kw["softmax_scale"], causal=True,
        window_size=(-1, -1), block_table=kw["block_table"], softcap=0.0, q_descale=None, k_descale=kw["k_descale"],
        v_descale=kw["v_descale"], output_scale=os_t, **extra)
    torch.cuda.synchronize()
    _assert_close_codes(ours, ref.cpu().view(torch.uint8), 0.02, "vs triton output_scale")


def test_impl_forward_with_output_scale_and_argument_checks():
    from vllm_triton_backend_b200 import _cabi, ops
    from vllm_triton_backend_b200.backend.attn import B200AttentionImpl, B200AttentionMetadata

    dev = "cuda"
    Hq, Hkv, D, bs = 32, 8, 128, 16
    lens = [(1, 700), (1, 33), (1, 3000)]
    inp = make_paged_inputs(lens, Hq, Hkv, D, bs, seed=17)
    T, B = 3, 3
    kv = torch.stack([inp["k_cache"], inp["v_cache"]]).to(dev)
    ws = torch.zeros(ops.workspace_bytes(T, B, Hq, Hkv, D), dtype=torch.uint8, device=dev)
    meta = B200AttentionMetadata(num_actual_tokens=T, max_query_len=1, query_start_loc=inp["cu_seqlens_q"].to(dev),
                                 max_seq_len=3000, seq_lens=inp["seq_lens"].to(dev), block_table=inp["block_table"].to(dev),
                                 slot_mapping=torch.full((T,), -1, dtype=torch.int64, device=dev), workspace=ws)
    impl = B200AttentionImpl(Hq, D, D ** -0.5, Hkv, None, None, "auto")

    class L:
        _k_scale = torch.ones(1, device=dev)
        _v_scale = torch.ones(1, device=dev)

    os_t = torch.tensor([0.02], dtype=torch.float32, device=dev)
    out = torch.zeros(T, Hq * D, dtype=torch.float8_e4m3fn, device=dev)
    impl.forward(L, inp["q"].to(dev), None, None, kv, meta, output=out, output_scale=os_t)
    torch.cuda.synchronize()
    ref = unified_attention_ref(inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"],
                                inp["block_table"], D ** -0.5, output_scale=0.02)
    _assert_close_codes(out.cpu().view(torch.uint8).view(T, Hq, D), ref.view(torch.uint8), 0.02, "impl.forward fp8 output")
    # a 16-bit output with a scale, or an fp8 output without one, is a caller bug
    with pytest.raises(ValueError, match="float8_e4m3fn"):
        impl.forward(L, inp["q"].to(dev), None, None, kv, meta, output=torch.zeros(T, Hq * D, dtype=torch.bfloat16, device=dev),
                     output_scale=os_t)
    with pytest.raises(ValueError, match="out dtype"):
        impl.forward(L, inp["q"].to(dev), None, None, kv, meta, output=out)
    # rows that are not 8-byte aligned are refused by the library
    bad = torch.zeros(T * Hq * D + 8, dtype=torch.uint8, device=dev)[4:4 + T * Hq * D].view(torch.float8_e4m3fn).view(T, Hq, D)
    with pytest.raises(_cabi.B200AttnError, match="ALIGNMENT"):
        ops.unified_attention(
            q=inp["q"].to(dev), k=kv[0], v=kv[1], out=bad, cu_seqlens_q=meta.query_start_loc, max_seqlen_q=1,
            seqused_k=meta.seq_lens, max_seqlen_k=3000, softmax_scale=D ** -0.5, causal=True, window_size=(-1, -1),
            block_table=meta.block_table, softcap=0, q_descale=None, k_descale=None, v_descale=None, output_scale=os_t)

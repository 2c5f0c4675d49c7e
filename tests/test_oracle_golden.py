"""Pins the CPU oracle to the reference: golden vectors were produced by the
reference's own Triton kernels (CPU interpreter) and torch functions
(oracle/make_golden.py)."""
import numpy as np
import pytest
import torch

from oracle.cache_write import reshape_and_cache_flash_ref, reshape_and_cache_flash_c, load_c_lib
from oracle.paged_attention import (
    paged_attn_torch_ref, reduce_segments_ref, split_kv_partials_ref, unified_attention_ref,
)
from helpers import GOLDEN_DIR, assert_close, golden_attn_files, load_golden_attn


@pytest.mark.parametrize("path", golden_attn_files(), ids=lambda p: p.split("attn_")[-1][:-4])
def test_oracle_matches_reference_triton_kernels(path):
    c = load_golden_attn(path)
    out = unified_attention_ref(
        c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["seq_lens"], c["block_table"], c["scale"],
        sliding_window=c["window"], softcap=c["softcap"], alibi_slopes=c["alibi_slopes"],
        round_p_to=c["dtype"] if c["dtype"] != torch.float32 else None, out_dtype=torch.float32,
    )
    # fp32 cases: summation-order noise only; fp16 cases: the kernel rounds the output to fp16
    atol = 2e-5 if c["dtype"] == torch.float32 else 1e-3
    assert_close(out, c["out_triton"], atol=atol, rtol=1e-5 if c["dtype"] == torch.float32 else 2e-3, what=c["name"])


@pytest.mark.parametrize("path", [p for p in golden_attn_files() if "alibi" not in p],
                         ids=lambda p: p.split("attn_")[-1][:-4])
def test_oracle_matches_reference_torch_function(path):
    c = load_golden_attn(path, dtype=torch.float32)
    assert c["out_torch"] is not None
    q_lens = (c["cu_seqlens_q"][1:] - c["cu_seqlens_q"][:-1]).tolist()
    out = paged_attn_torch_ref(
        c["q"], c["k_cache"], c["v_cache"], q_lens, c["seq_lens"].tolist(), c["block_table"], c["scale"],
        sliding_window=c["window"] if c["window"] > 0 else None, soft_cap=c["softcap"] if c["softcap"] > 0 else None,
    )
    assert_close(out, c["out_torch"], atol=1e-6, rtol=1e-6, what=c["name"])
    # and the kernel-semantics oracle agrees with the reference torch function too
    out2 = unified_attention_ref(
        c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["seq_lens"], c["block_table"], c["scale"],
        sliding_window=c["window"], softcap=c["softcap"], out_dtype=torch.float32,
    )
    assert_close(out2, c["out_torch"], atol=2e-5, rtol=1e-5, what=c["name"] + " (kernel semantics)")


def test_split_kv_then_reduce_equals_dense():
    """kernel_unified_attention_3d + reduce_segments restatement == dense oracle == golden."""
    c = load_golden_attn(f"{GOLDEN_DIR}/attn_decode3d_f32.npz")
    acc, M, L = split_kv_partials_ref(c["q"], c["k_cache"], c["v_cache"], c["seq_lens"], c["block_table"], c["scale"])
    out = reduce_segments_ref(acc, M, L)
    assert_close(out, c["out_triton"], atol=2e-5, rtol=1e-5, what="3d restatement")


def test_cache_write_oracle_matches_reference_torch_function():
    z = np.load(f"{GOLDEN_DIR}/cache_write_bf16.npz")
    key = torch.from_numpy(z["key"]).view(torch.bfloat16)
    value = torch.from_numpy(z["value"]).view(torch.bfloat16)
    slots = torch.from_numpy(z["slot_mapping"])
    kc = torch.zeros(z["k_cache"].shape, dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    reshape_and_cache_flash_ref(key, value, kc, vc, slots)
    assert np.array_equal(kc.view(torch.int16).numpy(), z["k_cache"])
    assert np.array_equal(vc.view(torch.int16).numpy(), z["v_cache"])
    if load_c_lib() is not None:
        kc2 = torch.zeros_like(kc)
        vc2 = torch.zeros_like(kc)
        reshape_and_cache_flash_c(key, value, kc2, vc2, slots)
        assert torch.equal(kc2.view(torch.int16), kc.view(torch.int16))
        assert torch.equal(vc2.view(torch.int16), vc.view(torch.int16))


def test_cache_write_skips_negative_slots_and_uses_slot_mapping_length():
    key = torch.randn(6, 2, 16).bfloat16()
    value = torch.randn(6, 2, 16).bfloat16()
    kc = torch.full((4, 16, 2, 16), 7.0).bfloat16()
    vc = kc.clone()
    slots = torch.tensor([3, -1, 20, 63], dtype=torch.int64)  # 4 < key.shape[0]
    reshape_and_cache_flash_ref(key, value, kc, vc, slots)
    assert torch.equal(kc[0, 3], key[0]) and torch.equal(kc[1, 4], key[2]) and torch.equal(vc[3, 15], value[3])
    untouched = torch.ones(4, 16, dtype=torch.bool)
    untouched[0, 3] = untouched[1, 4] = untouched[3, 15] = False
    assert (kc[untouched] == 7.0).all() and (vc[untouched] == 7.0).all()


def test_output_scale_quantises_the_fp32_result():
    """`output_scale` (upstream epilogue: acc * (1/scale), clamp +-448, e4m3 store) applied to a golden case."""
    c = load_golden_attn(f"{GOLDEN_DIR}/attn_decode3d_f32.npz")
    args = (c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["seq_lens"], c["block_table"], c["scale"])
    dense = unified_attention_ref(*args, out_dtype=torch.float32)
    q8 = unified_attention_ref(*args, output_scale=0.01)
    assert q8.dtype == torch.float8_e4m3fn and q8.shape == dense.shape
    x = dense / 0.01
    step = torch.exp2(torch.floor(torch.log2(x.abs().clamp_min(2.0 ** -6))) - 3)
    assert ((q8.float() - x).abs() <= 0.5 * step * (1 + 1e-5)).all(), "not the nearest e4m3 code"
    sat = unified_attention_ref(*args, output_scale=1e-12).view(torch.uint8)
    assert set(sat.unique().tolist()) <= {0x7E, 0xFE}, "saturates to +-448, never NaN"


def test_qq_bias_is_an_additive_bias_on_the_query_query_block():
    """Zero bias changes nothing; -inf between query i and an earlier query j of the same step equals dropping key ctx+j."""
    c = load_golden_attn(f"{GOLDEN_DIR}/attn_mixed2d_f32.npz")
    args = (c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["seq_lens"], c["block_table"], c["scale"])
    q_max = int((c["cu_seqlens_q"][1:] - c["cu_seqlens_q"][:-1]).max())
    base = unified_attention_ref(*args, out_dtype=torch.float32)
    same = unified_attention_ref(*args, out_dtype=torch.float32, qq_bias=torch.zeros(q_max, q_max))
    assert torch.equal(base, same)
    if q_max >= 2:
        bias = torch.zeros(q_max, q_max)
        bias[1:, 0] = float("-inf")  # nobody but the first query token sees the first query token's key
        cut = unified_attention_ref(*args, out_dtype=torch.float32, qq_bias=bias)
        cu = c["cu_seqlens_q"].tolist()
        first_rows = torch.tensor([cu[b] for b in range(len(cu) - 1) if cu[b + 1] > cu[b]])
        assert torch.allclose(cut[first_rows], base[first_rows], atol=1e-6)  # row 0 of every sequence is unaffected
        multi = [b for b in range(len(cu) - 1) if cu[b + 1] - cu[b] >= 2]
        assert multi and any(not torch.allclose(cut[cu[b] + 1], base[cu[b] + 1], atol=1e-6) for b in multi)

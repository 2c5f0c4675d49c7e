"""`b200attn_decode_attention_static_scratch`: the decode launch whose scratch is a device array inside the library
(include/b200attn.h; for callers that cannot allocate).  Called through the C ABI and checked against the existing entry
point `b200attn_unified_attention` on the same inputs, on two calls in a row (the second with other lengths: the array
keeps the first call's partial records and must not let them through).

Named test_zz_* so that pytest collects it last: a failure here must not hide the
other tests."""
import ctypes

import pytest
import torch

from helpers import BF16_ATOL, BF16_RTOL, FP8_ATOL, assert_close, make_paged_inputs

pytestmark = pytest.mark.gpu


def _params(inp, out, scale, workspace, k_scale=None):
    from vllm_triton_backend_b200 import _cabi, ops

    sl = inp["seq_lens"]
    return ops.build_attention_params(
        inp["q"], inp["k_cache"], inp["v_cache"], out, inp["cu_seqlens_q"], 1, sl, int(sl.max()), scale, (-1, -1),
        inp["block_table"], 0.0, k_scale, k_scale, None, _cabi.PATH_DECODE, workspace)


def _both(lens, Hq, Hkv, D, seed, fp8=False):
    """(static-scratch result, existing entry point's result, oracle-free) on one seeded decode batch."""
    from vllm_triton_backend_b200 import _cabi, ops

    lib = _cabi.load_library()
    inp = make_paged_inputs([(1, n) for n in lens], Hq, Hkv, D, 16, dtype=torch.bfloat16, seed=seed, device="cuda",
                            kv_dtype=torch.float8_e4m3fn if fp8 else None, k_scale=0.4, v_scale=0.4)
    sc = torch.tensor([0.4], dtype=torch.float32, device="cuda") if fp8 else None
    scale = D ** -0.5
    stream = torch.cuda.current_stream().cuda_stream
    # existing entry point, caller-owned workspace
    ws = torch.zeros(ops.workspace_bytes(len(lens), len(lens), Hq, Hkv, D), dtype=torch.uint8, device="cuda")
    ref = torch.full_like(inp["q"], float("nan"))
    p_ref = _params(inp, ref, scale, ws, sc)
    _cabi.check(lib.b200attn_unified_attention(ctypes.byref(p_ref), stream))
    # new entry point: no workspace in the parameter block at all
    out = torch.full_like(inp["q"], float("nan"))
    p = _params(inp, out, scale, None, sc)
    assert not p.workspace and p.workspace_bytes == 0
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.b200attn_decode_attention_static_scratch(ctypes.byref(p), stream)
    _cabi.check(rc)
    assert _cabi.last_kernel() == "attn_decode_split_kv"
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before  # nothing went through torch's allocator either
    return out.cpu(), ref.cpu()


def test_static_scratch_matches_the_existing_entry_point_on_two_calls_in_a_row():
    # first call: long sequences, every item split across many CTAs (partials + merge through the array)
    out, ref = _both([4096, 700, 33, 2500, 1, 8000, 257, 4096], 32, 8, 128, seed=1)
    assert not torch.isnan(out).any()
    assert_close(out, ref, atol=BF16_ATOL, rtol=BF16_RTOL, what="static scratch, call 1")
    # second call straight after it: other lengths, other head geometry, the array still holds call 1's records
    out2, ref2 = _both([5000, 64, 900, 3, 1200], 16, 2, 128, seed=2)
    assert not torch.isnan(out2).any()
    assert_close(out2, ref2, atol=BF16_ATOL, rtol=BF16_RTOL, what="static scratch, call 2")
    # the split is static, so the two entry points do the same arithmetic in the same order
    assert torch.equal(out, ref) and torch.equal(out2, ref2)


def test_static_scratch_fp8_cache_and_repeat_of_the_same_call():
    a, ref = _both([6000, 128, 3000, 17], 8, 1, 128, seed=3, fp8=True)
    b, _ = _both([6000, 128, 3000, 17], 8, 1, 128, seed=3, fp8=True)
    assert_close(a, ref, atol=FP8_ATOL, rtol=BF16_RTOL, what="static scratch, fp8 cache")
    assert torch.equal(a, b)


def test_static_scratch_rejects_what_does_not_fit_or_is_not_decode():
    from vllm_triton_backend_b200 import _cabi, ops

    lib = _cabi.load_library()
    stream = torch.cuda.current_stream().cuda_stream
    # 64 query heads in groups of 8 at head size 128: 44 MiB of scratch, the array holds 24
    Hq, Hkv, D = 64, 8, 128
    assert ops.workspace_bytes(4, 4, Hq, Hkv, D) > 24 << 20
    inp = make_paged_inputs([(1, 100)] * 4, Hq, Hkv, D, 16, dtype=torch.bfloat16, seed=4, device="cuda")
    out = torch.full_like(inp["q"], 7.0)
    p = _params(inp, out, D ** -0.5, None)
    rc = lib.b200attn_decode_attention_static_scratch(ctypes.byref(p), stream)
    assert rc == 4  # B200ATTN_E_WORKSPACE
    msg = lib.b200attn_last_error().decode()
    assert "B200ATTN_STATIC_SCRATCH_BYTES" in msg and str(24 << 20) in msg
    torch.cuda.synchronize()
    assert (out == 7.0).all()  # nothing ran
    # a batch with a multi-token query is not this entry point's
    inp = make_paged_inputs([(1, 100), (5, 60)], 8, 2, 128, 16, dtype=torch.bfloat16, seed=5, device="cuda")
    out = torch.empty_like(inp["q"])
    p = ops.build_attention_params(inp["q"], inp["k_cache"], inp["v_cache"], out, inp["cu_seqlens_q"], 5, inp["seq_lens"], 100,
                                   0.1, (-1, -1), inp["block_table"], 0.0, None, None, None, _cabi.PATH_AUTO, None)
    assert lib.b200attn_decode_attention_static_scratch(ctypes.byref(p), stream) == 5  # B200ATTN_E_UNSUPPORTED
    assert lib.b200attn_decode_attention_static_scratch(None, stream) == 7  # B200ATTN_E_NULL

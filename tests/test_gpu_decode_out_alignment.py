"""The split-KV decode kernel writes four outputs per store when the output rows are 8-byte aligned and falls back to
element stores otherwise (csrc/attn_decode.cu, run epilogue): both must give the same bits, and the fallback must stay
inside its rows."""
import pytest
import torch

from helpers import make_paged_inputs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("lens", [[700, 33, 4096, 1, 257], [5000] * 3])
def test_misaligned_output_rows_match_aligned_ones(dtype, lens):
    from vllm_triton_backend_b200 import _cabi, ops

    Hq, Hkv, D = 16, 4, 128
    inp = make_paged_inputs([(1, n) for n in lens], Hq, Hkv, D, 16, dtype=dtype, seed=11, device="cuda")
    B = len(lens)
    kw = dict(q=inp["q"], k=inp["k_cache"], v=inp["v_cache"], cu_seqlens_q=inp["cu_seqlens_q"], max_seqlen_q=1,
              seqused_k=inp["seq_lens"], max_seqlen_k=max(lens), softmax_scale=D ** -0.5, causal=True, window_size=(-1, -1),
              block_table=inp["block_table"], softcap=0.0, q_descale=None, k_descale=None, v_descale=None)
    aligned = torch.full((B, Hq, D), float("nan"), dtype=dtype, device="cuda")
    ops.unified_attention(out=aligned, **kw)
    assert _cabi.last_kernel() == "attn_decode_split_kv"
    # rows that start 2 bytes off an 8-byte boundary, with a head stride that is not a multiple of four elements
    wide = torch.full((B, Hq, D + 6), 123.0, dtype=dtype, device="cuda")
    out = wide[:, :, 1:D + 1]
    assert out.data_ptr() % 8 != 0 and out.stride(1) % 4 != 0 and out.stride(2) == 1
    ops.unified_attention(out=out, **kw)
    assert _cabi.last_kernel() == "attn_decode_split_kv"
    torch.cuda.synchronize()
    assert torch.equal(out, aligned)
    assert (wide[:, :, 0] == 123.0).all() and (wide[:, :, D + 1:] == 123.0).all()  # nothing outside the rows

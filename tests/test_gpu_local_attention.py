"""Chunked local attention (iRoPE / Llama-4).  The reference cuts the batch into
virtual sequences with vLLM's `make_local_attention_virtual_batches` and runs its unchanged kernel
on them (ibm-triton-lib/ibm_triton_lib/backend/triton_attn.py:155-190, 423-444); vLLM 0.22 does the
cutting in a wrapper around the backend's metadata builder.  Either way the kernels see many short
sequences with a freshly gathered block table: run exactly that through `B200AttentionImpl.forward`
and compare with the oracle's direct statement of the mask (`local_chunk=`)."""
import pytest
import torch

from helpers import BF16_ATOL, BF16_RTOL, FP8_ATOL, assert_close, make_paged_inputs
from oracle.paged_attention import unified_attention_ref

pytestmark = pytest.mark.gpu


def _virtual_batches(chunk, inp, bs, dev):
    try:
        from vllm.v1.attention.backend import CommonAttentionMetadata
        from vllm.v1.attention.backends.utils import make_local_attention_virtual_batches
    except Exception as e:  # pragma: no cover - depends on the installed vLLM
        pytest.skip(f"vLLM helpers not importable: {type(e).__name__}: {e}")
    cu, sl = inp["cu_seqlens_q"], inp["seq_lens"]
    T = int(cu[-1])
    common = CommonAttentionMetadata(
        query_start_loc=cu.to(dev), query_start_loc_cpu=cu, seq_lens=sl.to(dev), num_reqs=len(sl), num_actual_tokens=T,
        max_query_len=int((cu[1:] - cu[:-1]).max()), max_seq_len=int(sl.max()), block_table_tensor=inp["block_table"].to(dev),
        slot_mapping=torch.full((T,), -1, dtype=torch.int64, device=dev), _seq_lens_cpu=sl,
        _num_computed_tokens_cpu=sl - (cu[1:] - cu[:-1]),
    )
    local, _ = make_local_attention_virtual_batches(chunk, common, bs)
    return local


@pytest.mark.parametrize("chunk", [64, 256])
@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("lens", [
    [(1, 700), (1, 33), (1, 64), (1, 65), (1, 1)],                        # decode: one virtual sequence each
    [(300, 300), (1, 555), (64, 1064), (5, 40), (513, 2049), (200, 263)],  # prefill / extend / decode rows
], ids=["decode", "mixed"])
def test_virtual_batches_equal_the_chunk_mask(chunk, fp8, lens):
    from vllm_triton_backend_b200 import ops
    from vllm_triton_backend_b200.backend.attn import B200AttentionImpl, B200AttentionMetadata

    dev = "cuda"
    Hq, Hkv, D, bs = 16, 4, 128, 16
    ks, vs = (0.37, 0.61) if fp8 else (1.0, 1.0)
    inp = make_paged_inputs(lens, Hq, Hkv, D, bs, seed=23, kv_dtype=torch.float8_e4m3fn if fp8 else None, k_scale=ks, v_scale=vs)
    local = _virtual_batches(chunk, inp, bs, dev)
    B, T = local.seq_lens.shape[0], inp["q"].shape[0]
    assert B >= len(lens)
    ws = torch.zeros(ops.workspace_bytes(T, B, Hq, Hkv, D) + ops.fp8_staging_bytes(B, 4096, bs, Hkv, D), dtype=torch.uint8, device=dev)
    meta = B200AttentionMetadata(
        num_actual_tokens=T, max_query_len=int(local.max_query_len), query_start_loc=local.query_start_loc,
        max_seq_len=int(local.max_seq_len), seq_lens=local.seq_lens, block_table=local.block_table_tensor,
        slot_mapping=local.slot_mapping, workspace=ws)
    impl = B200AttentionImpl(Hq, D, D ** -0.5, Hkv, None, None, "fp8" if fp8 else "auto")

    class L:
        _k_scale = torch.tensor([ks], dtype=torch.float32, device=dev)
        _v_scale = torch.tensor([vs], dtype=torch.float32, device=dev)

    if fp8:
        kv = torch.stack([inp["k_cache"].view(torch.uint8), inp["v_cache"].view(torch.uint8)]).to(dev)
    else:
        kv = torch.stack([inp["k_cache"], inp["v_cache"]]).to(dev)
    out = torch.full((T, Hq * D), float("nan"), dtype=torch.bfloat16, device=dev)
    impl.forward(L, inp["q"].to(dev), None, None, kv, meta, output=out)
    torch.cuda.synchronize()
    ref = unified_attention_ref(inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"],
                                inp["block_table"], D ** -0.5, k_scale=ks, v_scale=vs, out_dtype=torch.float32,
                                local_chunk=chunk)
    assert_close(out.cpu().view(T, Hq, D), ref, atol=FP8_ATOL if fp8 else BF16_ATOL, rtol=BF16_RTOL,
                 what=f"chunked local attention, chunk {chunk}")

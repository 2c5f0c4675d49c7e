"""The oracle's chunked-local-attention mask (`local_chunk=`) against the reference's formulation:
vLLM's `make_local_attention_virtual_batches` (what ibm_triton_lib/backend/triton_attn.py:155-190
calls) cuts the batch into virtual sequences; the plain oracle on those must give the same numbers."""
import pytest
import torch

from helpers import make_paged_inputs
from oracle.paged_attention import unified_attention_ref


@pytest.mark.parametrize("chunk", [32, 64, 256])
def test_chunk_mask_equals_virtual_batches(chunk):
    try:
        from vllm.v1.attention.backend import CommonAttentionMetadata
        from vllm.v1.attention.backends.utils import make_local_attention_virtual_batches
    except Exception as e:  # pragma: no cover - depends on the installed vLLM
        pytest.skip(f"vLLM helpers not importable: {type(e).__name__}: {e}")
    lens = [(300, 300), (1, 555), (64, 1064), (5, 40), (200, 263), (1, 64), (1, 65)]
    Hq, Hkv, D, bs = 4, 2, 32, 16
    inp = make_paged_inputs(lens, Hq, Hkv, D, bs, seed=1, dtype=torch.float32)
    cu, sl = inp["cu_seqlens_q"], inp["seq_lens"]
    T = int(cu[-1])
    common = CommonAttentionMetadata(
        query_start_loc=cu, query_start_loc_cpu=cu, seq_lens=sl, num_reqs=len(sl), num_actual_tokens=T,
        max_query_len=int((cu[1:] - cu[:-1]).max()), max_seq_len=int(sl.max()), block_table_tensor=inp["block_table"],
        slot_mapping=torch.zeros(T, dtype=torch.int64), _seq_lens_cpu=sl, _num_computed_tokens_cpu=sl - (cu[1:] - cu[:-1]))
    local, _ = make_local_attention_virtual_batches(chunk, common, bs)
    assert local.seq_lens.shape[0] > len(lens)
    args = (inp["q"], inp["k_cache"], inp["v_cache"])
    virt = unified_attention_ref(*args, local.query_start_loc_cpu, local.seq_lens, local.block_table_tensor, D ** -0.5,
                                 out_dtype=torch.float32)
    mask = unified_attention_ref(*args, cu, sl, inp["block_table"], D ** -0.5, out_dtype=torch.float32, local_chunk=chunk)
    assert (virt - mask).abs().max().item() < 1e-6

"""Host-side logic of the per-(token, head) quantised cache (no GPU): the plugin's block layout and dtype plumbing."""
import pytest
import torch

from vllm_triton_backend_b200 import ops


@pytest.mark.parametrize("D,Hkv,bs,NB", [(128, 8, 16, 5), (64, 2, 32, 3), (256, 1, 16, 4), (128, 3, 16, 2)])
@pytest.mark.parametrize("kv_dtype", [torch.float8_e4m3fn, torch.int8])
def test_block_layout_tiles_the_allocation_exactly(D, Hkv, bs, NB, kv_dtype):
    """Data [bs][Hkv][D] first, then the block's scales [Hkv][bs] fp32: together they are the head_size + 4 bytes per row
    that vLLM budgets (AttentionSpec.page_size_bytes), nothing overlaps and nothing is left over."""
    raw = torch.zeros(2, NB, bs, Hkv, D + 4, dtype=torch.uint8)
    data, scales = ops.per_token_head_cache_views(raw[1], D, kv_dtype)
    assert data.shape == (NB, bs, Hkv, D) and data.dtype == kv_dtype and data.stride() == (bs * Hkv * (D + 4), Hkv * D, D, 1)
    assert scales.shape == (NB, bs, Hkv) and scales.dtype == torch.float32
    assert scales.stride() == (bs * Hkv * (D + 4) // 4, 1, bs)  # a head's bs scales are contiguous: one 64-byte line per 16 tokens
    data.view(torch.uint8).fill_(1)
    scales.view(torch.int32).fill_(0x02020202)
    plane = raw[1].reshape(NB, -1)
    assert int((plane == 1).sum()) == NB * bs * Hkv * D           # the data region
    assert int((plane == 2).sum()) == NB * bs * Hkv * 4           # the scale region
    assert int((plane == 0).sum()) == 0                            # and nothing else in the block
    assert int(raw[0].sum()) == 0                                  # the other plane is untouched
    # within a block: data first, scales behind it
    assert (plane[:, : bs * Hkv * D] == 1).all() and (plane[:, bs * Hkv * D:] == 2).all()


def test_block_layout_rejects_what_it_cannot_address():
    with pytest.raises(ValueError, match="head_size \\+ 4"):
        ops.per_token_head_cache_views(torch.zeros(2, 16, 2, 128, dtype=torch.uint8), 128, torch.int8)
    with pytest.raises(ValueError, match="contiguous"):
        ops.per_token_head_cache_views(torch.zeros(2, 2, 16, 132, dtype=torch.uint8).transpose(1, 2), 128, torch.int8)  # HND
    with pytest.raises(ValueError, match="one-byte|head_size"):
        ops.per_token_head_cache_views(torch.zeros(2, 16, 2, 132, dtype=torch.float16), 128, torch.int8)


def test_backend_declares_the_padded_shape_and_dtypes():
    pytest.importorskip("vllm")
    from vllm_triton_backend_b200.backend.attn import PER_TOKEN_HEAD_DTYPES, B200AttentionBackend, B200AttentionImpl

    for name in ("fp8_per_token_head", "int8_per_token_head"):
        assert name in B200AttentionBackend.supported_kv_cache_dtypes and name in PER_TOKEN_HEAD_DTYPES
        assert B200AttentionBackend.get_kv_cache_shape(7, 16, 4, 128, name) == (2, 7, 16, 4, 132)
        impl = B200AttentionImpl(8, 128, 0.1, 4, None, None, name)
        assert impl._per_token_head and not impl.fused_rope_kvcache_supported()
    assert B200AttentionBackend.get_kv_cache_shape(7, 16, 4, 128, "fp8") == (2, 7, 16, 4, 128)
    with pytest.raises(NotImplementedError):
        B200AttentionImpl(8, 128, 0.1, 4, None, None, "fp8_e5m2")


def test_operators_refuse_cpu_tensors():
    k = torch.zeros(4, 2, 128)
    kc = torch.zeros(2, 16, 2, 128, dtype=torch.int8)
    sc = torch.ones(2, 16, 2)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.reshape_and_cache_flash_per_token_head(k, k, kc, kc, sc, sc, torch.zeros(4, dtype=torch.int64))

"""The vLLM plugin surface mirrors the reference plugin's (backend/triton_attn.py,
backend/platform.py) against the vLLM installed in this image.  No GPU needed."""
import pytest
import torch

vllm = pytest.importorskip("vllm")

from vllm.v1.attention.backend import AttentionBackend, AttentionImpl, AttentionMetadataBuilder, AttentionType  # noqa: E402

from vllm_triton_backend_b200.backend import register, register_backend  # noqa: E402
from vllm_triton_backend_b200.backend.attn import (  # noqa: E402
    BACKEND_PATH, B200AttentionBackend, B200AttentionImpl, B200AttentionMetadataBuilder, split_kv_cache,
)


def test_entry_points_resolve():
    import importlib.util

    mod, cls = register().rsplit(".", 1)
    # vllm.platforms.cuda needs libcuda, so without a GPU we only resolve the module's file
    spec = importlib.util.find_spec(mod)
    assert spec is not None and f"class {cls}(" in open(spec.origin).read()
    register_backend()
    from vllm.v1.attention.backends.registry import AttentionBackendEnum

    assert AttentionBackendEnum.CUSTOM.get_path() == BACKEND_PATH
    assert AttentionBackendEnum.CUSTOM.get_class() is B200AttentionBackend
    # Attention.__init__ maps the backend back to its enum member by name (attention.py:386)
    assert AttentionBackendEnum[B200AttentionBackend.get_name()] is AttentionBackendEnum.CUSTOM


def test_backend_class_surface():
    B = B200AttentionBackend
    assert issubclass(B, AttentionBackend) and issubclass(B.get_impl_cls(), AttentionImpl)
    assert issubclass(B.get_builder_cls(), AttentionMetadataBuilder) and B.get_builder_cls() is B200AttentionMetadataBuilder
    assert B.get_kv_cache_shape(10, 16, 8, 128) == (2, 10, 16, 8, 128)
    assert B.get_kv_cache_shape(10, 32, 8, 128, cache_dtype_str="fp8") == (2, 10, 32, 8, 128)
    with pytest.raises(ValueError, match="multiple of 16"):
        B.get_kv_cache_shape(10, 8, 8, 128)
    assert B.use_cascade_attention() is False
    assert B.forward_includes_kv_cache_update is False
    assert {"auto", "bfloat16", "fp8", "fp8_e4m3"} <= set(B.supported_kv_cache_dtypes)
    assert B.get_kv_cache_block_dim(16, 8, 128) == 1


def test_validate_configuration_accepts_the_target_and_rejects_the_rest():
    from vllm.platforms.interface import DeviceCapability

    base = dict(head_size=128, dtype=torch.bfloat16, kv_cache_dtype="auto", block_size=16, use_mla=False, has_sink=False,
                use_sparse=False, use_mm_prefix=False, use_per_head_quant_scales=False, attn_type="decoder",
                use_non_causal=False, use_batch_invariant=False)
    ok = B200AttentionBackend.validate_configuration(device_capability=DeviceCapability(10, 0), **base)
    assert ok == []
    assert B200AttentionBackend.validate_configuration(device_capability=DeviceCapability(9, 0), **base) != []
    assert B200AttentionBackend.validate_configuration(device_capability=DeviceCapability(10, 0), **{**base, "use_mla": True}) != []
    assert B200AttentionBackend.validate_configuration(device_capability=DeviceCapability(10, 0), **{**base, "attn_type": "encoder"}) != []


def test_impl_init_mirrors_reference_argument_handling():
    impl = B200AttentionImpl(32, 128, 0.088, 8, None, 4096, "auto", logits_soft_cap=None)
    assert impl.sliding_window == (4095, 0) and impl.logits_soft_cap == 0 and impl.num_queries_per_kv == 4
    assert B200AttentionImpl(32, 128, 0.088, 8, None, None, "auto").sliding_window == (-1, -1)
    assert impl.supports_quant_query_input is False
    with pytest.raises(ValueError, match="block-sparse"):
        B200AttentionImpl(32, 128, 1.0, 8, None, None, "auto", blocksparse_params={})
    with pytest.raises(ValueError, match="Head size 100"):
        B200AttentionImpl(32, 100, 1.0, 8, None, None, "auto")
    with pytest.raises(NotImplementedError, match="Encoder"):
        B200AttentionImpl(32, 128, 1.0, 8, None, None, "auto", attn_type=AttentionType.ENCODER)


def test_profiling_run_and_fused_output_quant():
    impl = B200AttentionImpl(4, 64, 1.0, 4, None, None, "auto")
    out = torch.ones(3, 4 * 64)
    assert impl.forward(None, None, None, None, None, None, output=out).abs().sum() == 0  # attn_metadata is None
    with pytest.raises(NotImplementedError, match="block_scale output quantization"):
        impl.forward(None, None, None, None, None, object(), output=out, output_block_scale=torch.ones(1))
    # static per-tensor FP8 of the output is fused into the kernels (vLLM 0.22 triton_attn.py:453-454)
    from vllm.model_executor.layers.quantization.utils.quant_utils import kFp8DynamicTokenSym, kFp8StaticTensorSym
    assert impl.fused_output_quant_supported(kFp8StaticTensorSym)
    assert not impl.fused_output_quant_supported(kFp8DynamicTokenSym)
    with pytest.raises(AssertionError, match="Output tensor"):
        impl.forward(None, None, None, None, None, object(), output=None)


def test_split_kv_cache_views_for_both_allocations():
    ours = torch.zeros(2, 6, 16, 8, 128)
    k, v = split_kv_cache(ours)
    assert k.shape == (6, 16, 8, 128) and k.stride(0) == 16 * 8 * 128
    stock = torch.zeros(6, 2, 16, 8, 128)
    k, v = split_kv_cache(stock)
    assert k.shape == (6, 16, 8, 128) and k.stride(0) == 2 * 16 * 8 * 128 and v.data_ptr() != k.data_ptr()


@pytest.mark.gpu
def test_platform_selects_our_backend_unconditionally():
    from vllm_triton_backend_b200.backend.platform import B200Platform

    assert B200Platform.get_attn_backend_cls(None, None, 32) == BACKEND_PATH


def test_kv_cache_stride_order_follows_the_configured_layout():
    from vllm.v1.attention.backends import utils as vutils

    try:
        vutils.set_kv_cache_layout("NHD")
        assert B200AttentionBackend.get_kv_cache_stride_order() == (0, 1, 2, 3, 4)
        vutils.set_kv_cache_layout("HND")
        order = B200AttentionBackend.get_kv_cache_stride_order()
        assert order == (0, 1, 3, 2, 4)
        assert B200AttentionBackend.get_kv_cache_stride_order(include_num_layers_dimension=True) == (2, 4, 0, 1, 3, 5)
    finally:
        vutils.set_kv_cache_layout(None)
    # what the model runner does with it (gpu_model_runner.py:6935-6984): allocate permuted, view back
    shape = B200AttentionBackend.get_kv_cache_shape(6, 16, 4, 64)
    phys = torch.zeros(tuple(shape[i] for i in order))
    inv = [order.index(i) for i in range(len(order))]
    kv = phys.permute(*inv)
    assert tuple(kv.shape) == shape
    k, v = split_kv_cache(kv)
    assert k.shape == (6, 16, 4, 64) and k.stride(3) == 1 and k.stride(1) == 64 and k.stride(2) == 16 * 64

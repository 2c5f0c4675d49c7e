"""e4m3 quantisation oracle: torch and plain C restatements agree byte for byte
(reference arithmetic: quant_utils.cuh:479-502, __NV_SATFINITE / __NV_E4M3)."""
import pytest
import torch

from oracle.cache_write import load_c_lib, quantize_e4m3_c, quantize_e4m3_torch


def _all_bf16():
    bits = torch.arange(0, 65536, dtype=torch.int32).to(torch.int16)
    return bits.view(torch.bfloat16).float()


@pytest.mark.skipif(load_c_lib() is None, reason="oracle C library not built (make -C oracle)")
@pytest.mark.parametrize("scale", [1.0, 0.5, 0.37, 3.3, 1 / 64])
def test_c_and_torch_e4m3_agree_on_every_bf16(scale):
    x = _all_bf16()
    assert torch.equal(quantize_e4m3_torch(x, scale), quantize_e4m3_c(x, scale))


def test_e4m3_known_answers():
    x = torch.tensor([0.0, 1.0, -1.0, 448.0, 1e9, -1e9, 2.0 ** -9, 2.0 ** -10, 0.0625 + 0.00390625, float("nan")])
    got = quantize_e4m3_torch(x, 1.0).tolist()
    #            0     1     -1    448   sat   -sat  min-sub  tie->0 (RNE)  tie->even   nan
    assert got == [0x00, 0x38, 0xB8, 0x7E, 0x7E, 0xFE, 0x01, 0x00, 0x18, 0x7F]


@pytest.mark.skipif(load_c_lib() is None, reason="oracle C library not built (make -C oracle)")
@pytest.mark.parametrize("scale", [1.0, 0.37, 2.5])
def test_scaled_fp8_quant_oracle_is_reciprocal_multiply_then_the_same_conversion(scale):
    """K4's oracle (vllm .../fp8/common.cuh:40-60): x * fl(1/scale), fmin/fmax clamp, then the e4m3 conversion that the
    plain-C restatement pins; NaN goes to +448 (fminf returns the number), unlike the cache write's NaN -> 0x7f."""
    from oracle.cache_write import scaled_fp8_quant_ref

    x = _all_bf16()
    inv = (torch.tensor(1.0) / torch.tensor(scale, dtype=torch.float32))
    y = x * inv
    finite = ~torch.isnan(y)
    exp = quantize_e4m3_c(torch.where(finite, y, torch.full_like(y, 448.0)), 1.0)
    assert torch.equal(scaled_fp8_quant_ref(x, scale), exp)
    assert scaled_fp8_quant_ref(torch.tensor([float("nan"), float("inf"), float("-inf")]), scale).tolist() == [0x7E, 0x7E, 0xFE]

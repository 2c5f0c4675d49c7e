"""K4, `b200attn_scaled_fp8_quant`: byte-exact against the oracle's restatement of vLLM's static per-tensor FP8
quantisation (reciprocal multiply, vllm/csrc/quantization/w8a8/fp8/common.cu:11-29, common.cuh:40-60) - the call the
reference makes on the query at ibm-triton-lib/ibm_triton_lib/backend/triton_attn.py:417."""
import pytest
import torch

from oracle.cache_write import quantize_e4m3_torch, scaled_fp8_quant_ref

pytestmark = pytest.mark.gpu


def _inputs(dtype, rows, n, seed, spread=600.0):
    g = torch.Generator().manual_seed(seed)
    x = (torch.rand(rows, n, generator=g) * 2 - 1) * spread
    # every magnitude class: values that saturate, subnormal e4m3 results, exact ties, zeros of both signs
    x[0, :8] = torch.tensor([0.0, -0.0, 448.0, -448.0, 1e6, -1e6, 2 ** -9, -(2 ** -10)])
    x[1 % rows, :6] = torch.tensor([0.0009765625, 0.0029296875, 17.0, 18.0, 19.0, 464.0])
    return x.to(dtype)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("scale", [1.0, 0.37, 2.5, 1 / 64])
@pytest.mark.parametrize("rows,n", [(7, 4096), (64, 128), (3, 100)])
def test_bytes_match_the_oracle(dtype, scale, rows, n):
    from vllm_triton_backend_b200 import ops

    spread = 600.0 if dtype != torch.float16 else 300.0
    x = _inputs(dtype, rows, n, seed=rows + n, spread=spread)
    s = torch.tensor([scale], dtype=torch.float32, device="cuda")
    out, s_out = ops.scaled_fp8_quant(x.cuda(), s)
    torch.cuda.synchronize()
    assert out.dtype == torch.float8_e4m3fn and s_out is s
    exp = scaled_fp8_quant_ref(x, scale)
    assert torch.equal(out.cpu().view(torch.uint8), exp)


def test_strided_rows_and_nan():
    from vllm_triton_backend_b200 import ops

    big = _inputs(torch.bfloat16, 16, 3 * 256, seed=3)
    big[2, 300] = float("nan")
    big[5, 256] = float("inf")
    big[5, 257] = float("-inf")
    x = big[:, 256:512]  # a column slice: row stride 768, 256 elements per row
    s = torch.tensor([0.5], dtype=torch.float32, device="cuda")
    out, _ = ops.scaled_fp8_quant(big.cuda()[:, 256:512], s)
    torch.cuda.synchronize()
    exp = scaled_fp8_quant_ref(x, 0.5)
    got = out.cpu().view(torch.uint8)
    assert torch.equal(got, exp)
    assert got[2, 44] == 0x7E and got[5, 0] == 0x7E and got[5, 1] == 0xFE  # NaN and +inf -> +448, -inf -> -448


def _rounding_boundary_inputs(scale):
    """fp32 inputs within a few ulps of `midpoint * scale` for every midpoint between two adjacent e4m3 values: where
    x * fl(1/scale) and x / scale fall on different sides of a rounding boundary."""
    import numpy as np

    vals = torch.arange(0, 127, dtype=torch.uint8).view(torch.float8_e4m3fn).float()  # the positive finite e4m3 values
    x0 = (((vals[:-1] + vals[1:]) / 2).double() * scale).float().numpy()
    xs, up, dn = [x0], x0.copy(), x0.copy()
    for _ in range(4):
        up, dn = np.nextafter(up, np.float32(np.inf)), np.nextafter(dn, np.float32(-np.inf))
        xs += [up.copy(), dn.copy()]
    x = torch.from_numpy(np.concatenate(xs))
    return torch.cat([x, -x]).view(2, -1)


@pytest.mark.parametrize("scale", [0.37, 0.21, 3.3])
def test_reciprocal_multiply_is_not_the_cache_writes_division(scale):
    """K4 multiplies by fl(1/scale), the cache write divides: next to a rounding boundary the two land on different
    bytes.  Both conventions are pinned (this test fails if K4 were to divide)."""
    from vllm_triton_backend_b200 import ops

    x = _rounding_boundary_inputs(scale)
    mul, div = scaled_fp8_quant_ref(x, scale), quantize_e4m3_torch(x, scale)
    assert int((mul != div).sum()) >= 16
    s = torch.tensor([scale], dtype=torch.float32, device="cuda")
    out, _ = ops.scaled_fp8_quant(x.cuda(), s)
    assert torch.equal(out.cpu().view(torch.uint8), mul)


def test_matches_vllm_compiled_op_when_available():
    from vllm_triton_backend_b200 import ops

    try:
        from vllm import _custom_ops as vops
        import vllm._C  # noqa: F401
    except Exception as e:  # pragma: no cover
        pytest.skip(f"vllm._C not loadable: {e}")
    x = _inputs(torch.bfloat16, 33, 1024, seed=5).cuda()
    s = torch.tensor([0.21], dtype=torch.float32, device="cuda")
    theirs, _ = vops.scaled_fp8_quant(x, s)
    ours, _ = ops.scaled_fp8_quant(x, s)
    assert torch.equal(ours.view(torch.uint8), theirs.view(torch.uint8))

"""K1 parity on the GPU: bit-exact against the CPU oracle, the reference golden
vector and (when it loads) vLLM's own reshape_and_cache_flash.  Calls go through
the C ABI (ops.reshape_and_cache_flash -> libb200attn.so)."""
import numpy as np
import pytest
import torch

from helpers import GOLDEN_DIR
from oracle.cache_write import reshape_and_cache_flash_ref

pytestmark = pytest.mark.gpu


def _ops():
    from vllm_triton_backend_b200 import ops

    return ops


def test_golden_reference_vector_bit_exact():
    z = np.load(f"{GOLDEN_DIR}/cache_write_bf16.npz")
    key = torch.from_numpy(z["key"]).view(torch.bfloat16).cuda()
    value = torch.from_numpy(z["value"]).view(torch.bfloat16).cuda()
    slots = torch.from_numpy(z["slot_mapping"]).cuda()
    kc = torch.zeros(z["k_cache"].shape, dtype=torch.bfloat16, device="cuda")
    vc = torch.zeros_like(kc)
    _ops().reshape_and_cache_flash(key, value, kc, vc, slots, "auto", None, None)
    torch.cuda.synchronize()
    assert np.array_equal(kc.view(torch.int16).cpu().numpy(), z["k_cache"])
    assert np.array_equal(vc.view(torch.int16).cpu().numpy(), z["v_cache"])


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("head_size", [64, 80, 128, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
@pytest.mark.parametrize("layout", ["planes", "interleaved", "hnd"])
def test_copy_bit_exact(dtype, head_size, block_size, layout):
    torch.manual_seed(0)
    T, H, NB = 42, 8, 12
    # key/value are slices of a fused QKV output: row stride > H*D
    qkv = torch.randn(T + 3, 3 * H * head_size, device="cuda").to(dtype)
    key = qkv[:, H * head_size : 2 * H * head_size].view(-1, H, head_size)
    value = qkv[:, 2 * H * head_size :].view(-1, H, head_size)
    slots = torch.randperm(NB * block_size, device="cuda")[:T].to(torch.int64)
    slots[5] = -1
    slots[17] = -1
    if layout == "planes":
        kv = torch.full((2, NB, block_size, H, head_size), 3.0, device="cuda").to(dtype)
        kc, vc = kv.unbind(0)
    elif layout == "interleaved":
        kv = torch.full((NB, 2, block_size, H, head_size), 3.0, device="cuda").to(dtype)
        kc, vc = kv.unbind(1)
    else:  # HND physical layout viewed as NHD
        kv = torch.full((2, NB, H, block_size, head_size), 3.0, device="cuda").to(dtype)
        kc, vc = (x.permute(0, 2, 1, 3) for x in kv.unbind(0))
    kc_ref, vc_ref = kc.cpu().clone(), vc.cpu().clone()
    reshape_and_cache_flash_ref(key.cpu(), value.cpu(), kc_ref, vc_ref, slots.cpu())
    _ops().reshape_and_cache_flash(key, value, kc, vc, slots, "auto", None, None)
    torch.cuda.synchronize()
    assert torch.equal(kc.cpu(), kc_ref) and torch.equal(vc.cpu(), vc_ref)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("head_size", [64, 80, 128])
@pytest.mark.parametrize("scales", [(1.0, 1.0), (0.5, 0.37), (1 / 64, 3.3)])
def test_fp8_quantise_byte_exact(dtype, head_size, scales):
    torch.manual_seed(1)
    T, H, NB, bs = 37, 8, 6, 16
    key = (torch.randn(T, H, head_size, device="cuda") * 3).to(dtype)
    value = (torch.randn(T, H, head_size, device="cuda") * 40).to(dtype)
    key[0, 0, :4] = torch.tensor([1e4, -1e4, 0.0, 2.0 ** -10], dtype=dtype)  # saturation, zero, tiny
    slots = torch.randperm(NB * bs, device="cuda")[:T].to(torch.int64)
    slots[3] = -1
    kv = torch.zeros(NB, 2, bs, H, head_size, dtype=torch.uint8, device="cuda")
    kc, vc = kv.unbind(1)
    ks = torch.tensor(scales[0], dtype=torch.float32, device="cuda")
    vs = torch.tensor(scales[1], dtype=torch.float32, device="cuda")
    kc_ref, vc_ref = kc.cpu().clone(), vc.cpu().clone()
    reshape_and_cache_flash_ref(key.cpu(), value.cpu(), kc_ref, vc_ref, slots.cpu(), "fp8", scales[0], scales[1])
    _ops().reshape_and_cache_flash(key, value, kc, vc, slots, "fp8", ks, vs)
    torch.cuda.synchronize()
    assert torch.equal(kc.cpu(), kc_ref) and torch.equal(vc.cpu(), vc_ref)


def test_token_count_comes_from_slot_mapping():
    key = torch.randn(10, 2, 64, device="cuda").bfloat16()
    value = torch.randn(10, 2, 64, device="cuda").bfloat16()
    kc = torch.zeros(4, 16, 2, 64, device="cuda").bfloat16()
    vc = torch.zeros_like(kc)
    slots = torch.tensor([0, 1, 2], dtype=torch.int64, device="cuda")  # 3 < 10 (CUDA-graph padding)
    _ops().reshape_and_cache_flash(key, value, kc, vc, slots, "auto", None, None)
    torch.cuda.synchronize()
    assert torch.equal(kc[0, :3], key[:3]) and (kc[0, 3:] == 0).all() and (kc[1:] == 0).all()


def test_matches_vllm_compiled_op_when_available():
    try:
        import vllm._C  # noqa: F401
        op = torch.ops._C_cache_ops.reshape_and_cache_flash
    except Exception as e:  # pragma: no cover
        pytest.skip(f"vllm._C not loadable: {e}")
    torch.manual_seed(2)
    T, H, D, NB, bs = 129, 8, 128, 16, 16
    key = torch.randn(T, H, D, device="cuda").bfloat16()
    value = torch.randn(T, H, D, device="cuda").bfloat16()
    slots = torch.randperm(NB * bs, device="cuda")[:T].to(torch.int64)
    one = torch.ones((), dtype=torch.float32, device="cuda")
    for kvd, cdt, ks, vs in [("auto", torch.bfloat16, one, one), ("fp8", torch.uint8, one * 0.37, one * 0.5)]:
        a = torch.zeros(2, NB, bs, H, D, dtype=cdt, device="cuda")
        b = torch.zeros_like(a)
        op(key, value, a[0], a[1], slots, kvd, ks, vs)
        _ops().reshape_and_cache_flash(key, value, b[0], b[1], slots, kvd, ks, vs)
        torch.cuda.synchronize()
        assert torch.equal(a.view(torch.uint8), b.view(torch.uint8)), kvd


def test_cpu_tensors_are_rejected_not_emulated():
    key = torch.randn(2, 2, 64).bfloat16()
    kc = torch.zeros(1, 16, 2, 64).bfloat16()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _ops().reshape_and_cache_flash(key, key, kc, kc, torch.zeros(2, dtype=torch.int64), "auto", None, None)

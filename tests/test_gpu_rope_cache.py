"""K5 (fused rotary embedding + KV-cache write) on the GPU: bit-exact against the oracle
(oracle/rope.py + oracle/cache_write.py) and, where the installed vLLM's compiled ops load,
against the unfused pair it replaces (`_C.rotary_embedding` + `_C_cache_ops.reshape_and_cache_flash`)."""
import pytest
import torch

from oracle.cache_write import reshape_and_cache_flash_ref
from oracle.rope import rotary_embedding_ref

pytestmark = pytest.mark.gpu


def _inputs(T, Hq, Hkv, D, rot, bs, dtype, fp8, seed=0, pad=2):
    g = torch.Generator().manual_seed(seed)
    qkv = ((torch.rand(T, (Hq + 2 * Hkv) * D, generator=g) * 2 - 1) * 3).to(dtype)  # q/k/v are views, as in vLLM
    positions = torch.randint(0, 4096, (T,), generator=g, dtype=torch.int64)
    inv = 1.0 / (10000.0 ** (torch.arange(0, rot, 2, dtype=torch.float32) / rot))
    f = torch.arange(4096, dtype=torch.float32)[:, None] * inv[None]
    cos_sin = torch.cat([f.cos(), f.sin()], dim=-1).to(dtype)
    NB = (T + bs - 1) // bs + 5
    slots = torch.randperm(NB * bs, generator=g)[:T].to(torch.int64)
    slots[-pad:] = -1  # padded tokens: rotated, not cached
    cdt = torch.uint8 if fp8 else dtype
    kv = (torch.rand(2, NB, bs, Hkv, D, generator=g) * 100).to(cdt)
    return qkv, positions, cos_sin, slots, kv


def _split(qkv, Hq, Hkv, D):
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    return q.view(-1, Hq, D), k.view(-1, Hkv, D), v.view(-1, Hkv, D)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("Hq,Hkv,D,rot", [(32, 8, 128, 128), (8, 2, 64, 32), (4, 4, 256, 256)])
def test_fused_rope_cache_is_bit_exact(dtype, is_neox, fp8, Hq, Hkv, D, rot):
    from vllm_triton_backend_b200 import _cabi, ops

    T, bs = 37, 16
    ks, vs = (0.37, 0.61) if fp8 else (1.0, 1.0)
    qkv, positions, cos_sin, slots, kv = _inputs(T, Hq, Hkv, D, rot, bs, dtype, fp8, seed=Hq + D)
    # oracle
    q, k, v = _split(qkv, Hq, Hkv, D)
    q_ref, k_ref = rotary_embedding_ref(positions, q, k, cos_sin, is_neox)
    kc_ref, vc_ref = kv[0].clone(), kv[1].clone()
    reshape_and_cache_flash_ref(k_ref, v, kc_ref, vc_ref, slots, kv_cache_dtype="fp8" if fp8 else "auto", k_scale=ks, v_scale=vs)
    # device
    dev = "cuda"
    d_qkv, d_kv = qkv.to(dev), kv.to(dev)
    dq, dk, dv = _split(d_qkv, Hq, Hkv, D)
    ops.rope_and_cache_flash(dq, dk, dv, positions.to(dev), cos_sin.to(dev), is_neox, d_kv[0], d_kv[1], slots.to(dev),
                             "fp8" if fp8 else "auto", ks, vs)
    torch.cuda.synchronize()
    assert _cabi.last_kernel() == "rope_cache_write"
    assert torch.equal(dq.cpu(), q_ref), "rotated q"
    assert torch.equal(dk.cpu(), k_ref), "rotated k"
    assert torch.equal(dv.cpu(), v), "v must be untouched"
    assert torch.equal(d_kv[0].cpu(), kc_ref), "key cache"
    assert torch.equal(d_kv[1].cpu(), vc_ref), "value cache"


@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("fp8", [False, True])
def test_matches_vllm_unfused_ops_when_available(is_neox, fp8):
    try:
        import vllm._C  # noqa: F401
        from vllm import _custom_ops as vops

        rotary, cache_op = vops.rotary_embedding, torch.ops._C_cache_ops.reshape_and_cache_flash
    except Exception as e:  # pragma: no cover - depends on the image
        pytest.skip(f"vllm._C not loadable: {type(e).__name__}: {e}")
    from vllm_triton_backend_b200 import ops

    Hq, Hkv, D, rot, T, bs = 32, 8, 128, 128, 150, 16
    dtype, dev = torch.bfloat16, "cuda"
    qkv, positions, cos_sin, slots, kv = _inputs(T, Hq, Hkv, D, rot, bs, dtype, fp8, seed=3)
    sc = torch.tensor([0.37], dtype=torch.float32, device=dev), torch.tensor([0.61], dtype=torch.float32, device=dev)
    # unfused reference pair
    r_qkv, r_kv = qkv.to(dev), kv.to(dev)
    rq, rk, rv = _split(r_qkv, Hq, Hkv, D)
    rotary(positions.to(dev), rq, rk, D, cos_sin.to(dev), is_neox)
    cache_op(rk, rv, r_kv[0], r_kv[1], slots.to(dev), "fp8" if fp8 else "auto", sc[0], sc[1])
    # fused
    d_qkv, d_kv = qkv.to(dev), kv.to(dev)
    dq, dk, dv = _split(d_qkv, Hq, Hkv, D)
    ops.rope_and_cache_flash(dq, dk, dv, positions.to(dev), cos_sin.to(dev), is_neox, d_kv[0], d_kv[1], slots.to(dev),
                             "fp8" if fp8 else "auto", sc[0], sc[1])
    torch.cuda.synchronize()
    assert torch.equal(d_qkv, r_qkv), "q/k after rope differ from vLLM's rotary_embedding"
    assert torch.equal(d_kv, r_kv), "cache differs from vLLM's reshape_and_cache_flash"


def test_impl_method_and_error_paths():
    from vllm_triton_backend_b200 import _cabi, ops
    from vllm_triton_backend_b200.backend.attn import B200AttentionImpl

    dev = "cuda"
    Hq, Hkv, D, T, bs = 8, 2, 128, 9, 16
    qkv, positions, cos_sin, slots, kv = _inputs(T, Hq, Hkv, D, D, bs, torch.bfloat16, False, seed=1)
    impl = B200AttentionImpl(Hq, D, D ** -0.5, Hkv, None, None, "auto")
    assert impl.fused_rope_kvcache_supported()

    class L:
        _k_scale = torch.ones(1, device=dev)
        _v_scale = torch.ones(1, device=dev)

    q, k, v = _split(qkv, Hq, Hkv, D)
    q_ref, k_ref = rotary_embedding_ref(positions, q, k, cos_sin, True)
    d_qkv, d_kv = qkv.to(dev), kv.to(dev)
    dq, dk, dv = _split(d_qkv, Hq, Hkv, D)
    impl.do_rope_and_kv_cache_update(L, dq, dk, dv, positions.to(dev), cos_sin.to(dev), True, d_kv, slots.to(dev))
    torch.cuda.synchronize()
    assert torch.equal(dq.cpu(), q_ref) and torch.equal(dk.cpu(), k_ref)
    with pytest.raises(ValueError, match="int64"):
        ops.rope_and_cache_flash(dq, dk, dv, positions.to(dev).int(), cos_sin.to(dev), True, d_kv[0], d_kv[1],
                                 slots.to(dev), "auto", 1.0, 1.0)
    with pytest.raises(_cabi.B200AttnError, match="ALIGNMENT"):  # rot_dim/2 = 4 is not a 16-byte vector
        ops.rope_and_cache_flash(dq, dk, dv, positions.to(dev), cos_sin.to(dev)[:, :8].contiguous(), True, d_kv[0],
                                 d_kv[1], slots.to(dev), "auto", 1.0, 1.0)

"""oracle/rope.py: the rotation itself (against exact float64 math) and its rounding model."""
import math

import pytest
import torch

from oracle.rope import rotary_embedding_ref


def _cache(max_pos, rot, base=10000.0, dtype=torch.float32):
    inv = 1.0 / (base ** (torch.arange(0, rot, 2, dtype=torch.float64) / rot))
    f = torch.arange(max_pos, dtype=torch.float64)[:, None] * inv[None]
    return torch.cat([f.cos(), f.sin()], dim=-1).to(dtype)


@pytest.mark.parametrize("is_neox", [True, False])
@pytest.mark.parametrize("rot,D", [(64, 64), (32, 64)])
def test_rotation_matches_float64_and_preserves_norm(is_neox, rot, D):
    g = torch.Generator().manual_seed(0)
    T, Hq, Hkv = 7, 4, 2
    q = torch.rand(T, Hq, D, generator=g, dtype=torch.float64) * 2 - 1
    k = torch.rand(T, Hkv, D, generator=g, dtype=torch.float64) * 2 - 1
    pos = torch.tensor([0, 1, 5, 100, 4095, 17, 2], dtype=torch.int64)
    cs = _cache(4096, rot, dtype=torch.float64)
    q2, k2 = rotary_embedding_ref(pos, q, k, cs, is_neox)
    # exact complex rotation per pair
    half = rot // 2
    for x, y in ((q, q2), (k, k2)):
        a, b = (x[..., :half], x[..., half:rot]) if is_neox else (x[..., 0:rot:2], x[..., 1:rot:2])
        a2, b2 = (y[..., :half], y[..., half:rot]) if is_neox else (y[..., 0:rot:2], y[..., 1:rot:2])
        ang = torch.atan2(cs[pos][:, half:], cs[pos][:, :half])[:, None, :]
        z = torch.complex(a, b) * torch.polar(torch.ones_like(ang), ang)
        assert torch.allclose(a2, z.real, atol=1e-12) and torch.allclose(b2, z.imag, atol=1e-12)
        assert torch.equal(y[..., rot:], x[..., rot:])  # pass-through dims untouched
        assert torch.allclose(y.norm(dim=-1), x.norm(dim=-1), atol=1e-12)
    assert torch.equal(q2[0], q[0]) and torch.equal(k2[0], k[0])  # position 0 is the identity


def test_bf16_is_fp32_arithmetic_rounded_once():
    x = torch.tensor([[[0.3379, 0.7773]]], dtype=torch.bfloat16)
    cs = torch.tensor([[math.cos(1.0), math.sin(1.0)]], dtype=torch.bfloat16)
    out, _ = rotary_embedding_ref(torch.tensor([0]), x, x.clone(), cs, True)
    a, b, c, s = (float(v) for v in (x[0, 0, 0], x[0, 0, 1], cs[0, 0], cs[0, 1]))
    r = lambda v: torch.tensor(v, dtype=torch.float64).float().bfloat16().float().item()
    assert float(out[0, 0, 0]) == r(a * c - b * s)  # products of bf16 values are exact in fp32/fp64
    assert float(out[0, 0, 1]) == r(b * c + a * s)

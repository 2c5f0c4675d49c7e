"""Which rounding model does the installed vLLM's rotary_embedding follow? (GPU box only)"""
import torch
import vllm._C  # noqa
from vllm import _custom_ops as vops

dev = "cuda"
for dtype in (torch.bfloat16, torch.float16):
    g = torch.Generator().manual_seed(0)
    T, H, D = 64, 8, 128
    q = ((torch.rand(T, H, D, generator=g) * 2 - 1) * 3).to(dtype).to(dev)
    pos = torch.randint(0, 4096, (T,), generator=g).to(dev)
    inv = 1.0 / (10000.0 ** (torch.arange(0, D, 2, dtype=torch.float32) / D))
    f = torch.arange(4096, dtype=torch.float32)[:, None] * inv[None]
    cs = torch.cat([f.cos(), f.sin()], -1).to(dtype).to(dev)
    out = q.clone()
    vops.rotary_embedding(pos, out, None, D, cs, True)
    c, s = cs[pos][:, None, :D // 2], cs[pos][:, None, D // 2:]
    x, y = q[..., :D // 2], q[..., D // 2:]
    xf, yf, cf, sf = x.float(), y.float(), c.float(), s.float()
    models = {
        "per-op rounding in dtype": torch.cat([x * c - y * s, y * c + x * s], -1),
        "fp32, one rounding": torch.cat([(xf * cf - yf * sf).to(dtype), (yf * cf + xf * sf).to(dtype)], -1),
        "fp32 fma(x,c,-(y*s))": torch.cat([torch.addcmul(-(yf * sf), xf, cf).to(dtype), torch.addcmul(xf * sf, yf, cf).to(dtype)], -1),
        "fp64, one rounding": torch.cat([(x.double() * c.double() - y.double() * s.double()).to(dtype),
                                          (y.double() * c.double() + x.double() * s.double()).to(dtype)], -1),
        "products rounded, sum in fp32": torch.cat([((x * c).float() - (y * s).float()).to(dtype), ((y * c).float() + (x * s).float()).to(dtype)], -1),
    }
    for name, m in models.items():
        print(dtype, name, "mismatches:", int((m != out).sum()), "of", out.numel())

"""Out-of-bounds writes, checked by hand (compute-sanitizer is closed on this pool:
profiles/r02_compute_sanitizer_closed.txt says to use bounds checks of one's own): every buffer a kernel WRITES -
output, LSE, split-KV workspace, work plan, the KV cache (cache write), the rotated q / k - is carved out of one
arena with 64 KB guard bands on both sides, filled with a pattern; after the calls the bands must be intact and so
must everything in the written buffers that the call had no business touching (rows past num_tokens, cache slots
no token maps to).  Shapes are chosen to sit on tile and page edges."""
import pytest
import torch

from helpers import BF16_ATOL, BF16_RTOL, FP8_ATOL, assert_close, make_paged_inputs
from oracle.paged_attention import unified_attention_ref

pytestmark = pytest.mark.gpu
GUARD = 64 << 10
PATTERN = 0xA5


class Arena:
    def __init__(self, dev="cuda"):
        self.dev, self.parts = dev, []

    def carve(self, shape, dtype, fill=None):
        n = 1
        for s in shape:
            n *= s
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        pad = (-nbytes) % 256
        buf = torch.full((GUARD + nbytes + pad + GUARD,), PATTERN, dtype=torch.uint8, device=self.dev)
        view = buf[GUARD:GUARD + nbytes].view(dtype).view(shape)
        if fill is not None:
            view.copy_(fill.to(self.dev)) if isinstance(fill, torch.Tensor) else view.fill_(fill)
        self.parts.append((buf, nbytes))
        return view

    def check(self, what):
        torch.cuda.synchronize()
        for i, (buf, nbytes) in enumerate(self.parts):
            lo, hi = buf[:GUARD], buf[GUARD + nbytes + ((-nbytes) % 256):]
            assert bool((lo == PATTERN).all()) and bool((hi == PATTERN).all()), f"{what}: guard band of buffer {i} was written"


CASES = [
    # lens, Hq, Hkv, D, bs, fp8
    ([(1, 4096), (1, 17), (1, 1), (1, 0), (1, 4095), (1, 257)], 32, 8, 128, 16, False),      # split-KV decode
    ([(1, 8191)] * 3 + [(1, 33)], 8, 1, 128, 16, True),                                        # fp8 shard shape, merges
    ([(1, 700), (300, 300), (5, 18), (1, 77), (129, 463), (12, 900), (1, 1)], 32, 8, 128, 16, False),  # mixed: 3 kernels
    ([(1, 700), (300, 300), (64, 1064), (513, 2049)], 32, 8, 128, 16, True),                   # fp8 staging
    ([(255, 255), (257, 513), (1, 64), (128, 128)], 40, 8, 128, 16, False),                    # padded group, tile edges
    ([(255, 255), (257, 513), (1, 64), (128, 128)], 16, 8, 256, 32, False),                    # head size 256
    ([(255, 255), (65, 129), (1, 64), (64, 64)], 64, 8, 64, 16, False),                        # head size 64
]


@pytest.mark.parametrize("lens,Hq,Hkv,D,bs,fp8", CASES)
def test_attention_writes_stay_inside_their_buffers(lens, Hq, Hkv, D, bs, fp8):
    from vllm_triton_backend_b200 import ops

    inp = make_paged_inputs(lens, Hq, Hkv, D, bs, dtype=torch.bfloat16, seed=len(lens) + D, kv_dtype=torch.float8_e4m3fn if fp8 else None,
                            k_scale=0.5, v_scale=0.5)
    T, B = inp["q"].shape[0], len(lens)
    ar = Arena()
    extra = 3  # rows past num_tokens that no kernel may touch
    out_full = ar.carve((T + extra, Hq, D), torch.bfloat16, fill=7.0)
    lse_full = ar.carve((T + extra, Hq), torch.float32, fill=7.0)
    ws_bytes = ops.workspace_bytes(T, B, Hq, Hkv, D) + (ops.fp8_staging_bytes(B, sum(kv for _, kv in lens), bs, Hkv, D) if fp8 else 0)
    ws = ar.carve((ws_bytes,), torch.uint8, fill=0)
    dev = "cuda"
    sc = torch.tensor([0.5], dtype=torch.float32, device=dev) if fp8 else None
    max_q = max(a for a, _ in lens)
    plan = ops.DecodePlan(dev, B, Hq, Hkv, D, bs, torch.bfloat16, torch.float8_e4m3fn if fp8 else torch.bfloat16)
    plan.buf = ar.carve((plan.buf.numel(),), torch.uint8, fill=0)
    cu, sl = inp["cu_seqlens_q"].to(dev), inp["seq_lens"].to(dev)
    for use_plan in (False, True):
        if use_plan:
            plan.build(cu, sl, max_q, T)
        ops.unified_attention(q=inp["q"].to(dev), k=inp["k_cache"].to(dev), v=inp["v_cache"].to(dev), out=out_full[:T], cu_seqlens_q=cu,
                              max_seqlen_q=max_q, seqused_k=sl, max_seqlen_k=max(b for _, b in lens), softmax_scale=D ** -0.5,
                              causal=True, window_size=(-1, -1), block_table=inp["block_table"].to(dev), softcap=0, q_descale=None,
                              k_descale=sc, v_descale=sc, workspace=ws, softmax_lse=lse_full[:T], decode_plan=plan if use_plan else None)
        ar.check(f"attention (plan={use_plan})")
        assert bool((out_full[T:] == 7.0).all()) and bool((lse_full[T:] == 7.0).all()), "rows past num_tokens were written"
        ref = unified_attention_ref(inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"], inp["block_table"],
                                    D ** -0.5, k_scale=0.5 if fp8 else 1.0, v_scale=0.5 if fp8 else 1.0, out_dtype=torch.float32)
        assert_close(out_full[:T].cpu(), ref, atol=FP8_ATOL if fp8 else BF16_ATOL, rtol=BF16_RTOL, what="guarded run")
    assert int(ws[: 1 << 20].view(torch.int32).abs().sum()) == 0, "counters not left at zero"


@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("T", [1, 63, 64, 1000, 4099])
def test_cache_write_touches_only_the_mapped_slots(T, fp8):
    from vllm_triton_backend_b200 import ops

    Hkv, D, bs = 8, 128, 16
    NB = (T + bs - 1) // bs + 5
    g = torch.Generator().manual_seed(T)
    ar = Arena()
    cdt = torch.uint8 if fp8 else torch.bfloat16
    kc = ar.carve((NB, bs, Hkv, D), cdt, fill=0x11 if fp8 else 3.0)
    vc = ar.carve((NB, bs, Hkv, D), cdt, fill=0x11 if fp8 else 3.0)
    key = ar.carve((T, Hkv, D), torch.bfloat16, fill=(torch.rand(T, Hkv, D, generator=g) * 2 - 1).bfloat16())
    val = ar.carve((T, Hkv, D), torch.bfloat16, fill=(torch.rand(T, Hkv, D, generator=g) * 2 - 1).bfloat16())
    slots = torch.randperm(NB * bs, generator=g)[:T].to(torch.int64)
    slots[::7] = -1  # padded tokens
    sc = torch.tensor([0.5], dtype=torch.float32, device="cuda")
    kcv = kc.view(torch.float8_e4m3fn) if fp8 else kc
    vcv = vc.view(torch.float8_e4m3fn) if fp8 else vc
    ops.reshape_and_cache_flash(key, val, kcv, vcv, slots.cuda(), "fp8" if fp8 else "auto", sc, sc)
    ar.check("cache write")
    mapped = torch.zeros(NB * bs, dtype=torch.bool)
    mapped[slots[slots >= 0]] = True
    untouched = kc.view(NB * bs, -1)[~mapped.cuda()]
    assert bool((untouched == (0x11 if fp8 else 3.0)).all()), "a slot no token maps to was written"
    assert bool((vc.view(NB * bs, -1)[~mapped.cuda()] == (0x11 if fp8 else 3.0)).all())

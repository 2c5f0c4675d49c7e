"""Per-(token, head) quantised KV caches (upstream KVQuantMode.FP8_PER_TOKEN_HEAD / INT8_PER_TOKEN_HEAD):
the cache write against the oracle bit for bit and against vLLM's compiled Triton kernel, attention through every
kernel that reads such a cache against the oracle, the fixtures of vLLM's own kernels and the compiled Triton kernel."""
import glob
import os

import numpy as np
import pytest
import torch

from helpers import BF16_RTOL, FP8_ATOL, GOLDEN_DIR, assert_close
from oracle.cache_write import reshape_and_cache_flash_per_token_head_ref
from oracle.paged_attention import unified_attention_ref

pytestmark = pytest.mark.gpu
GENERIC, PREFILL, AUTO = 1, 22, None
E4M3, I8 = torch.float8_e4m3fn, torch.int8
FILES = sorted(glob.glob(os.path.join(GOLDEN_DIR, "pth_*.npz")))


def make_case(lens, Hq, Hkv, D, bs, kv_dtype, dtype=torch.bfloat16, seed=0, layout="planes", spread=3.0):
    """Seeded batch whose caches were WRITTEN by the oracle's per-token-head quantiser: rows of very different
    magnitude (a per-tensor scale would lose the small ones).  `layout`: "planes" = separate dense caches and
    [NB, bs, Hkv] scale tensors; "blocks" = this backend's vLLM allocation (ops.per_token_head_cache_views)."""
    from vllm_triton_backend_b200 import ops

    g = torch.Generator().manual_seed(seed)
    q_lens, kv_lens = [a for a, _ in lens], [b for _, b in lens]
    T = sum(q_lens)
    pages = [(n + bs - 1) // bs for n in kv_lens]
    NB = sum(pages) + 5
    perm = torch.randperm(NB, generator=g)[: sum(pages)].to(torch.int32)
    bt = torch.zeros(len(lens), max(max(pages), 1), dtype=torch.int32)
    slots, o = [], 0
    for i, n in enumerate(pages):
        bt[i, :n] = perm[o: o + n]
        o += n
        slots += [int(bt[i, t // bs]) * bs + t % bs for t in range(kv_lens[i])]
    n_kv = len(slots)
    mag = torch.exp((torch.rand(n_kv, Hkv, 1, generator=g) * 2 - 1) * spread)
    key = ((torch.rand(n_kv, Hkv, D, generator=g) * 2 - 1) * mag).to(dtype)
    value = ((torch.rand(n_kv, Hkv, D, generator=g) * 2 - 1) * mag).to(dtype)
    q = (torch.rand(T, Hq, D, generator=g) * 2 - 1).to(dtype)
    if layout == "blocks":
        raw = torch.zeros(2, NB, bs, Hkv, D + 4, dtype=torch.uint8)
        kc, ks = ops.per_token_head_cache_views(raw[0], D, kv_dtype)
        vc, vs = ops.per_token_head_cache_views(raw[1], D, kv_dtype)
        ks.fill_(1.0), vs.fill_(1.0)
    else:
        raw = None
        kc, vc = torch.zeros(NB, bs, Hkv, D, dtype=kv_dtype), torch.zeros(NB, bs, Hkv, D, dtype=kv_dtype)
        ks, vs = torch.ones(NB, bs, Hkv), torch.ones(NB, bs, Hkv)
    slot_mapping = torch.tensor(slots, dtype=torch.int64)
    reshape_and_cache_flash_per_token_head_ref(key, value, kc, vc, ks, vs, slot_mapping)
    cu = torch.tensor([0] + list(np.cumsum(q_lens)), dtype=torch.int32)
    return dict(q=q, key=key, value=value, slot_mapping=slot_mapping, k_cache=kc, v_cache=vc, k_scale_cache=ks,
                v_scale_cache=vs, block_table=bt, cu_seqlens_q=cu, seq_lens=torch.tensor(kv_lens, dtype=torch.int32),
                raw=raw, NB=NB, bs=bs, Hkv=Hkv, D=D, kv_dtype=kv_dtype)


def to_device(c):
    """Device copies; for the "blocks" layout the views are re-made over the device copy of the allocation."""
    from vllm_triton_backend_b200 import ops

    d = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in c.items() if k != "raw"}
    if c["raw"] is not None:
        raw = c["raw"].cuda()
        d["k_cache"], d["k_scale_cache"] = ops.per_token_head_cache_views(raw[0], c["D"], c["kv_dtype"])
        d["v_cache"], d["v_scale_cache"] = ops.per_token_head_cache_views(raw[1], c["D"], c["kv_dtype"])
        d["raw"] = raw
    return d


def run_attention(c, scale, force=AUTO, window=0, softcap=0.0, sinks=None, alibi=None, lse=False, decode_plan=None):
    from vllm_triton_backend_b200 import _cabi, ops

    d = to_device(c)
    q = d["q"]
    out = torch.full_like(q, float("nan"))
    mq = int((c["cu_seqlens_q"][1:] - c["cu_seqlens_q"][:-1]).max())
    if force == PREFILL and mq == 1:
        force = 2
    lse_t = torch.full((q.shape[0], q.shape[1]), float("nan"), dtype=torch.float32, device="cuda") if lse else None
    ops.unified_attention(
        q=q, k=d["k_cache"], v=d["v_cache"], out=out, cu_seqlens_q=d["cu_seqlens_q"], max_seqlen_q=mq, seqused_k=d["seq_lens"],
        max_seqlen_k=int(c["seq_lens"].max()), softmax_scale=scale, causal=True,
        window_size=(window - 1, 0) if window else (-1, -1), block_table=d["block_table"], softcap=softcap, q_descale=None,
        k_descale=None, v_descale=None, force_selection=force, sinks=None if sinks is None else sinks.cuda(),
        alibi_slopes=None if alibi is None else alibi.cuda(), softmax_lse=lse_t,
        kv_quant_mode=2 if c["kv_dtype"] == I8 else 3, k_scale_cache=d["k_scale_cache"], v_scale_cache=d["v_scale_cache"])
    torch.cuda.synchronize()
    return out.cpu(), (lse_t.cpu() if lse else None), _cabi.last_kernel()


def oracle(c, scale, window=0, softcap=0.0, sinks=None, alibi=None, lse=False):
    return unified_attention_ref(
        c["q"], c["k_cache"], c["v_cache"], c["cu_seqlens_q"], c["seq_lens"], c["block_table"], scale, sliding_window=window,
        softcap=softcap, sinks=sinks, alibi_slopes=alibi, k_scale_cache=c["k_scale_cache"], v_scale_cache=c["v_scale_cache"],
        out_dtype=torch.float32, return_lse=lse)


def tol(ref):
    """BASELINE.json: FP8 KV within 6e-2 of the reference kernels on outputs of order one.  Values here are not of
    order one (rows span e^+-3, logits reach +-10), so the bound scales with the largest output magnitude - at half
    of BASELINE's figure - plus 1e-2 relative.  (The staged tcgen05 path forms round_to_bf16(K * scale) like the
    per-tensor mode does; with logits this large that rounding alone moves single outputs by 2 % of the scale.)"""
    return FP8_ATOL * max(1.0, float(ref.abs().max())) / 2, BF16_RTOL


# ---- cache write ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kv_dtype", [E4M3, I8], ids=["e4m3", "int8"])
@pytest.mark.parametrize("src", [torch.bfloat16, torch.float16, torch.float32], ids=["bf16", "f16", "f32"])
@pytest.mark.parametrize("D,Hkv,bs,layout", [(128, 8, 16, "planes"), (128, 2, 16, "blocks"), (256, 1, 32, "blocks"),
                                             (64, 4, 16, "planes"), (80, 3, 16, "planes"), (66, 2, 16, "planes")])
def test_cache_write_bit_exact_against_the_oracle(kv_dtype, src, D, Hkv, bs, layout):
    from vllm_triton_backend_b200 import _cabi, ops

    c = make_case([(1, 70), (1, 33), (1, 200)], Hkv, Hkv, D, bs, kv_dtype, dtype=src, seed=D + Hkv, layout=layout)
    d = to_device(c)
    # start from EMPTY device caches and write them with the kernel; tokens in a shuffled order, some padding slots
    for k in ("k_cache", "v_cache"):
        d[k].view(torch.uint8).zero_()
    d["k_scale_cache"].fill_(1.0), d["v_scale_cache"].fill_(1.0)
    g = torch.Generator().manual_seed(1)
    order = torch.randperm(c["key"].shape[0], generator=g)
    sm = c["slot_mapping"][order].clone()
    pad = torch.arange(0, sm.shape[0], 17)
    # a padding token's row must not be written: move its slot's expected content back to "never written"
    exp_k, exp_v = c["k_cache"].clone(), c["v_cache"].clone()
    exp_ks, exp_vs = c["k_scale_cache"].clone(), c["v_scale_cache"].clone()
    for sl in sm[pad].tolist():
        b, o = sl // bs, sl % bs
        exp_k.view(torch.uint8)[b, o] = 0
        exp_v.view(torch.uint8)[b, o] = 0
        exp_ks[b, o] = 1.0
        exp_vs[b, o] = 1.0
    sm[pad] = -1
    # key / value as views into a wider buffer (vLLM hands views of the fused qkv projection)
    wide = torch.zeros(sm.shape[0], 3 * Hkv * D + 8, dtype=src, device="cuda")
    key = wide[:, 8: 8 + Hkv * D].view(-1, Hkv, D) if (8 * wide.element_size()) % 16 == 0 and D % 4 == 0 else None
    if key is None:
        key, value = c["key"][order].cuda(), c["value"][order].cuda()
    else:
        value = wide[:, 8 + Hkv * D: 8 + 2 * Hkv * D].view(-1, Hkv, D)
        key.copy_(c["key"][order]), value.copy_(c["value"][order])
    ops.reshape_and_cache_flash_per_token_head(key, value, d["k_cache"], d["v_cache"], d["k_scale_cache"], d["v_scale_cache"],
                                               sm.cuda())
    torch.cuda.synchronize()
    assert _cabi.last_kernel() == ("cache_quant_per_token_head_vec" if D % 4 == 0 else "cache_quant_per_token_head")
    assert torch.equal(d["k_scale_cache"].cpu(), exp_ks) and torch.equal(d["v_scale_cache"].cpu(), exp_vs)
    assert torch.equal(d["k_cache"].cpu().view(torch.uint8), exp_k.view(torch.uint8))
    assert torch.equal(d["v_cache"].cpu().view(torch.uint8), exp_v.view(torch.uint8))
    if layout == "blocks":
        # nothing outside the data and scale regions of the allocation was touched (there is nothing else in it)
        assert d["raw"].numel() == 2 * c["NB"] * bs * Hkv * (D + 4)


@pytest.mark.parametrize("kv_dtype", [E4M3, I8], ids=["e4m3", "int8"])
def test_cache_write_matches_vllm_compiled_kernel(kv_dtype):
    """vLLM's own Triton kernel, compiled (not interpreted) on this GPU: same scales, same codes.  Its fp32 division
    is Triton's (not guaranteed IEEE), so a scale may differ in the last bit and then a code by one step; such rows
    are counted, not hidden."""
    try:
        from vllm.v1.attention.ops.triton_reshape_and_cache_flash import (
            triton_reshape_and_cache_flash_per_token_head_quant as theirs,
        )
    except Exception as e:  # pragma: no cover
        pytest.skip(f"vLLM kernel not importable: {e}")
    from vllm_triton_backend_b200 import ops

    T, H, D, bs, NB = 4096, 8, 128, 16, 300
    g = torch.Generator().manual_seed(3)
    key = (torch.randn(T, H, D, generator=g) * torch.exp(torch.randn(T, H, 1, generator=g) * 2)).bfloat16().cuda()
    value = (torch.randn(T, H, D, generator=g) * torch.exp(torch.randn(T, H, 1, generator=g) * 2)).bfloat16().cuda()
    sm = torch.randperm(NB * bs, generator=g)[:T].cuda()
    res = []
    for fn in (ops.reshape_and_cache_flash_per_token_head, theirs):
        kc = torch.zeros(NB, bs, H, D, dtype=kv_dtype, device="cuda")
        vc = torch.zeros_like(kc)
        ks = torch.ones(NB, bs, H, device="cuda")
        vs = torch.ones_like(ks)
        fn(key, value, kc, vc, ks, vs, sm)
        torch.cuda.synchronize()
        res.append((kc.view(torch.uint8).cpu(), vc.view(torch.uint8).cpu(), ks.cpu(), vs.cpu()))
    (kc0, vc0, ks0, vs0), (kc1, vc1, ks1, vs1) = res
    for s0, s1 in ((ks0, ks1), (vs0, vs1)):
        assert torch.allclose(s0, s1, rtol=2.5e-7, atol=0), float(((s0 - s1) / s1).abs().max())  # <= 2 ulp
    scale_same = float(((ks0 == ks1).float().mean() + (vs0 == vs1).float().mean()) / 2)
    code_same = float(((kc0 == kc1).float().mean() + (vc0 == vc1).float().mean()) / 2)
    print(f"scales identical: {scale_same:.6f}, codes identical: {code_same:.6f}")
    assert code_same > 0.999
    # codes never differ by more than one step
    for a, b in ((kc0, kc1), (vc0, vc1)):
        if kv_dtype == I8:
            assert int((a.view(torch.int8).int() - b.view(torch.int8).int()).abs().max()) <= 1
        else:
            fa, fb = a.view(E4M3).float(), b.view(E4M3).float()
            assert float(((fa - fb).abs() / fa.abs().clamp(min=2 ** -6)).max()) <= 0.126


# ---- fixtures of vLLM's own kernels ----------------------------------------------------------------------------
def _load_fixture(path):
    z = np.load(path)
    kv_dtype = I8 if str(z["kv_dtype"]) == "int8" else E4M3
    c = {k: torch.from_numpy(z[k]) for k in z.files if z[k].dtype.kind in "fiu" and z[k].ndim > 0}
    # the interpreter's e4m3 rounding is defective (oracle/make_golden_pth.py): the caches the attention fixture was
    # computed FROM are the fixture's own, so use those bytes as they are
    c["k_cache"], c["v_cache"] = c["k_cache"].view(kv_dtype), c["v_cache"].view(kv_dtype)
    c.update(kv_dtype=kv_dtype, raw=None, D=c["q"].shape[2], scale=float(z["scale"]), window=int(z["window"]),
             softcap=float(z["softcap"]))
    return c


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(p)[4:-4] for p in FILES])
@pytest.mark.parametrize("force", [AUTO, GENERIC, PREFILL], ids=["auto", "simt", "mma"])
def test_attention_matches_the_fixtures_of_vllms_kernels(path, force):
    c = _load_fixture(path)
    if force == PREFILL and c["D"] not in (64, 128, 256):
        pytest.skip("mma.sync kernel: head size")
    out, _, kern = run_attention(c, c["scale"], force, c["window"], c["softcap"])
    ref = c["out_triton"].float()
    assert_close(out, ref, atol=4e-3 * max(1.0, float(ref.abs().max())), rtol=BF16_RTOL, what=f"{kern} vs vLLM fixture")


# ---- every kernel that reads such a cache, against the oracle ---------------------------------------------------
DECODE = [(1, 700), (1, 33), (1, 4096), (1, 1), (1, 0), (1, 257), (1, 16), (1, 2000)]
MIXED = [(1, 700), (300, 300), (5, 18), (1, 77), (129, 463), (12, 900), (1, 0)]
SHORT = [(1, 700), (5, 18), (3, 463), (16, 64), (1, 0), (2, 2)]


@pytest.mark.parametrize("kv_dtype", [E4M3, I8], ids=["e4m3", "int8"])
@pytest.mark.parametrize("feature", ["plain", "window", "softcap", "sinks", "alibi"])
@pytest.mark.parametrize("D,Hq,Hkv,bs,layout", [(128, 32, 8, 16, "blocks"), (128, 16, 1, 16, "planes"), (256, 8, 2, 32, "blocks"),
                                                (128, 24, 2, 64, "planes")])
def test_decode_kernel(kv_dtype, feature, D, Hq, Hkv, bs, layout):
    c = make_case(DECODE, Hq, Hkv, D, bs, kv_dtype, seed=7, layout=layout)
    window = 300 if feature == "window" else 0
    softcap = 20.0 if feature == "softcap" else 0.0
    sinks = torch.linspace(-2.0, 4.0, Hq) if feature == "sinks" else None
    alibi = torch.tensor([2.0 ** (-(i + 1) * 8.0 / Hq) for i in range(Hq)]) if feature == "alibi" else None
    scale = D ** -0.5
    out, lse, kern = run_attention(c, scale, AUTO, window, softcap, sinks, alibi, lse=True)
    assert kern == "attn_decode_split_kv"
    ref, ref_lse = oracle(c, scale, window, softcap, sinks, alibi, lse=True)
    a, r = tol(ref)
    assert_close(out, ref, atol=a, rtol=r, what=f"decode {feature}")
    fin = torch.isfinite(ref_lse)
    # logits reach +-10 here; the staged path's round_to_bf16(K * scale) moves them by up to 0.4 %
    assert torch.equal(torch.isfinite(lse), fin)
    assert float((lse[fin] - ref_lse[fin]).abs().max()) < 1e-2 * max(1.0, float(ref_lse[fin].abs().max()))


@pytest.mark.parametrize("kv_dtype", [E4M3, I8], ids=["e4m3", "int8"])
@pytest.mark.parametrize("force,lens,D,Hq,Hkv", [
    (GENERIC, SHORT, 128, 8, 2), (GENERIC, SHORT, 96, 6, 3), (GENERIC, MIXED, 64, 4, 4),
    (PREFILL, SHORT, 128, 16, 4), (PREFILL, MIXED, 64, 8, 2), (PREFILL, MIXED, 256, 4, 1), (PREFILL, MIXED, 128, 12, 4),
    (AUTO, MIXED, 128, 32, 8), (AUTO, MIXED, 256, 8, 2), (AUTO, MIXED, 64, 8, 2), (AUTO, SHORT, 128, 16, 2),
    (AUTO, [(700, 700), (1500, 3000)], 128, 16, 4),
])
@pytest.mark.parametrize("feature", ["plain", "softcap_window"])
def test_prefill_and_mixed_batches(kv_dtype, force, lens, D, Hq, Hkv, feature):
    c = make_case(lens, Hq, Hkv, D, 16, kv_dtype, seed=11, layout="blocks" if D % 4 == 0 and D != 96 else "planes")
    window, softcap = (200, 30.0) if feature == "softcap_window" else (0, 0.0)
    scale = D ** -0.5
    out, lse, kern = run_attention(c, scale, force, window, softcap, lse=True)
    ref, ref_lse = oracle(c, scale, window, softcap, lse=True)
    a, r = tol(ref)
    assert_close(out, ref, atol=a, rtol=r, what=f"{kern} {feature}")
    fin = torch.isfinite(ref_lse)
    # logits reach +-10 here; the staged path's round_to_bf16(K * scale) moves them by up to 0.4 %
    assert torch.equal(torch.isfinite(lse), fin)
    assert float((lse[fin] - ref_lse[fin]).abs().max()) < 1e-2 * max(1.0, float(ref_lse[fin].abs().max()))


def test_auto_routes_long_prompts_through_staging_and_tcgen05():
    from vllm_triton_backend_b200 import _cabi

    c = make_case([(700, 700), (1, 900), (1500, 3000), (4, 100)], 16, 4, 128, 16, E4M3, seed=2, layout="blocks")
    _cabi.load_library().b200attn_reset_launch_count()
    out, _, kern = run_attention(c, 128 ** -0.5)
    # plan + decode (forked) | stage plan + dequant + tcgen05 | mma.sync for the 4-token extend
    assert _cabi.load_library().b200attn_launch_count() == 6
    ref = oracle(c, 128 ** -0.5)
    a, r = tol(ref)
    assert_close(out, ref, atol=a, rtol=r, what="mixed per-token-head batch")


def test_what_a_per_tensor_scale_would_lose():
    """The point of the mode: rows whose magnitudes span e^+-5 keep ~3 significant bits each; one scale for the whole
    tensor flushes the small rows to zero.  The per-token-head result stays close to attention over the unquantised
    values, the per-tensor one does not."""
    from vllm_triton_backend_b200 import ops

    lens = [(1, 512), (1, 100)]
    c = make_case(lens, 8, 2, 128, 16, E4M3, seed=5, spread=5.0)
    out, _, _ = run_attention(c, 128 ** -0.5)
    # unquantised reference from the source rows
    kc = torch.zeros(c["NB"], 16, 2, 128, dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    kc.view(-1, 2, 128)[c["slot_mapping"]] = c["key"]
    vc.view(-1, 2, 128)[c["slot_mapping"]] = c["value"]
    exact = unified_attention_ref(c["q"], kc, vc, c["cu_seqlens_q"], c["seq_lens"], c["block_table"], 128 ** -0.5,
                                  out_dtype=torch.float32)
    # per-tensor e4m3 of the same rows, through the per-tensor path of this library
    amax = float(max(c["key"].float().abs().max(), c["value"].float().abs().max()))
    s = torch.tensor([amax / 448.0], device="cuda")
    k8 = torch.zeros(c["NB"], 16, 2, 128, dtype=E4M3, device="cuda")
    v8 = torch.zeros_like(k8)
    ops.reshape_and_cache_flash(c["key"].cuda(), c["value"].cuda(), k8, v8, c["slot_mapping"].cuda(), "fp8", s, s)
    out_pt = torch.empty_like(c["q"]).cuda()
    ops.unified_attention(q=c["q"].cuda(), k=k8, v=v8, out=out_pt, cu_seqlens_q=c["cu_seqlens_q"].cuda(), max_seqlen_q=1,
                          seqused_k=c["seq_lens"].cuda(), max_seqlen_k=512, softmax_scale=128 ** -0.5, causal=True,
                          window_size=(-1, -1), block_table=c["block_table"].cuda(), softcap=0.0, q_descale=None, k_descale=s,
                          v_descale=s)
    err_pth = float((out.float() - exact).abs().max() / exact.abs().max())
    err_pt = float((out_pt.cpu().float() - exact).abs().max() / exact.abs().max())
    print(f"relative error vs unquantised: per-token-head {err_pth:.4f}, per-tensor {err_pt:.4f}")
    assert err_pth < 0.08 and err_pth < err_pt


def test_argument_checks():
    from vllm_triton_backend_b200 import ops

    c = make_case([(1, 40)], 4, 2, 128, 16, I8)
    d = to_device(c)
    kw = dict(q=d["q"], k=d["k_cache"], v=d["v_cache"], out=torch.empty_like(d["q"]), cu_seqlens_q=d["cu_seqlens_q"], max_seqlen_q=1,
              seqused_k=d["seq_lens"], max_seqlen_k=40, softmax_scale=0.1, causal=True, window_size=(-1, -1),
              block_table=d["block_table"], softcap=0.0, q_descale=None, k_descale=None, v_descale=None)
    with pytest.raises(AssertionError, match="requires k_scale_cache"):
        ops.unified_attention(**kw, kv_quant_mode=2)
    with pytest.raises(ValueError, match="int8 kv cache needs"):
        ops.unified_attention(**kw)
    with pytest.raises(ValueError, match="does not match the cache dtype"):
        ops.unified_attention(**kw, kv_quant_mode=3, k_scale_cache=d["k_scale_cache"], v_scale_cache=d["v_scale_cache"])
    with pytest.raises(NotImplementedError):
        ops.unified_attention(**kw, kv_quant_mode=4)


@pytest.mark.parametrize("kv_dtype", [E4M3, I8], ids=["e4m3", "int8"])
@pytest.mark.parametrize("lens,D,Hq,Hkv", [(DECODE, 128, 32, 8), (MIXED, 128, 16, 4), (SHORT, 64, 8, 2)])
def test_against_vllms_compiled_triton_kernel(kv_dtype, lens, D, Hq, Hkv):
    """The kernel this mode comes from, compiled on this GPU, on the same device tensors."""
    try:
        from vllm.v1.attention.ops.triton_unified_attention import unified_attention as theirs
        from vllm.v1.kv_cache_interface import KVQuantMode
    except Exception as e:  # pragma: no cover
        pytest.skip(f"vLLM kernel not importable: {e}")
    lens = [x for x in lens if x[1] > 0]
    c = make_case(lens, Hq, Hkv, D, 16, kv_dtype, seed=13)
    d = to_device(c)
    scale = D ** -0.5
    out, _, kern = run_attention(c, scale)
    ref = torch.zeros_like(d["q"])
    mq = max(a for a, _ in lens)
    theirs(q=d["q"], k=d["k_cache"], v=d["v_cache"], out=ref, cu_seqlens_q=d["cu_seqlens_q"], max_seqlen_q=mq,
           seqused_k=d["seq_lens"], max_seqlen_k=max(b for _, b in lens), softmax_scale=scale, causal=True, window_size=(-1, -1),
           block_table=d["block_table"], softcap=0.0, q_descale=None, k_descale=None, v_descale=None,
           kv_quant_mode=KVQuantMode.INT8_PER_TOKEN_HEAD if kv_dtype == I8 else KVQuantMode.FP8_PER_TOKEN_HEAD,
           k_scale_cache=d["k_scale_cache"], v_scale_cache=d["v_scale_cache"])
    torch.cuda.synchronize()
    ref = ref.cpu().float()
    a, r = tol(ref)
    assert_close(out, ref, atol=a, rtol=r, what=f"{kern} vs vLLM Triton (per-token-head)")


def test_backend_roundtrip_through_the_plugin_classes():
    """B200AttentionImpl with kv_cache_dtype="fp8_per_token_head": the padded allocation of get_kv_cache_shape, the cache
    write and the attention call the way vLLM drives them (do_kv_cache_update, then forward)."""
    pytest.importorskip("vllm")
    from types import SimpleNamespace

    from vllm_triton_backend_b200 import ops
    from vllm_triton_backend_b200.backend.attn import B200AttentionBackend, B200AttentionImpl, B200AttentionMetadata

    Hq, Hkv, D, bs = 16, 4, 128, 16
    for name, dt in (("fp8_per_token_head", E4M3), ("int8_per_token_head", I8)):
        c = make_case([(1, 300), (40, 40), (1, 77), (6, 50)], Hq, Hkv, D, bs, dt, seed=21)
        shape = B200AttentionBackend.get_kv_cache_shape(c["NB"], bs, Hkv, D, name)
        assert shape == (2, c["NB"], bs, Hkv, D + 4)
        kv_cache = torch.zeros(shape, dtype=torch.uint8, device="cuda")
        impl = B200AttentionImpl(Hq, D, D ** -0.5, Hkv, None, None, name)
        layer = SimpleNamespace(_k_scale=torch.ones(1, device="cuda"), _v_scale=torch.ones(1, device="cuda"))
        impl.do_kv_cache_update(layer, c["key"].cuda(), c["value"].cuda(), kv_cache, c["slot_mapping"].cuda())
        T = c["q"].shape[0]
        ws = torch.zeros(ops.workspace_bytes(T, 4, Hq, Hkv, D) + ops.fp8_staging_bytes(4, 1024, bs, Hkv, D), dtype=torch.uint8,
                         device="cuda")
        md = B200AttentionMetadata(
            num_actual_tokens=T, max_query_len=40, query_start_loc=c["cu_seqlens_q"].cuda(), max_seq_len=300,
            seq_lens=c["seq_lens"].cuda(), block_table=c["block_table"].cuda(), slot_mapping=c["slot_mapping"].cuda(),
            workspace=ws)
        out = torch.empty(T, Hq * D, dtype=torch.bfloat16, device="cuda")
        impl.forward(layer, c["q"].cuda().view(T, Hq * D), None, None, kv_cache, md, output=out)
        torch.cuda.synchronize()
        ref = oracle(c, D ** -0.5)
        a, r = tol(ref)
        assert_close(out.view(T, Hq, D).cpu(), ref, atol=a, rtol=r, what=f"plugin {name}")

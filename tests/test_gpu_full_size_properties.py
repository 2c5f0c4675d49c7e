"""Parity at BASELINE.json's full sizes (C2, C3, C4) through properties that do not need the CPU oracle,
which takes minutes at these sizes:

* retrieval: one key per (sequence, kv head) carries a logit so large that every other softmax weight
  underflows to exactly 0 -- the output row must then equal, bit for bit, the value row the block table
  points at.  That is the index arithmetic of the path (page lookup, in-page slot, head offset, split
  merge, GQA packing) checked over the whole cache;
* constant values: with every visible value row equal to c the output is c up to 16-bit rounding, and
  with NaN in every slot past seq_len any out-of-range read shows up as NaN;
* relabelling the physical pages (cache moved accordingly) leaves every output bit unchanged;
* linearity in V;
* cache write: written tokens read back bit-exact through the slot mapping, untouched slots stay untouched.

Inputs are generated on the device; all calls go through the C ABI (ops.*)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda"
C2 = dict(B=64, ctx=4096, Hq=32, Hkv=8, D=128, bs=16, fp8=False)             # BASELINE.json configs[1]
C4 = dict(B=64, ctx=32768, Hq=64, Hkv=8, D=128, bs=16, fp8=True)             # configs[3], all 8 kv heads on one GPU
C4_SHARD = dict(B=64, ctx=32768, Hq=8, Hkv=1, D=128, bs=16, fp8=True)        # configs[3], the 8-GPU shard


def _gen(seed):
    return torch.Generator(device=DEV).manual_seed(seed)


def _paged(cfg, seq_lens, seed=0, extra_pages=64):
    """Block table (random permutation of the physical pages, as the bandwidth runs use) for `seq_lens`."""
    bs = cfg["bs"]
    pages = [(n + bs - 1) // bs for n in seq_lens]
    NB = sum(pages) + extra_pages
    perm = torch.randperm(NB, generator=_gen(seed), device=DEV)[: sum(pages)].to(torch.int32)
    bt = torch.zeros(len(seq_lens), max(pages), dtype=torch.int32, device=DEV)
    o = 0
    for i, n in enumerate(pages):
        bt[i, :n] = perm[o:o + n]
        o += n
    return NB, bt


def _cache_dtype(cfg):
    return torch.float8_e4m3fn if cfg["fp8"] else torch.bfloat16


def _rand_cache(cfg, NB, seed, lo=-1.0, hi=1.0):
    shape = (NB, cfg["bs"], cfg["Hkv"], cfg["D"])
    x = torch.empty(shape, dtype=torch.bfloat16, device=DEV).uniform_(lo, hi, generator=_gen(seed))
    return x.to(_cache_dtype(cfg))


def _attend(cfg, q, kc, vc, cu, seq_lens_t, bt, max_q, max_k, window=0, k_scale=1.0, v_scale=1.0, scale=None):
    from vllm_triton_backend_b200 import ops

    out = torch.full_like(q, float("nan"))
    sc = [torch.tensor([s], dtype=torch.float32, device=DEV) for s in (k_scale, v_scale)]
    ops.unified_attention(
        q=q, k=kc, v=vc, out=out, cu_seqlens_q=cu, max_seqlen_q=max_q, seqused_k=seq_lens_t, max_seqlen_k=max_k,
        softmax_scale=cfg["D"] ** -0.5 if scale is None else scale, causal=True,
        window_size=(window - 1, 0) if window > 0 else (-1, -1), block_table=bt, softcap=0.0, q_descale=None,
        k_descale=sc[0], v_descale=sc[1])
    torch.cuda.synchronize()
    return out


def _decode_meta(seq_lens):
    B = len(seq_lens)
    cu = torch.arange(B + 1, dtype=torch.int32, device=DEV)
    sl = torch.tensor(seq_lens, dtype=torch.int32, device=DEV)
    return cu, sl


_ONE = {True: 0x38, False: 0x3F80}  # 1.0 as e4m3 / bf16 bits


def _bits(cache):
    """Integer view of a cache (index_put / gather are not implemented for float8 tensors)."""
    return cache.view(torch.uint8 if cache.element_size() == 1 else torch.int16)


def _nan_fill_unused(cfg, cache, bt, seq_lens):
    """NaN into every slot no sequence owns: slots past seq_len in a last page and all unallocated pages."""
    bs = cfg["bs"]
    used = torch.zeros(cache.shape[0] * bs, dtype=torch.bool, device=DEV)
    for b, n in enumerate(seq_lens):
        pos = torch.arange(n, device=DEV)
        used[bt[b, pos // bs].long() * bs + pos % bs] = True
    flat = _bits(cache).view(cache.shape[0] * bs, -1)
    flat[~used] = 0x7F if cfg["fp8"] else 0x7FC0  # e4m3 NaN / bf16 NaN
    return cache


# ---------------------------------------------------------------------------------------------
# retrieval, decode (C2, C4, the C4 8-GPU shard, a ragged batch)
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name,cfg,ragged", [("c2", C2, False), ("c2-ragged", C2, True), ("c4", C4, False),
                                             ("c4-shard", C4_SHARD, False), ("c4-shard-ragged", C4_SHARD, True)])
def test_decode_retrieves_the_row_the_block_table_points_at(name, cfg, ragged):
    B, Hq, Hkv, D, bs = cfg["B"], cfg["Hq"], cfg["Hkv"], cfg["D"], cfg["bs"]
    g = Hq // Hkv
    if ragged:  # 1 .. ctx, including one-token and page-boundary lengths
        seq_lens = [1, bs, bs + 1, cfg["ctx"], cfg["ctx"] - 1] + torch.randint(
            1, cfg["ctx"] + 1, (B - 5,), generator=torch.Generator().manual_seed(3)).tolist()
    else:
        seq_lens = [cfg["ctx"]] * B
    NB, bt = _paged(cfg, seq_lens, seed=1)
    cu, sl = _decode_meta(seq_lens)
    v_scale = 0.5 if cfg["fp8"] else 1.0  # a power of two: the expected row stays exact
    vc = _nan_fill_unused(cfg, _rand_cache(cfg, NB, seed=2, lo=-4.0, hi=4.0), bt, seq_lens)
    kc = torch.zeros(NB, bs, Hkv, D, dtype=torch.bfloat16, device=DEV).to(_cache_dtype(cfg))
    # the hot key of (b, h): random, but the first / last position for some rows
    pstar = (torch.rand(B, Hkv, generator=_gen(4), device=DEV) * sl[:, None]).long()
    pstar = torch.minimum(pstar, sl[:, None].long() - 1)
    pstar[0::7] = 0
    pstar[1::7] = (sl[1::7, None].long() - 1).expand(-1, Hkv)
    page = torch.gather(bt.long(), 1, pstar // bs)  # [B, Hkv]
    slot = pstar % bs
    h_idx = torch.arange(Hkv, device=DEV)[None].expand(B, -1)
    _bits(kc)[page, slot, h_idx, 0] = _ONE[cfg["fp8"]]
    q = torch.zeros(B, Hq, D, dtype=torch.bfloat16, device=DEV)
    q[:, :, 0] = 2048.0  # logit 181 at the hot key, 0 elsewhere: every other weight is exactly 0
    out = _attend(cfg, q, kc, vc, cu, sl, bt, 1, max(seq_lens), v_scale=v_scale)
    want = (_bits(vc)[page, slot, h_idx].view(vc.dtype).float() * v_scale).to(torch.bfloat16)  # [B, Hkv, D]
    want = want[:, :, None, :].expand(B, Hkv, g, D).reshape(B, Hq, D)
    assert torch.equal(out, want), f"{name}: {(out != want).any(-1).sum().item()} of {B * Hq} rows differ"


# ---------------------------------------------------------------------------------------------
# retrieval, C3-shaped mixed batch: decode rows + fresh 8k prompt + chunked continuations + prefix hits.
# Key p of every sequence is one-hot in dimension p % 128, query at position pos asks for pos % 128, and a
# sliding window of 128 leaves exactly one matching key in sight: pos itself.  One call, three kernels.
# ---------------------------------------------------------------------------------------------
def _c3_lens():
    import random

    rnd = random.Random(0)
    lens = [(1, rnd.randint(128, 8192)) for _ in range(32)]                 # decode rows
    lens += [(8192, 8192), (2048, 6144), (1024, 7168)]                       # fresh prompt, chunked continuations
    lens += [(rnd.randint(128, 1024), 0) for _ in range(4)]                  # prefix-cache hits
    lens = [(q, kv if kv else q + 1024 * (1 + i % 2)) for i, (q, kv) in enumerate(lens)]
    lens += [(5, 40), (17, 17), (64, 1064)]                                  # short extends
    rnd.shuffle(lens)
    return lens


@pytest.mark.parametrize("fp8", [False, True])
def test_c3_mixed_batch_retrieves_its_own_position(fp8):
    from vllm_triton_backend_b200 import _cabi

    cfg = dict(C2, fp8=fp8)
    Hq, Hkv, D, bs = cfg["Hq"], cfg["Hkv"], cfg["D"], cfg["bs"]
    g = Hq // Hkv
    lens = _c3_lens()
    q_lens, seq_lens = [a for a, _ in lens], [b for _, b in lens]
    T = sum(q_lens)
    NB, bt = _paged(cfg, seq_lens, seed=5)
    cu = torch.tensor([0] + torch.tensor(q_lens).cumsum(0).tolist(), dtype=torch.int32, device=DEV)
    sl = torch.tensor(seq_lens, dtype=torch.int32, device=DEV)
    v_scale = 0.5 if fp8 else 1.0
    vc = _nan_fill_unused(cfg, _rand_cache(cfg, NB, seed=6, lo=-4.0, hi=4.0), bt, seq_lens)
    kc = torch.zeros(NB, bs, Hkv, D, dtype=torch.bfloat16, device=DEV).to(_cache_dtype(cfg))
    q = torch.zeros(T, Hq, D, dtype=torch.bfloat16, device=DEV)
    want = torch.empty(T, Hq, D, dtype=torch.bfloat16, device=DEV)
    kb, vb = _bits(kc), _bits(vc)
    for b, (ql, n) in enumerate(lens):
        pos = torch.arange(n, device=DEV)
        pages, slots = bt[b, pos // bs].long(), pos % bs
        kb[pages, slots, :, pos % 128] = _ONE[fp8]
        qpos = torch.arange(n - ql, n, device=DEV)  # absolute positions of this step's query tokens
        t0 = int(cu[b])
        q[t0 + torch.arange(ql, device=DEV), :, qpos % 128] = 2048.0
        rows = vb[pages[qpos], slots[qpos]].view(vc.dtype).float() * v_scale  # [ql, Hkv, D]
        want[t0:t0 + ql] = rows.to(torch.bfloat16)[:, :, None, :].expand(ql, Hkv, g, D).reshape(ql, Hq, D)
    _cabi.reset_launch_count()
    out = _attend(cfg, q, kc, vc, cu, sl, bt, max(q_lens), max(seq_lens), window=128, v_scale=v_scale)
    assert _cabi.launch_count() >= 3, "long prompts, short extends and decode rows take different kernels"
    bad = (out != want).any(-1)
    assert not bad.any(), f"{bad.sum().item()} of {T * Hq} rows differ; first token {bad.any(-1).nonzero()[0].item()}"


# ---------------------------------------------------------------------------------------------
# constant values, page relabelling, linearity in V (C2 / C4 decode, 8k prompt)
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name,cfg,q_len", [("c2", C2, 1), ("c4", C4, 1), ("prompt8k", dict(C2, B=2, ctx=8192), 8192),
                                            ("prompt8k-fp8", dict(C2, B=2, ctx=8192, fp8=True), 8192)])
def test_constant_values_come_back_and_nothing_past_seq_len_is_read(name, cfg, q_len):
    B, Hq, Hkv, D, bs = cfg["B"], cfg["Hq"], cfg["Hkv"], cfg["D"], cfg["bs"]
    seq_lens = [cfg["ctx"] - 3 * i for i in range(B)]  # ragged ends: NaN slots inside the last pages
    q_lens = [min(q_len, n) for n in seq_lens]
    NB, bt = _paged(cfg, seq_lens, seed=7)
    cu = torch.tensor([0] + torch.tensor(q_lens).cumsum(0).tolist(), dtype=torch.int32, device=DEV)
    sl = torch.tensor(seq_lens, dtype=torch.int32, device=DEV)
    c = torch.empty(Hkv, D, dtype=torch.bfloat16, device=DEV).uniform_(-3, 3, generator=_gen(8)).to(_cache_dtype(cfg))
    vc = _nan_fill_unused(cfg, _bits(c)[None, None].expand(NB, bs, Hkv, D).contiguous().view(c.dtype), bt, seq_lens)
    kc = _nan_fill_unused(cfg, _rand_cache(cfg, NB, seed=9), bt, seq_lens)
    q = torch.empty(sum(q_lens), Hq, D, dtype=torch.bfloat16, device=DEV).uniform_(-1, 1, generator=_gen(10))
    out = _attend(cfg, q, kc, vc, cu, sl, bt, max(q_lens), max(seq_lens), k_scale=0.5 if cfg["fp8"] else 1.0)
    want = c.float()[:, None, :].expand(Hkv, Hq // Hkv, D).reshape(1, Hq, D)
    assert not torch.isnan(out).any(), f"{name}: a slot past seq_len leaked into the output"
    err = (out.float() - want).abs().max().item()
    assert err <= 2e-2, f"{name}: softmax weights do not sum to one (max abs err {err})"


def test_relabelling_physical_pages_changes_no_bit():
    cfg = C2
    B = cfg["B"]
    seq_lens = [cfg["ctx"] - 5 * i for i in range(B)]
    NB, bt = _paged(cfg, seq_lens, seed=11)
    cu, sl = _decode_meta(seq_lens)
    kc, vc = _rand_cache(cfg, NB, seed=12), _rand_cache(cfg, NB, seed=13)
    q = torch.empty(B, cfg["Hq"], cfg["D"], dtype=torch.bfloat16, device=DEV).uniform_(-1, 1, generator=_gen(14))
    out1 = _attend(cfg, q, kc, vc, cu, sl, bt, 1, max(seq_lens))
    perm = torch.randperm(NB, generator=_gen(15), device=DEV)
    kc2, vc2 = torch.empty_like(kc), torch.empty_like(vc)
    kc2[perm], vc2[perm] = kc, vc
    out2 = _attend(cfg, q, kc2, vc2, cu, sl, perm[bt.long()].to(torch.int32), 1, max(seq_lens))
    assert torch.equal(out1, out2)


def test_output_is_linear_in_the_values():
    cfg = C2
    B = cfg["B"]
    seq_lens = [cfg["ctx"]] * B
    NB, bt = _paged(cfg, seq_lens, seed=16)
    cu, sl = _decode_meta(seq_lens)
    kc = _rand_cache(cfg, NB, seed=17)
    va, vb = _rand_cache(cfg, NB, seed=18), _rand_cache(cfg, NB, seed=19, lo=-2.0, hi=2.0)
    q = torch.empty(B, cfg["Hq"], cfg["D"], dtype=torch.bfloat16, device=DEV).uniform_(-2, 2, generator=_gen(20))
    oa = _attend(cfg, q, kc, va, cu, sl, bt, 1, cfg["ctx"]).float()
    ob = _attend(cfg, q, kc, vb, cu, sl, bt, 1, cfg["ctx"]).float()
    oc = _attend(cfg, q, kc, (va.float() * 2 - vb.float()).to(torch.bfloat16), cu, sl, bt, 1, cfg["ctx"]).float()
    err = (oc - (2 * oa - ob)).abs().max().item()
    assert err <= 2e-2, f"max abs deviation from linearity {err}"


# ---------------------------------------------------------------------------------------------
# cache write at full size: a C3 step's 8192 tokens and a C2 decode step's 64 into a C2-sized cache
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("T", [64, 8192])
@pytest.mark.parametrize("fp8", [False, True])
def test_cache_write_reads_back_through_the_slot_mapping(T, fp8):
    from vllm_triton_backend_b200 import ops

    cfg = dict(C2, fp8=fp8)
    Hkv, D, bs = cfg["Hkv"], cfg["D"], cfg["bs"]
    NB = cfg["B"] * cfg["ctx"] // bs + 64
    sentinel = 0x55 if fp8 else 0x5555
    kc = torch.full((NB, bs, Hkv, D), sentinel, dtype=torch.uint8 if fp8 else torch.int16, device=DEV).view(_cache_dtype(cfg))
    vc = kc.clone()
    qkv = torch.empty(T, (32 + 2 * Hkv) * D, dtype=torch.bfloat16, device=DEV).uniform_(-600, 600, generator=_gen(21))
    key, value = qkv[:, 32 * D:(32 + Hkv) * D].view(T, Hkv, D), qkv[:, (32 + Hkv) * D:].view(T, Hkv, D)  # strided views
    slots = torch.randperm(NB * bs, generator=_gen(22), device=DEV)[:T]
    slots[T // 3::5] = -1  # padding tokens
    ks, vs = (2.0, 0.75) if fp8 else (1.0, 1.0)
    sc = [torch.tensor([s], dtype=torch.float32, device=DEV) for s in (ks, vs)]
    ops.reshape_and_cache_flash(key, value, kc, vc, slots, "fp8" if fp8 else "auto", sc[0], sc[1])
    torch.cuda.synchronize()
    live = slots >= 0
    for cache, src, s in ((kc, key, ks), (vc, value, vs)):
        flat = _bits(cache).view(NB * bs, Hkv, D)
        want = (src.float() / s).clamp(-448, 448).to(torch.float8_e4m3fn) if fp8 else src
        assert torch.equal(flat[slots[live]], _bits(want.contiguous())[live]), "written rows"
        untouched = torch.ones(NB * bs, dtype=torch.bool, device=DEV)
        untouched[slots[live]] = False
        assert (flat[untouched] == sentinel).all(), "rows no slot points at must keep their bytes"

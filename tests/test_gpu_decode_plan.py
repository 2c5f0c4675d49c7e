"""The split-KV work plan (b200attn_decode_plan; csrc/attn_decode.cu): the plan kernel's output is read back
and checked against a restatement in numpy - the static ranges and the pool chunks must tile the flattened
(sequence, head group, step) space exactly once, and every cursor must resolve its flat position - and a launch
that is handed a plan must produce the bits of a launch that plans for itself."""
import numpy as np
import pytest
import torch

from helpers import BF16_ATOL, BF16_RTOL, FP8_ATOL, assert_close, make_paged_inputs
from oracle.paged_attention import unified_attention_ref

pytestmark = pytest.mark.gpu

MAX_CTAS, ROUNDS = 160, 3  # csrc: kMaxDecodeCtas, kPoolRounds
MAX_CHUNKS = MAX_CTAS * ROUNDS


def _read_plan(plan):
    raw = plan.buf.cpu().numpy()
    hdr = raw[:32].view(np.int32)
    size, off = raw[32:48].view(np.int16).astype(int), raw[48:64].view(np.int16).astype(int)
    cur = raw[64:64 + 32 * (MAX_CTAS + MAX_CHUNKS)].view(np.int32).reshape(-1, 8)
    h = dict(total=int(hdr[0]), per=int(hdr[1]), per_static=int(hdr[2]), n_rounds=int(hdr[3]), n_chunks=int(hdr[4]),
             grid=int(hdr[5]), has_zero_rows=int(hdr[6]), snapped=int(hdr[7]))
    o = 64 + 32 * (MAX_CTAS + MAX_CHUNKS)
    h["bound"] = raw[o:o + 4 * (MAX_CTAS + 8)].view(np.int32).astype(int)  # Plan::bound (meaningful when snapped)
    return h, size, off, cur[:MAX_CTAS], cur[MAX_CTAS:]


def _check_plan(plan, q_lens, kv_lens, window, only_q1):
    h, size, off, cta, chunk = _read_plan(plan)
    # tokens per step and head groups are the kernel's choice: recover them from the plan itself
    some = [c for c in list(cta[:h["grid"]]) + list(chunk[:h["n_chunks"]]) if c[0] >= 0]
    if h["total"] == 0:
        assert not some
        return h
    tps = None
    for shift in range(4, 10):
        t = 1 << shift
        if all(int(c[5]) == -(-int(c[3]) // t) - (max(0, int(c[3]) - window) // t if window > 0 else 0) for c in some):
            tps = t
            break
    assert tps is not None, "no power-of-two step size explains the cursors"
    n = []
    first = []
    for ql, sl in zip(q_lens, kv_lens):
        if sl <= 0 or ql <= 0 or (only_q1 and ql != 1):
            n.append(0), first.append(0)
            continue
        f = (max(0, sl - window) // tps) if window > 0 else 0
        first.append(f), n.append(-(-sl // tps) - f)
    steps = sum(n)
    assert h["total"] % steps == 0
    HG = h["total"] // steps
    prefix = np.concatenate([[0], np.cumsum(n)])
    cu = np.concatenate([[0], np.cumsum(q_lens)])

    def resolve(f):
        b = int(np.searchsorted(prefix * HG, f, side="right") - 1)
        rem = f - prefix[b] * HG
        return b, rem // n[b], first[b] + rem % n[b]

    def check_cursor(c, f, live):
        if not live:
            assert int(c[0]) == -1, c
            return
        b, hg, st = resolve(f)
        assert tuple(int(x) for x in c[:3]) == (b, hg, st), (f, c)
        assert int(c[3]) == kv_lens[b] and int(c[4]) == first[b] and int(c[5]) == n[b]
        assert int(c[6]) == cu[b] and int(c[7]) == f

    grid, per, ps, T = h["grid"], h["per"], h["per_static"], h["n_rounds"]
    assert per == -(-h["total"] // grid) and 1 <= ps <= per and h["n_chunks"] == T * grid and 0 <= T <= ROUNDS
    # the rounds cut the tail of a range exactly: sizes never grow, offsets are their running sum
    assert sum(size[:T]) == per - ps and all(size[r] >= 1 for r in range(T))
    assert all(off[r] == sum(size[:r]) for r in range(T)) and all(size[r] <= size[r - 1] for r in range(1, T))
    covered = np.zeros(h["total"], dtype=np.int32)
    assert h["snapped"] == (1 if (T == 0 and per >= 4) else 0)
    if h["snapped"]:
        # boundaries one step inside an item have moved onto the item's edge; all others are where the division puts them
        bound = h["bound"]
        starts = set((prefix[:-1, None] * HG + np.arange(HG)[None, :] * np.array(n)[:, None]).flatten().tolist()) | {h["total"]}
        assert bound[0] == 0 and bound[grid] == h["total"]
        for c in range(grid + 1):
            f = min(c * per, h["total"])
            assert abs(bound[c] - f) <= 1 and (c == 0 or bound[c] >= bound[c - 1])
            if bound[c] != f:
                assert bound[c] in starts, (c, f, bound[c])
            if 0 < bound[c] < h["total"]:
                b_, hg_, st_ = resolve(bound[c])
                o_ = st_ - first[b_]
                assert n[b_] <= 2 or (o_ != 1 and o_ != n[b_] - 1) or (bound[c] in starts), (c, bound[c], o_, n[b_])
    for c in range(grid):
        r0, r1 = c * per, min((c + 1) * per, h["total"])
        if h["snapped"]:
            r0, r1 = int(h["bound"][c]), int(h["bound"][c + 1])
        b, e = r0, (r1 if h["snapped"] else min(r0 + ps, r1))
        check_cursor(cta[c], b, b < e)
        covered[b:max(b, e)] += 1
        for r in range(T):
            b = r0 + ps + off[r]
            e = min(b + size[r], r1)
            check_cursor(chunk[r * grid + c], b, b < e)
            covered[b:max(b, e)] += 1
    assert int(covered.min()) == 1 and int(covered.max()) == 1, "static parts and pool chunks must tile the step space once"
    zero = any(sl <= 0 and ql == 1 for ql, sl in zip(q_lens, kv_lens))
    assert bool(h["has_zero_rows"]) == zero
    return h


CASES = [
    ([(1, 4096)] * 64, 32, 8, False, 0),
    ([(1, 32768)] * 8, 8, 1, True, 0),
    ([(1, n) for n in [1, 15, 16, 17, 300, 0, 4095, 4097, 5, 640]], 32, 8, False, 0),
    ([(1, n) for n in [3000, 90, 5000, 12]], 16, 2, False, 256),
    ([(1, 700), (33, 900), (1, 64), (200, 200), (1, 0), (1, 5000)], 32, 8, False, 0),  # mixed: only the 1-token rows
    ([(1, 50 + 7 * i) for i in range(300)], 32, 8, False, 0),
    ([(1, 3)], 32, 8, False, 0),
    ([(1, 100 + (37 * i) % 4000) for i in range(1500)], 8, 8, False, 0),
]


@pytest.mark.parametrize("pool", [0, 3, 8])
@pytest.mark.parametrize("lens,Hq,Hkv,fp8,window", CASES)
def test_plan_tiles_the_step_space_exactly_once(lens, Hq, Hkv, fp8, window, pool):
    from vllm_triton_backend_b200 import _cabi, ops

    _cabi.set_tuning(0, pool, 0, 0)
    try:
        _plan_case(lens, Hq, Hkv, fp8, window, pool)
    finally:
        _cabi.set_tuning(0, 0, 0, 0)


def _plan_case(lens, Hq, Hkv, fp8, window, pool):
    from vllm_triton_backend_b200 import ops

    dev = "cuda"
    q_lens, kv_lens = [a for a, _ in lens], [b for _, b in lens]
    cu = torch.tensor(np.concatenate([[0], np.cumsum(q_lens)]), dtype=torch.int32, device=dev)
    sl = torch.tensor(kv_lens, dtype=torch.int32, device=dev)
    plan = ops.DecodePlan(dev, len(lens), Hq, Hkv, 128, 16, torch.bfloat16, torch.float8_e4m3fn if fp8 else torch.bfloat16, window)
    key = plan.build(cu, sl, max(q_lens))
    torch.cuda.synchronize()
    assert key != 0
    h = _check_plan(plan, q_lens, kv_lens, window, only_q1=max(q_lens) > 1)
    assert h["grid"] == min(torch.cuda.get_device_properties(0).multi_processor_count, MAX_CTAS)
    assert pool > 0 or h["n_rounds"] == 0


def _attend(inp, dev, plan=None, k_scale=None, Hq=32, max_q=1, window=0):
    from vllm_triton_backend_b200 import _cabi, ops

    q = inp["q"].to(dev)
    out = torch.full_like(q, float("nan"))
    sc = None if k_scale is None else torch.tensor([k_scale], dtype=torch.float32, device=dev)
    ops.unified_attention(
        q=q, k=inp["k_cache"].to(dev), v=inp["v_cache"].to(dev), out=out, cu_seqlens_q=inp["cu_seqlens_q"].to(dev),
        max_seqlen_q=max_q, seqused_k=inp["seq_lens"].to(dev), max_seqlen_k=int(inp["seq_lens"].max()),
        softmax_scale=q.shape[-1] ** -0.5, causal=True, window_size=(window - 1, 0) if window else (-1, -1),
        block_table=inp["block_table"].to(dev), softcap=0, q_descale=None, k_descale=sc, v_descale=sc, decode_plan=plan)
    torch.cuda.synchronize()
    return out.cpu(), _cabi.launch_count()


@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("lens", [[4096] * 16, [1, 17, 3000, 0, 640, 5000, 2048, 9], [8192] * 2])
def test_launch_with_a_plan_equals_the_self_planned_launch(lens, fp8):
    from vllm_triton_backend_b200 import _cabi, ops

    dev = "cuda"
    Hq, Hkv, D, bs = 32, 8, 128, 16
    kvd = torch.float8_e4m3fn if fp8 else None
    inp = make_paged_inputs([(1, n) for n in lens], Hq, Hkv, D, bs, dtype=torch.bfloat16, seed=11, kv_dtype=kvd, k_scale=0.5,
                            v_scale=0.5)
    _cabi.set_tuning(0, 0, 0, 0)  # static split (the default): the cut, hence the fp32 summation order, is a function of the plan alone
    try:
        _cabi.reset_launch_count()
        ref, n_self = _attend(inp, dev, None, 0.5 if fp8 else None)
        plan = ops.DecodePlan(dev, len(lens), Hq, Hkv, D, bs, torch.bfloat16, torch.float8_e4m3fn if fp8 else torch.bfloat16)
        assert plan.build(inp["cu_seqlens_q"].to(dev), inp["seq_lens"].to(dev), 1) != 0
        _cabi.reset_launch_count()
        out, n_plan = _attend(inp, dev, plan, 0.5 if fp8 else None)
    finally:
        _cabi.set_tuning(0, 0, 0, 0)
    assert n_self == 2 and n_plan == 1, (n_self, n_plan)  # plan kernel + decode vs decode alone
    assert torch.equal(out.view(torch.int16), ref.view(torch.int16))
    exp = unified_attention_ref(inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"],
                                inp["block_table"], D ** -0.5, k_scale=0.5 if fp8 else 1.0, v_scale=0.5 if fp8 else 1.0,
                                out_dtype=torch.float32)
    assert_close(out, exp, atol=FP8_ATOL if fp8 else BF16_ATOL, rtol=BF16_RTOL, what="planned decode")


def test_a_plan_for_another_geometry_is_ignored():
    """The key ties a plan to the launch geometry: one built for other head counts / another window must not be used."""
    from vllm_triton_backend_b200 import _cabi, ops

    dev = "cuda"
    lens = [3000, 17, 4096, 900]
    inp = make_paged_inputs([(1, n) for n in lens], 32, 8, 128, 16, dtype=torch.bfloat16, seed=5)
    wrong = ops.DecodePlan(dev, len(lens), 32, 8, 128, 16, torch.bfloat16, torch.bfloat16, sliding_window=128)
    assert wrong.build(inp["cu_seqlens_q"].to(dev), inp["seq_lens"].to(dev), 1) != 0
    _cabi.reset_launch_count()
    out, n = _attend(inp, dev, wrong)
    assert n == 2  # planned for itself
    exp = unified_attention_ref(inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"],
                                inp["block_table"], 128 ** -0.5, out_dtype=torch.float32)
    assert_close(out, exp, atol=BF16_ATOL, rtol=BF16_RTOL, what="ignored plan")


@pytest.mark.parametrize("pool", [0, 1, 3, 8, 12])
@pytest.mark.parametrize("fp8", [False, True])
def test_result_does_not_depend_on_the_pool_share(pool, fp8):
    """No pool (equal static split), the default and an oversized pool: same attention, different cuts."""
    from vllm_triton_backend_b200 import _cabi

    dev = "cuda"
    lens = [4096, 100, 1777, 16, 3000, 15, 2048, 640, 8000, 8000]
    kvd = torch.float8_e4m3fn if fp8 else None
    inp = make_paged_inputs([(1, n) for n in lens], 32, 8, 128, 16, dtype=torch.bfloat16, seed=3, kv_dtype=kvd, k_scale=0.37,
                            v_scale=0.21)
    try:
        _cabi.set_tuning(0, pool, 0, 1024 if pool == 12 else 0)
        out, _ = _attend(inp, dev, None, None)
        if fp8:
            from vllm_triton_backend_b200 import ops

            q = inp["q"].to(dev)
            out = torch.full_like(q, float("nan"))
            ks = torch.tensor([0.37], dtype=torch.float32, device=dev)
            vs = torch.tensor([0.21], dtype=torch.float32, device=dev)
            ops.unified_attention(q=q, k=inp["k_cache"].to(dev), v=inp["v_cache"].to(dev), out=out,
                                  cu_seqlens_q=inp["cu_seqlens_q"].to(dev), max_seqlen_q=1, seqused_k=inp["seq_lens"].to(dev),
                                  max_seqlen_k=max(lens), softmax_scale=128 ** -0.5, causal=True, window_size=(-1, -1),
                                  block_table=inp["block_table"].to(dev), softcap=0, q_descale=None, k_descale=ks, v_descale=vs)
            torch.cuda.synchronize()
            out = out.cpu()
    finally:
        _cabi.set_tuning(0, 0, 0, 0)
    exp = unified_attention_ref(inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"],
                                inp["block_table"], 128 ** -0.5, k_scale=0.37 if fp8 else 1.0, v_scale=0.21 if fp8 else 1.0,
                                out_dtype=torch.float32)
    assert_close(out, exp, atol=FP8_ATOL if fp8 else BF16_ATOL, rtol=BF16_RTOL, what=f"pool={pool}")


@pytest.mark.parametrize("pool", [0, 3])
def test_repeated_launches_leave_the_counters_clean(pool):
    """Race check by repetition (compute-sanitizer is closed on this pool, profiles/r02_compute_sanitizer_closed.txt):
    200 launches of a ragged batch on one workspace.  Without the pool (static split) the cut and the merge order
    are fixed, so every output must be the same bits.  With it, which CTA steals which chunk differs from launch to
    launch and with it the fp32 summation order: outputs agree to rounding, every one within tolerance of the
    oracle.  Either way tickets, claim words and the done counter end at zero."""
    from vllm_triton_backend_b200 import _cabi, ops

    dev = "cuda"
    lens = [5000, 33, 4096, 1, 700, 8191, 64, 2500, 12000, 300, 4097, 16]
    Hq, Hkv, D, bs = 32, 8, 128, 16
    inp = make_paged_inputs([(1, n) for n in lens], Hq, Hkv, D, bs, dtype=torch.bfloat16, seed=21, device=dev)
    ws = torch.zeros(ops.workspace_bytes(len(lens), len(lens), Hq, Hkv, D), dtype=torch.uint8, device=dev)
    outs = []
    try:
        _cabi.set_tuning(0, pool, 0, 0)
        for i in range(200):
            out = torch.empty_like(inp["q"])
            ops.unified_attention(q=inp["q"], k=inp["k_cache"], v=inp["v_cache"], out=out, cu_seqlens_q=inp["cu_seqlens_q"],
                                  max_seqlen_q=1, seqused_k=inp["seq_lens"], max_seqlen_k=max(lens), softmax_scale=D ** -0.5,
                                  causal=True, window_size=(-1, -1), block_table=inp["block_table"], softcap=0, q_descale=None,
                                  k_descale=None, v_descale=None, workspace=ws)
            outs.append(out)
        torch.cuda.synchronize()
    finally:
        _cabi.set_tuning(0, 0, 0, 0)
    exp = unified_attention_ref(inp["q"].cpu(), inp["k_cache"].cpu(), inp["v_cache"].cpu(), inp["cu_seqlens_q"].cpu(),
                                inp["seq_lens"].cpu(), inp["block_table"].cpu(), D ** -0.5, out_dtype=torch.float32)
    for o in outs[1:]:
        if pool == 0:
            assert torch.equal(o.view(torch.int16), outs[0].view(torch.int16))
        else:
            assert float((o.float() - outs[0].float()).abs().max()) <= 4e-3
    for o in (outs[0], outs[-1], outs[57]):
        assert_close(o, exp, atol=BF16_ATOL, rtol=BF16_RTOL, what=f"pool={pool}")
    assert int(ws[: 1 << 20].view(torch.int32).abs().sum()) == 0, "ticket / claim words not left at zero"

"""`softmax_lse`: every kernel's epilogue also writes ln sum_j exp(s_ij) of each (token, head) row (SURVEY.md 8f-3;
what `AttentionImpl.can_return_lse_for_decode` promises, vllm/vllm/attention/backends/abstract.py:130-153).  Checked
against the oracle's `return_lse`, and by the use it exists for: two partial attentions over disjoint halves of the
keys, combined with their LSEs, must equal attention over all keys."""
import pytest
import torch

from helpers import BF16_ATOL, BF16_RTOL, FP8_ATOL, assert_close, make_paged_inputs
from oracle.paged_attention import unified_attention_ref

pytestmark = pytest.mark.gpu
GENERIC, PREFILL, AUTO, PREFILL_TC = 1, 22, None, 4


def _run(inp, scale, force, window=0, softcap=0.0, sinks=None, k_scale=None, max_q=None, lens=None, stride=None):
    from vllm_triton_backend_b200 import _cabi, ops

    dev = "cuda"
    q = inp["q"].to(dev)
    T, Hq = q.shape[0], q.shape[1]
    out = torch.full_like(q, float("nan"))
    lse_buf = torch.full((T, stride or Hq), float("nan"), dtype=torch.float32, device=dev)
    lse = lse_buf[:, :Hq]
    sc = None if k_scale is None else torch.tensor([k_scale], dtype=torch.float32, device=dev)
    sl = inp["seq_lens"] if lens is None else torch.tensor(lens, dtype=torch.int32)
    mq = max_q if max_q is not None else int((inp["cu_seqlens_q"][1:] - inp["cu_seqlens_q"][:-1]).max())
    if force == PREFILL and mq == 1:
        force = 2
    ops.unified_attention(q=q, k=inp["k_cache"].to(dev), v=inp["v_cache"].to(dev), out=out, cu_seqlens_q=inp["cu_seqlens_q"].to(dev),
                          max_seqlen_q=mq, seqused_k=sl.to(dev), max_seqlen_k=int(sl.max()), softmax_scale=scale, causal=True,
                          window_size=(window - 1, 0) if window else (-1, -1), block_table=inp["block_table"].to(dev),
                          softcap=softcap, q_descale=None, k_descale=sc, v_descale=sc, force_selection=force,
                          sinks=None if sinks is None else sinks.to(dev), softmax_lse=lse)
    torch.cuda.synchronize()
    return out.cpu(), lse.cpu(), _cabi.last_kernel()


def _check_lse(lse, ref_lse, what):
    both_inf = torch.isinf(lse) & torch.isinf(ref_lse) & (lse < 0) & (ref_lse < 0)
    assert not torch.isnan(lse).any(), what
    err = torch.where(both_inf, torch.zeros_like(lse), (lse - ref_lse).abs())
    # logits come from bf16 operands with fp32 accumulation in both; ex2.approx / lg2.approx are the difference
    assert float(err.max()) <= 5e-3, f"{what}: lse max abs err {float(err.max()):.3e}"


@pytest.mark.parametrize("force,lens", [
    (AUTO, [(1, 700), (1, 33), (1, 4096), (1, 1), (1, 0), (1, 257)]),          # split-KV decode: direct and merged items
    (GENERIC, [(1, 700), (5, 18), (129, 463), (1, 0)]),
    (PREFILL, [(1, 700), (5, 18), (129, 463), (64, 64), (1, 0)]),               # mma.sync
    (PREFILL_TC, [(300, 300), (129, 463), (64, 1064), (513, 2049)]),             # tcgen05
    (AUTO, [(1, 700), (300, 300), (5, 18), (1, 77), (129, 463), (12, 900)]),     # mixed: three kernels, one LSE tensor
])
@pytest.mark.parametrize("feature", ["plain", "window", "softcap", "sinks", "fp8"])
@pytest.mark.parametrize("D", [128, 256])
def test_lse_matches_oracle(force, lens, feature, D):
    Hq, Hkv = 16, 4
    fp8 = feature == "fp8"
    if fp8 and force in (GENERIC, PREFILL_TC):
        pytest.skip("fp8 KV reaches the tcgen05 kernel through staging (the mixed AUTO case); the SIMT kernel adds nothing")
    inp = make_paged_inputs(lens, Hq, Hkv, D, 16, dtype=torch.bfloat16, seed=len(lens), kv_dtype=torch.float8_e4m3fn if fp8 else None,
                            k_scale=0.37, v_scale=0.37)
    window = 128 if feature == "window" else 0
    softcap = 20.0 if feature == "softcap" else 0.0
    sinks = torch.linspace(-2.0, 4.0, Hq) if feature == "sinks" else None
    scale = D ** -0.5 * (3.0 if softcap else 1.0)
    out, lse, kern = _run(inp, scale, force, window, softcap, sinks, 0.37 if fp8 else None, stride=Hq + 3)
    ref, ref_lse = unified_attention_ref(inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"],
                                         inp["block_table"], scale, sliding_window=window, softcap=softcap, sinks=sinks,
                                         k_scale=0.37 if fp8 else 1.0, v_scale=0.37 if fp8 else 1.0, out_dtype=torch.float32,
                                         return_lse=True)
    assert_close(out, ref, atol=FP8_ATOL if fp8 else BF16_ATOL, rtol=BF16_RTOL, what=f"{kern} {feature}")
    _check_lse(lse, ref_lse, f"{kern} {feature}")


@pytest.mark.parametrize("D,Hq,Hkv", [(128, 32, 8), (64, 8, 1), (128, 40, 8)])
def test_two_halves_combined_with_their_lse_equal_the_whole(D, Hq, Hkv):
    """Decode context parallelism in one process: keys [0, L/2) and [L/2, L) attended separately (the second half
    through the sliding window), then out = sum_i softmax_i(lse) * out_i."""
    L = [4096, 1000, 64, 8190]
    inp = make_paged_inputs([(1, n) for n in L], Hq, Hkv, D, 16, dtype=torch.bfloat16, seed=5)
    scale = D ** -0.5
    full, lse_full, _ = _run(inp, scale, AUTO)
    a, lse_a, _ = _run(inp, scale, AUTO, lens=[n // 2 for n in L])
    b, lse_b, _ = _run(inp, scale, AUTO, window=0, lens=L)  # placeholder to keep shapes; replaced below per sequence
    # the second half of sequence i is a window of L_i - L_i // 2 keys: windows differ per sequence, so run them one by one
    outs, lses = [], []
    for i, n in enumerate(L):
        sub = dict(inp)
        sub["q"] = inp["q"][i:i + 1]
        sub["cu_seqlens_q"] = torch.tensor([0, 1], dtype=torch.int32)
        sub["seq_lens"] = inp["seq_lens"][i:i + 1]
        sub["block_table"] = inp["block_table"][i:i + 1]
        o, l, _ = _run(sub, scale, AUTO, window=n - n // 2)
        outs.append(o), lses.append(l)
    b, lse_b = torch.cat(outs), torch.cat(lses)
    w = torch.softmax(torch.stack([lse_a, lse_b]), dim=0).unsqueeze(-1)
    comb = w[0] * a.float() + w[1] * b.float()
    assert_close(comb, full.float(), atol=BF16_ATOL, rtol=BF16_RTOL, what="LSE combine")
    assert float((torch.logaddexp(lse_a, lse_b) - lse_full).abs().max()) <= 5e-3


def test_impl_advertises_lse_and_refuses_dcp():
    from vllm_triton_backend_b200.backend.attn import B200AttentionImpl

    assert B200AttentionImpl.can_return_lse_for_decode is True

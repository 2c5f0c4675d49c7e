"""Full-CUDA-graph path (reference triton_attn.py:107, 120-128, 229-233):
the cache write and the attention call are captured once with seq_lens = 1 (what
`build_for_cudagraph_capture` hands the backend) and replayed after the metadata
tensors were rewritten in place.  Grids depend on host ints only, the library never
allocates or synchronises, so the replay must equal an eager call and the oracle."""
import pytest
import torch

from helpers import BF16_ATOL, BF16_RTOL, FP8_ATOL, assert_close, make_paged_inputs
from oracle.cache_write import reshape_and_cache_flash_ref
from oracle.paged_attention import unified_attention_ref

pytestmark = pytest.mark.gpu


class _Layer:
    def __init__(self, dev, k=1.0, v=1.0):
        self._k_scale = torch.tensor([k], dtype=torch.float32, device=dev)
        self._v_scale = torch.tensor([v], dtype=torch.float32, device=dev)


def _new_token_slots(inp, bs):
    slots = []
    for b, (ql, kl) in enumerate(zip(inp["q_lens"], inp["kv_lens"])):
        for pos in range(kl - ql, kl):
            slots.append(int(inp["block_table"][b, pos // bs]) * bs + pos % bs)
    return torch.tensor(slots, dtype=torch.int64)


@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("lens_capture,lens_replay", [
    # decode graph: B fixed (vLLM pads the batch to a capture size), lengths change
    ([(1, 1)] * 6, [(1, 700), (1, 33), (1, 4096), (1, 1), (1, 257), (1, 1500)]),
    # uniform multi-token graph (spec-decode style): T = B * q
    ([(4, 4)] * 3, [(4, 900), (4, 64), (4, 2049)]),
])
def test_captured_step_replays_with_new_metadata(fp8, lens_capture, lens_replay):
    from vllm_triton_backend_b200 import ops
    from vllm_triton_backend_b200.backend.attn import B200AttentionImpl, B200AttentionMetadata

    dev = "cuda"
    Hq, Hkv, D, bs = 32, 8, 128, 16
    scale = D ** -0.5
    ks, vs = (0.37, 0.61) if fp8 else (1.0, 1.0)
    inp = make_paged_inputs(lens_replay, Hq, Hkv, D, bs, seed=21, kv_dtype=torch.float8_e4m3fn if fp8 else None,
                            k_scale=ks, v_scale=vs)
    T, B = inp["q"].shape[0], len(lens_replay)
    assert sum(a for a, _ in lens_capture) == T and len(lens_capture) == B
    g = torch.Generator().manual_seed(5)
    key = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).bfloat16()
    value = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).bfloat16()
    slots = _new_token_slots(inp, bs)

    # oracle: write the new tokens, then attend
    kc_ref, vc_ref = inp["k_cache"].clone(), inp["v_cache"].clone()
    reshape_and_cache_flash_ref(key, value, kc_ref, vc_ref, slots, kv_cache_dtype="fp8" if fp8 else "auto",
                                k_scale=ks, v_scale=vs)
    ref = unified_attention_ref(inp["q"], kc_ref, vc_ref, inp["cu_seqlens_q"], inp["seq_lens"], inp["block_table"], scale,
                                k_scale=ks, v_scale=vs, out_dtype=torch.float32)

    # persistent device buffers, as vLLM's model runner keeps them
    if fp8:  # vLLM hands fp8 caches as uint8
        kv = torch.stack([inp["k_cache"].view(torch.uint8), inp["v_cache"].view(torch.uint8)]).to(dev)
    else:
        kv = torch.stack([inp["k_cache"], inp["v_cache"]]).to(dev)  # (2, NB, bs, Hkv, D)
    d_q, d_key, d_val = inp["q"].to(dev), key.to(dev), value.to(dev)
    d_out = torch.zeros_like(d_q)
    max_blocks = inp["block_table"].shape[1]
    d_bt = torch.zeros(B, max_blocks, dtype=torch.int32, device=dev)
    d_sl = torch.ones(B, dtype=torch.int32, device=dev)  # capture sees seq_lens = 1
    d_cu = inp["cu_seqlens_q"].to(dev)
    d_slots = torch.full((T,), -1, dtype=torch.int64, device=dev)  # capture sees padding slots
    max_q = max(a for a, _ in lens_replay)
    max_k = max(b for _, b in lens_replay)
    ws = torch.zeros(ops.workspace_bytes(T, B, Hq, Hkv, D), dtype=torch.uint8, device=dev)
    meta = B200AttentionMetadata(num_actual_tokens=T, max_query_len=max_q, query_start_loc=d_cu, max_seq_len=max_k,
                                 seq_lens=d_sl, block_table=d_bt, slot_mapping=d_slots, workspace=ws)
    impl = B200AttentionImpl(Hq, D, scale, Hkv, None, None, "fp8" if fp8 else "auto")
    layer = _Layer(dev, ks, vs)

    def step():
        impl.do_kv_cache_update(layer, d_key, d_val, kv, d_slots)
        impl.forward(layer, d_q, d_key, d_val, kv, meta, output=d_out)

    step()  # warm-up outside the capture (module load, function attributes)
    torch.cuda.synchronize()
    kv_before = kv.clone()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    assert torch.equal(kv, kv_before), "padding slots (-1) must not touch the cache"

    # the step's real metadata arrives; same tensors, new contents
    d_bt.copy_(inp["block_table"].to(dev))
    d_sl.copy_(inp["seq_lens"].to(dev))
    d_slots.copy_(slots.to(dev))
    d_out.fill_(float("nan"))
    graph.replay()
    torch.cuda.synchronize()
    out_graph = d_out.clone()
    as_bytes = (lambda t: t.view(torch.uint8)) if fp8 else (lambda t: t)
    assert torch.equal(kv[0].cpu(), as_bytes(kc_ref)), "key cache after replay != oracle (bit-exact required)"
    assert torch.equal(kv[1].cpu(), as_bytes(vc_ref)), "value cache after replay != oracle"
    assert_close(out_graph.cpu(), ref, atol=FP8_ATOL if fp8 else BF16_ATOL, rtol=BF16_RTOL, what="graph replay vs oracle")

    # replaying again (cache already holds the tokens) and an eager call give the same bits
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(d_out, out_graph), "second replay differs (workspace counters not self-cleaning?)"
    d_out.fill_(0)
    step()
    torch.cuda.synchronize()
    assert torch.equal(d_out, out_graph), "eager call differs from graph replay"

"""Debug helper: where does the tcgen05 prefill kernel differ from the oracle?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from helpers import make_paged_inputs
from oracle.paged_attention import unified_attention_ref
from vllm_triton_backend_b200 import ops

CASES = {
    "a": ([(1000, 1000)], 64, 8, 16, 72),
    "b": ([(100, 700), (31, 31)], 16, 1, 256, 272),
    "c": ([(512, 1536), (33, 33), (1, 700), (1, 9)], 32, 8, 32, 64),
    "d": ([(1000, 1000)], 8, 8, 16, 72),
    "e": ([(1000, 1000)], 8, 1, 16, 72),
}
for name in sys.argv[1:] or list(CASES):
    lens, Hq, Hkv, bs, seed = CASES[name]
    inp = make_paged_inputs(lens, Hq, Hkv, 128, bs, dtype=torch.bfloat16, seed=seed)
    scale = 128 ** -0.5
    ref = unified_attention_ref(inp["q"], inp["k_cache"], inp["v_cache"], inp["cu_seqlens_q"], inp["seq_lens"],
                                inp["block_table"], scale, out_dtype=torch.float32)
    dev = "cuda"
    q = inp["q"].to(dev); out = torch.full_like(q, float("nan"))
    for force in (4, 22):
        ops.unified_attention(q=q, k=inp["k_cache"].to(dev), v=inp["v_cache"].to(dev), out=out,
                              cu_seqlens_q=inp["cu_seqlens_q"].to(dev), max_seqlen_q=max(a for a, _ in lens),
                              seqused_k=inp["seq_lens"].to(dev), max_seqlen_k=max(b for _, b in lens), softmax_scale=scale,
                              causal=True, window_size=(-1, -1), block_table=inp["block_table"].to(dev), softcap=0,
                              q_descale=None, k_descale=None, v_descale=None, force_selection=force)
        torch.cuda.synchronize()
        err = (out.float().cpu() - ref).abs()           # [T, Hq, D]
        bad = err > (2e-2 + 1e-2 * ref.abs())
        rows = bad.any(-1)                                # [T, Hq]
        print(f"case {name} force {force}: max err {err.max():.4f}, bad elems {int(bad.sum())}, bad rows {int(rows.sum())}")
        if force == 4 and rows.any():
            toks = rows.any(-1).nonzero().flatten().tolist()
            print("  bad tokens:", toks[:80], "... total", len(toks))
            heads = rows.any(0).nonzero().flatten().tolist()
            print("  bad heads:", heads)
            t = toks[0]
            print("  first bad token row-max-err per head:", [round(float(x), 3) for x in err[t].max(-1).values])
            t = toks[-1]
            print("  last bad token row-max-err per head:", [round(float(x), 3) for x in err[t].max(-1).values])
            # per-token max error profile (every 8th token)
            prof = err.max(-1).values.max(-1).values
            print("  per-token max err (stride 16):", [round(float(x), 3) for x in prof[::16]])

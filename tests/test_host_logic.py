"""Host-side logic that needs no GPU: argument checks, byte/FLOP formulas,
head sharding, plugin class surface."""
import pytest
import torch

from vllm_triton_backend_b200 import ops, roofline
from vllm_triton_backend_b200.sharding import shard_heads


def test_c2_algorithmic_bytes_match_the_survey_figure():
    # 1 073 741 824 + 524 288 + 524 288 + 65 536 + 516
    assert roofline.decode_bytes([4096] * 64, 32, 8, 128, 16, 2) == 1_073_741_824 + 524_288 + 524_288 + 65_536 + 516


def test_c4_kv_bytes_and_prefill_flops():
    kv = roofline.decode_bytes([32768] * 64, 64, 8, 128, 16, 1)
    assert abs(kv - 4.295e9) / 4.295e9 < 2e-3
    # one full 8192-token causal prompt, Hq 32, D 128 = 0.550 TFLOP
    f = roofline.attention_flops([8192], [8192], 32, 128)
    assert abs(f - 0.550e12) / 0.550e12 < 2e-3
    assert roofline.cache_write_bytes(64, 8, 128, 2, 2) == 64 * 8 * 128 * 8 + 512


def test_ops_refuse_cpu_tensors():
    q = torch.zeros(1, 4, 64, dtype=torch.bfloat16)
    kc = torch.zeros(2, 16, 4, 64, dtype=torch.bfloat16)
    bt = torch.zeros(1, 1, dtype=torch.int32)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.unified_attention(q=q, k=kc, v=kc, out=torch.empty_like(q), cu_seqlens_q=torch.tensor([0, 1], dtype=torch.int32),
                              max_seqlen_q=1, seqused_k=torch.tensor([1], dtype=torch.int32), max_seqlen_k=1,
                              softmax_scale=1.0, causal=True, window_size=(-1, -1), block_table=bt, softcap=0,
                              q_descale=None, k_descale=None, v_descale=None)


def test_reference_assertions_come_first():
    q = torch.zeros(1, 4, 64, dtype=torch.bfloat16)
    with pytest.raises(AssertionError, match="Only causal"):
        ops.unified_attention(q, q, q, q, None, 1, None, 1, 1.0, False, (-1, -1), None, 0, None, None, None)
    with pytest.raises(AssertionError, match="Q scales"):
        ops.unified_attention(q, q, q, q, None, 1, None, 1, 1.0, True, (-1, -1), None, 0, torch.ones(1), None, None)


@pytest.mark.parametrize("world,hq,hkv,expect", [
    (1, 64, 8, [(0, 64, 0, 8)]),
    (2, 64, 8, [(0, 32, 0, 4), (32, 64, 4, 8)]),
    (8, 64, 8, [(8 * r, 8 * r + 8, r, r + 1) for r in range(8)]),
    (4, 32, 2, [(0, 8, 0, 1), (8, 16, 0, 1), (16, 24, 1, 2), (24, 32, 1, 2)]),  # KV heads replicated
])
def test_shard_heads(world, hq, hkv, expect):
    got = []
    for r in range(world):
        qs, ks = shard_heads(r, world, hq, hkv)
        got.append((qs.start, qs.stop, ks.start, ks.stop))
    assert got == expect


def test_shard_heads_rejects_uneven_splits():
    with pytest.raises(ValueError):
        shard_heads(0, 3, 32, 8)


def test_bench_helpers_and_kbench_namespace_contract():
    """bench.py's prefill section drives scripts/kbench.py with a SimpleNamespace: every attribute
    kbench.prefill_case reads without a default must be in it (a missing one broke the default bench run once)."""
    import ast
    import importlib.util
    import os
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    assert bench._percentiles([5, 1, 3, 2, 4], (0, 50, 100)) == [1, 3, 5]
    # mixed batch: 1e12 FLOP at 1000 TFLOP/s = 1 ms, 1e9 B at 1000 GB/s = 1 ms; measured 4 ms
    mb = bench._mixed_bounds(dict(flops=1e12, decode_row_kv_bytes=1e9, ms=4.0, decode_rows=3), dict(bf16_tflops=1000.0, hbm_gbs=1000.0))
    assert mb["bound_ms_overlapped"] == 1.0 and mb["bound_ms_serial"] == 2.0 and mb["frac_of_serial_bound"] == 0.5
    t = bench.committed_ncu_traffic()
    assert t is None or 1.0e9 < t < 1.2e9  # C2 decode kernel: ~1.075 GB algorithmic
    src = open(os.path.join(root, "bench.py")).read()
    ns = re.search(r"types\.SimpleNamespace\(([^)]*)\)", src).group(1)
    provided = set(re.findall(r"(\w+)=", ns))
    ksrc = open(os.path.join(root, "scripts", "kbench.py")).read()
    fn = next(n for n in ast.parse(ksrc).body if isinstance(n, ast.FunctionDef) and n.name == "prefill_case")
    needed = {n.attr for n in ast.walk(fn) if isinstance(n, ast.Attribute) and isinstance(n.value, ast.Name) and n.value.id == "args"}
    assert needed <= provided, f"bench.py's namespace lacks {needed - provided}"


def test_fp8_staging_bytes_covers_what_the_library_carves():
    # csrc/kv_stage.cu plan_kv_stage: index (256-aligned) + 2 planes of pages; any size is valid, 0 disables
    assert ops.fp8_staging_bytes(4, 0, 16, 8, 128) == 0
    n = ops.fp8_staging_bytes(4, 4096, 16, 8, 128)
    pages = 4096 // 16 + 4
    assert n >= 256 + 2 * pages * 16 * 8 * 128 * 2

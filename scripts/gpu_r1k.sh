#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_rope_cache.py -q -m gpu -p no:cacheprovider > gpurun_out/test_rope.log 2>&1; echo "[rope tests] exit=$? $(tail -1 gpurun_out/test_rope.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)|Error|assert" gpurun_out/test_rope.log | head -12 | cut -c1-250
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider -x --deselect tests/test_gpu_rope_cache.py > gpurun_out/test_all_r1k.log 2>&1; echo "[all other gpu tests] exit=$? $(tail -1 gpurun_out/test_all_r1k.log | cut -c1-200)"
K="timeout 300 python scripts/kbench.py decode"
: > gpurun_out/kbench_r1k.jsonl
for a in "--workload c4" "--workload c4 --hq 32 --hkv 4" "--workload c4 --hq 16 --hkv 2" "--workload c4 --hq 8 --hkv 1" "--workload c4 --ctx 4096" "--workload c2"; do
  $K $a >> gpurun_out/kbench_r1k.jsonl 2>> gpurun_out/kbench_r1k.err
done
python - <<'PY'
import json
for l in open("gpurun_out/kbench_r1k.jsonl"):
    d=json.loads(l); print(d["workload"], "hkv",d["hkv"],"ctx",d["ctx"], d["ms"], d["gbs"])
PY
python - <<'PY'
# K5 timing against the unfused pair (vLLM's compiled ops), 8192 prefill tokens and 64 decode tokens
import torch, sys, json
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from vllm_triton_backend_b200 import ops
import test_gpu_rope_cache as t
try:
    import vllm._C
    rotary, cache_op = torch.ops._C.rotary_embedding, torch.ops._C_cache_ops.reshape_and_cache_flash
except Exception as e:
    rotary = None
dev = "cuda"
for T in (64, 8192):
    Hq, Hkv, D, bs = 32, 8, 128, 16
    qkv, pos, cs, slots, kv = t._inputs(T, Hq, Hkv, D, D, bs, torch.bfloat16, False, seed=2, pad=1)
    qkv, pos, cs, slots, kv = qkv.to(dev), pos.to(dev), cs.to(dev), slots.to(dev), kv.to(dev)
    q, k, v = t._split(qkv, Hq, Hkv, D)
    sc = torch.ones(1, device=dev)
    def fused(): ops.rope_and_cache_flash(q, k, v, pos, cs, True, kv[0], kv[1], slots, "auto", 1.0, 1.0)
    def unfused():
        rotary(pos, q, k, D, cs, True); cache_op(k, v, kv[0], kv[1], slots, "auto", sc, sc)
    def ours_unfused_cache(): ops.reshape_and_cache_flash(k, v, kv[0], kv[1], slots, "auto", None, None)
    def tm(fn, n=200):
        for _ in range(10): fn()
        torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n): fn()
        e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n * 1e3
    bytes_ = T * ((Hq + Hkv) * D * 2 * 2 + Hkv * D * 2 + 2 * Hkv * D * 2)  # q,k read+write, v read, cache writes
    r = dict(kind="rope_cache", tokens=T, fused_us=round(tm(fused), 2), algorithmic_gbs=None)
    r["algorithmic_gbs"] = round(bytes_ / r["fused_us"] / 1e3, 1)
    if rotary is not None: r["vllm_rotary_plus_cache_us"] = round(tm(unfused), 2)
    print(json.dumps(r))
PY

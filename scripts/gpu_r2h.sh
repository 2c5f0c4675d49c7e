#!/bin/bash
# Round 2, call H: per-range claim words: owner continues into its own tail chunks, idle CTAs steal from the back.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode_plan.py tests/test_gpu_attention.py tests/test_gpu_cuda_graph.py tests/test_gpu_fuzz.py tests/test_gpu_full_size_numeric.py -x -q -p no:cacheprovider > gpurun_out/r2h_tests.log 2>&1; echo "[tests] exit=$? $(tail -4 gpurun_out/r2h_tests.log | tr '\n' ' ' | cut -c1-600)"
K="timeout 300 python scripts/kbench.py"
: > gpurun_out/r2h_kbench.jsonl
for a in "decode --workload c4 --hq 8 --hkv 1 --plan --trace" "decode --workload c4 --hq 8 --hkv 1 --plan --pool 100" \
         "decode --workload c4 --hq 8 --hkv 1 --plan --pool 2" "decode --workload c4 --hq 8 --hkv 1 --plan --pool 5" "decode --workload c4 --hq 8 --hkv 1 --plan --pool 8" \
         "decode --workload c4 --hq 8 --hkv 1 --plan --pool 3 --split 1024" "decode --workload c4 --hq 8 --hkv 1 --plan --pool 5 --split 1024" \
         "decode --workload c4 --plan --trace" "decode --workload c4 --plan --pool 100" "decode --workload c4 --hq 16 --hkv 2 --plan" "decode --workload c4 --hq 16 --hkv 2 --plan --pool 100" \
         "decode --workload c4 --hq 32 --hkv 4 --plan" "decode --workload c4 --hq 32 --hkv 4 --plan --pool 100" \
         "decode --workload c2 --plan --trace" "decode --workload c2 --plan --pool 100" "decode --workload c2 --plan --pool 2" "decode --workload c2 --plan --pool 5" \
         "decode --workload c2var --plan" "decode --workload c2var --plan --pool 100" "decode --workload c2 --batch 8 --plan" "decode --workload c2 --batch 8 --plan --pool 100" \
         "decode --workload c2 --ctx 640 --plan" "decode --workload c2 --ctx 640 --plan --pool 100" "decode --workload c2 --ctx 16384 --plan" "decode --workload c2 --ctx 16384 --plan --pool 100"; do
  $K $a >> gpurun_out/r2h_kbench.jsonl 2>> gpurun_out/r2h_kbench.err || echo "[kbench $a] failed"
done
python - <<'PY'
import json
for l in open("gpurun_out/r2h_kbench.jsonl"):
    d = json.loads(l)
    if d["kind"] == "decode_trace":
        print({k: (v["mean"], v["max"]) if isinstance(v, dict) else v for k, v in d.items() if k.endswith("_us") or k in ("runs_per_cta",)})
    else:
        print(d["workload"], "B", d["B"], "ctx", d["ctx"], "hkv", d["hkv"], "pool", d["ctas"], "split", d["split"], d["ms"], d["gbs"])
PY
tail -3 gpurun_out/r2h_kbench.err

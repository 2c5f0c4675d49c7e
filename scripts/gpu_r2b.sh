#!/bin/bash
# Round 2, call B: the plan-driven decode kernel (run descriptors + dynamic pool): tests first, then kbench.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_decode_plan.py -x -q -p no:cacheprovider > gpurun_out/r2b_plan_tests.log 2>&1; echo "[plan tests] exit=$? $(tail -3 gpurun_out/r2b_plan_tests.log | tr '\n' ' ')"
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider -x > gpurun_out/r2b_tests.log 2>&1; echo "[tests] exit=$? $(tail -3 gpurun_out/r2b_tests.log | tr '\n' ' ')"
K="timeout 300 python scripts/kbench.py"
: > gpurun_out/r2b_kbench.jsonl
for a in "decode --workload c4 --hq 8 --hkv 1 --plan --trace" "decode --workload c4 --hq 8 --hkv 1 --plan --pool 100 --trace" \
         "decode --workload c4 --hq 8 --hkv 1 --plan --pool 2" "decode --workload c4 --hq 8 --hkv 1 --plan --pool 4" "decode --workload c4 --hq 8 --hkv 1 --plan --pool 6" \
         "decode --workload c4 --hq 8 --hkv 1" "decode --workload c4 --hq 8 --hkv 1 --plan --graph" \
         "decode --workload c4 --plan --trace" "decode --workload c4 --plan --pool 100" "decode --workload c4 --hq 16 --hkv 2 --plan" "decode --workload c4 --hq 32 --hkv 4 --plan" \
         "decode --workload c2 --plan --trace" "decode --workload c2 --plan --pool 100" "decode --workload c2" "decode --workload c2 --plan --pool 5" \
         "decode --workload c2var --plan" "decode --workload c2var --plan --pool 100" "decode --workload c2 --batch 8 --plan --trace" "decode --workload c2 --batch 4 --ctx 1024 --plan" \
         "decode --workload c2 --ctx 640 --plan" "decode --workload c2 --ctx 16384 --plan"; do
  $K $a >> gpurun_out/r2b_kbench.jsonl 2>> gpurun_out/r2b_kbench.err || echo "[kbench $a] failed"
done
python - <<'PY'
import json
for l in open("gpurun_out/r2b_kbench.jsonl"):
    d = json.loads(l)
    if d["kind"] == "decode_trace":
        print("  trace span", d["kernel_span_us"], "first_run", d["first_run_us"]["mean"], "first_data", d["first_data_us"]["mean"], "stream", d["streaming_us"]["mean"], d["streaming_us"]["max"], "tail", d["tail_us"]["mean"], "end_skew", d["end_skew_us"]["mean"], "runs", d["runs_per_cta"]["mean"])
    else:
        print(d["workload"], "B", d["B"], "ctx", d["ctx"], "hkv", d["hkv"], "plan", d.get("plan"), "pool", d["ctas"], "graph", d.get("cuda_graph"), d["ms"], d["gbs"])
PY
tail -5 gpurun_out/r2b_kbench.err

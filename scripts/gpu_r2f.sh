#!/bin/bash
# Round 2, call F: next-chunk block-table prefetch; remaining new tests.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode_plan.py tests/test_gpu_scaled_fp8_quant.py tests/test_gpu_full_size_numeric.py tests/test_gpu_vs_triton.py -q -p no:cacheprovider > gpurun_out/r2f_tests.log 2>&1; echo "[tests] exit=$? $(tail -8 gpurun_out/r2f_tests.log | tr '\n' ' ' | cut -c1-1500)"
K="timeout 300 python scripts/kbench.py"
: > gpurun_out/r2f_kbench.jsonl
for a in "decode --workload c4 --hq 8 --hkv 1 --plan --trace" "decode --workload c4 --hq 8 --hkv 1 --plan --pool 100" \
         "decode --workload c4 --hq 8 --hkv 1 --plan --pool 2" "decode --workload c4 --hq 8 --hkv 1 --plan --pool 4" \
         "decode --workload c4 --hq 8 --hkv 1 --plan --pool 3 --split 1024" "decode --workload c4 --hq 8 --hkv 1 --plan --pool 3 --split 2048" \
         "decode --workload c4 --plan --trace" "decode --workload c4 --plan --pool 100" "decode --workload c4 --hq 16 --hkv 2 --plan" "decode --workload c4 --hq 16 --hkv 2 --plan --pool 100" \
         "decode --workload c2 --plan --trace" "decode --workload c2 --plan --pool 100" "decode --workload c2 --plan --pool 2" \
         "decode --workload c2var --plan" "decode --workload c2var --plan --pool 100"; do
  $K $a >> gpurun_out/r2f_kbench.jsonl 2>> gpurun_out/r2f_kbench.err || echo "[kbench $a] failed"
done
python - <<'PY'
import json
for l in open("gpurun_out/r2f_kbench.jsonl"):
    d = json.loads(l)
    if d["kind"] == "decode_trace":
        print("  trace span", d["kernel_span_us"], "first_run", d["first_run_us"]["mean"], "first_data", d["first_data_us"]["mean"], "stream", d["streaming_us"]["mean"], d["streaming_us"]["max"], "tail", d["tail_us"]["mean"], d["tail_us"]["max"], "end_skew", d["end_skew_us"]["mean"], "runs", d["runs_per_cta"]["mean"])
    else:
        print(d["workload"], "B", d["B"], "ctx", d["ctx"], "hkv", d["hkv"], "plan", d.get("plan"), "pool", d["ctas"], "split", d["split"], d["ms"], d["gbs"])
PY
tail -5 gpurun_out/r2f_kbench.err

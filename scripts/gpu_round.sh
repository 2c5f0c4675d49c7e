#!/bin/bash
# tests + kernel benches in one call (no ncu)
cd "$(dirname "$0")/.."
BENCH_STEPS=${BENCH_STEPS:-300} bash scripts/gpu_check.sh
K="python scripts/kbench.py"
: > gpurun_out/kbench.jsonl
for a in "decode --workload c2" "decode --workload c2var" "decode --workload c4" "decode --workload c4 --hq 8 --hkv 1" \
         "decode --workload c4 --hq 16 --hkv 2" "decode --workload c2 --batch 8" "decode --workload c2 --batch 1 --ctx 32768" \
         "prefill --shape full8k" "prefill --shape full2k" "prefill --shape chunk" "prefill --shape mixed" "prefill --shape full8k --force 22"; do
  timeout 300 $K $a >> gpurun_out/kbench.jsonl 2>> gpurun_out/kbench.err || echo "[kbench $a] failed: $(tail -2 gpurun_out/kbench.err)"
done
cat gpurun_out/kbench.jsonl | cut -c1-420

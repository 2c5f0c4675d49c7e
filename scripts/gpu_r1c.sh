#!/bin/bash
# Session re-entry check: full GPU parity suite, bench (both arms), kernel benches,
# ncu launch list + full captures (scripts/gpu_profile.sh), prefill shapes.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/test_all.log 2>&1; echo "[tests] exit=$? $(tail -1 gpurun_out/test_all.log)"
bash scripts/gpu_profile.sh
: > gpurun_out/kbench_tc.jsonl
for sh in full8k full2k chunk mixed; do
  timeout 300 python scripts/kbench.py prefill --shape $sh >> gpurun_out/kbench_tc.jsonl 2>> gpurun_out/kbench.err || echo "[kbench $sh] failed"
done
cut -c1-400 gpurun_out/kbench_tc.jsonl

#!/bin/bash
# Baseline re-check after container re-creation: the driver's own pytest line,
# smoke, default bench, and the weak decode/prefill shapes.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/test_all.log 2>&1; echo "[pytest gpu] exit=$? $(tail -1 gpurun_out/test_all.log)"
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "[smoke] exit=$? $(tail -1 gpurun_out/smoke.log)"
timeout 900 python bench.py > gpurun_out/bench.log 2>&1; echo "[bench] exit=$? $(tail -c 3000 gpurun_out/bench.log)"
K="python scripts/kbench.py"
: > gpurun_out/kbench.jsonl
for a in "decode --workload c2 --batch 8" "decode --workload c2 --batch 1 --ctx 32768" "decode --workload c2 --batch 1 --ctx 4096" "decode --workload c2 --batch 256 --ctx 1024" \
  "decode --workload c2var" "decode --workload c4 --hq 8 --hkv 1" "decode --workload c4 --hq 16 --hkv 2" "decode --workload c4 --hq 32 --hkv 4" "decode --workload c4" \
  "prefill --shape full8k" "prefill --shape full2k" "prefill --shape chunk" "prefill --shape mixed" "cache --tokens 8192" "cache --tokens 8192 --fp8" "cache --tokens 64"; do
  timeout 300 $K $a >> gpurun_out/kbench.jsonl 2>> gpurun_out/kbench.err || echo "[kbench $a] failed"
done
cut -c1-400 gpurun_out/kbench.jsonl

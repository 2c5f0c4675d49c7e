#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 python -m pytest "$@" -q -m gpu -p no:cacheprovider > gpurun_out/test_$name.log 2>&1; echo "[$name] exit=$? $(tail -1 gpurun_out/test_$name.log)"; }
run tc tests/test_gpu_attention.py -k "prefill_tc or mixed_batch or workspace_counters"
run decode tests/test_gpu_attention.py -k "decode"
timeout 900 python scripts/kbench.py sweep > gpurun_out/sweep.jsonl 2> gpurun_out/sweep.err; echo "[sweep] exit=$? lines=$(wc -l < gpurun_out/sweep.jsonl)"
cat gpurun_out/sweep.jsonl
for a in "prefill --shape full8k" "prefill --shape full2k" "prefill --shape chunk" "prefill --shape mixed"; do timeout 300 python scripts/kbench.py $a 2>&1 | cut -c1-330; done

#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
run() { name=$1; shift; timeout 900 python -m pytest "$@" -q -m gpu -p no:cacheprovider > gpurun_out/test_$name.log 2>&1; echo "[$name] exit=$? $(tail -1 gpurun_out/test_$name.log)"; }
[ -z "$NOTEST" ] && run tc tests/test_gpu_attention.py -k "prefill_tc or mixed_batch or workspace_counters or golden_tensor_core"
for v in ${TRACE_VARIANTS:-1}; do timeout 60 ./scripts/ubench/tc_trace $v | grep -E "kernel|per-CTA|period|^ *2[4-5] "; done
: > gpurun_out/kbench_tc.jsonl
for t in ${TUNES:-16}; do
for sh in ${SHAPES:-full8k full2k chunk mixed}; do
  timeout 300 python scripts/kbench.py prefill --shape $sh --tune $t >> gpurun_out/kbench_tc.jsonl 2>> gpurun_out/kbench.err || echo "[kbench $sh] failed"
done; done
cut -c1-400 gpurun_out/kbench_tc.jsonl | sed 's/"hq.*"fp8": false, //; s/"tokens_per_s.*//'

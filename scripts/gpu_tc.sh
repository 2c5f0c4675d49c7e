#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { name=$1; shift; timeout 900 python -m pytest "$@" -q -m gpu -p no:cacheprovider > gpurun_out/test_$name.log 2>&1; echo "[$name] exit=$? $(tail -1 gpurun_out/test_$name.log)"; }
run tc tests/test_gpu_attention.py -k "prefill_tc or mixed_batch or golden_tensor_core"
for a in "prefill --shape full8k" "prefill --shape full2k" "prefill --shape full1k" "prefill --shape chunk" "prefill --shape mixed" "prefill --shape full8k --hq 64 --hkv 8" "prefill --shape full8k --hq 8 --hkv 8" "prefill --shape full8k --bs 128"; do timeout 300 python scripts/kbench.py $a 2>&1 | cut -c1-330; done

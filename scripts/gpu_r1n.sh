#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -p no:cacheprovider -x > gpurun_out/test_all_r1n.log 2>&1; echo "[all gpu tests] exit=$? $(tail -1 gpurun_out/test_all_r1n.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)" gpurun_out/test_all_r1n.log | head
for a in "prefill --shape mixed" "prefill --shape full8k" "prefill --shape full2k" "prefill --shape chunk" "prefill --shape full8k --fp8" "prefill --shape mixed --fp8"; do
  timeout 300 python scripts/kbench.py $a 2>&1 | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["kind"], d.get("shape"), "fp8",d.get("fp8"), d["ms"], d.get("tflops"), d["kernel"])'
done

#!/bin/bash
# Round-1 step t: mma.sync prefill kernel (K3b) loader / softmax rework.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/test_r1t.log 2>&1; echo "[suite] exit=$? $(tail -1 gpurun_out/test_r1t.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)|^E  " gpurun_out/test_r1t.log | head -20 | cut -c1-250
K="python scripts/kbench.py"
for a in "prefill --shape full8k --force 22" "prefill --shape full8k --hq 64 --hkv 8 --force 22" "prefill --shape full2k --force 22" "prefill --shape full8k --force 22 --softcap 50" "prefill --shape full8k --fp8 --no-staging" "prefill --shape mixed" "prefill --shape mixed --fp8"; do
  timeout 300 $K $a 2>&1 | tail -1 | tee -a gpurun_out/kbench_r1t.jsonl | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["kind"], d.get("shape"), d.get("hq"), "fp8", d.get("fp8"), "cap", d.get("softcap"), d["ms"], d.get("tflops"), d["kernel"])'
done

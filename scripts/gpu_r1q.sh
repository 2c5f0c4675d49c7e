#!/bin/bash
# Round-1 step q: softcap + sinks inside the tcgen05 prefill kernel.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/test_r1q.log 2>&1; echo "[suite] exit=$? $(tail -1 gpurun_out/test_r1q.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)|^E  " gpurun_out/test_r1q.log | head -20 | cut -c1-250
for a in "prefill --shape full8k" "prefill --shape full8k --softcap 50" "prefill --shape full8k --softcap 50 --tune 96" "prefill --shape full8k --softcap 50 --force 22" "prefill --shape full8k --softcap 50 --impl triton" "prefill --shape mixed --softcap 50"; do
  timeout 300 python scripts/kbench.py $a 2>&1 | tail -1 | tee -a gpurun_out/kbench_r1q.jsonl | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["kind"], d.get("shape"), "softcap", d.get("softcap"), "tune", d.get("tune"), d["ms"], d.get("tflops"), d["kernel"])'
done

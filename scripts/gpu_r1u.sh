#!/bin/bash
# Round-1 step u: decode kernel requests the next item's query before merging the current one.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/test_r1u.log 2>&1; echo "[suite] exit=$? $(tail -1 gpurun_out/test_r1u.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)|^E  " gpurun_out/test_r1u.log | head -20 | cut -c1-250
K="python scripts/kbench.py"
for a in "decode --workload c2" "decode --workload c4" "decode --workload c2 --ctx 640" "decode --workload c2 --ctx 256" "decode --workload c2 --batch 4 --ctx 1024" "decode --workload c2var" "decode --workload c4 --hq 8 --hkv 1" "decode --workload c2 --ctx 640 --trace"; do
  timeout 300 $K $a 2>&1 | tail -2 | tee -a gpurun_out/kbench_r1u.jsonl | python -c '
import json,sys
for l in sys.stdin:
    d=json.loads(l)
    if d["kind"]=="decode": print(d["kind"], d.get("workload"), "B", d.get("B"), "ctx", d.get("ctx"), "hkv", d.get("hkv"), d["ms"], d.get("gbs"))
    else: print({k:d[k] for k in ("kernel_span_us","partition_us","first_data_us","streaming_us","tail_us") if k in d})'
done

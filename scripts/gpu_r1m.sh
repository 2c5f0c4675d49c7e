#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_rope_cache.py -q -m gpu -p no:cacheprovider -rs > gpurun_out/test_rope3.log 2>&1; echo "[rope tests] exit=$? $(tail -1 gpurun_out/test_rope3.log | cut -c1-200)"; grep -E "SKIPPED|FAILED|Error" gpurun_out/test_rope3.log | head -5 | cut -c1-300
for a in "prefill --shape mixed" "prefill --shape mixed --force 4" "decode --workload c2 --force 4" "decode --workload c2 --ctx 640 --force 4" "decode --workload c2var --force 4"; do
  timeout 300 python scripts/kbench.py $a 2>&1 | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["kind"], d.get("shape",d.get("workload")), "force",d.get("force"), d["ms"], d.get("tflops",d.get("gbs")), d["kernel"])'
done

#!/bin/bash
# Round-1 step r: programmatic dependent launch between the cache write and the decode kernel.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider -x > gpurun_out/test_r1r.log 2>&1; echo "[suite] exit=$? $(tail -1 gpurun_out/test_r1r.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)|^E  " gpurun_out/test_r1r.log | head -20 | cut -c1-250
for t in 0 512 0 512; do
  timeout 600 python bench.py --steps 1000 --no-cpu-baseline --no-prefill --tune $t > gpurun_out/bench_r1r_$t.log 2>&1; echo "[bench tune=$t] exit=$? $(tail -1 gpurun_out/bench_r1r_$t.log | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], "roof", d["roofline"]["achieved"])')"
done
for a in "decode --workload c2" "decode --workload c2 --split 512" "decode --workload c4" "decode --workload c4 --split 512" "decode --workload c2 --batch 4 --ctx 1024" "decode --workload c2 --batch 4 --ctx 1024 --split 512"; do
  timeout 300 python scripts/kbench.py $a 2>&1 | tail -1 | tee -a gpurun_out/kbench_r1r.jsonl | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["kind"], d.get("workload"), "split", d.get("split"), d["ms"], d.get("gbps"), d["kernel"])'
done

#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_vs_triton.py -q -m gpu -p no:cacheprovider -k "sinks" > gpurun_out/test_r1i.log 2>&1; echo "[sinks tests] exit=$? $(tail -1 gpurun_out/test_r1i.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)|Error" gpurun_out/test_r1i.log | head -8
timeout 900 python -m pytest tests -q -m gpu -p no:cacheprovider -x > gpurun_out/test_all_r1i.log 2>&1; echo "[all gpu tests] exit=$? $(tail -1 gpurun_out/test_all_r1i.log | cut -c1-200)"
K="timeout 300 python scripts/kbench.py decode"
for a in "--workload c2 --graph" "--workload c2 --ctx 640 --graph" "--workload c4 --graph"; do $K $a 2>/dev/null | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("graph", d["workload"], d["ctx"], d["ms"], d["gbs"])'; done
E="python scripts/e2e_vllm.py"
for b in b200 triton; do
timeout 600 $E --backend $b --layers 4 --num-prompts 16 --output-len 64 --kv-cache-dtype fp8 --save gpurun_out/e2e_${b}_fp8_small.json > gpurun_out/e2e_${b}_fp8_small.log 2>&1; echo "[e2e $b fp8 small] exit=$? $(grep '^{"metric"' gpurun_out/e2e_${b}_fp8_small.log | tail -1 | cut -c1-300)"; grep -n "Error" gpurun_out/e2e_${b}_fp8_small.log | tail -3 | cut -c1-300
done
$E --compare gpurun_out/e2e_b200_fp8_small.json gpurun_out/e2e_triton_fp8_small.json

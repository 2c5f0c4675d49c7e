#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
E="python scripts/e2e_vllm.py"
for b in b200 triton; do
timeout 900 $E --backend $b --kv-cache-dtype fp8 --save gpurun_out/e2e_${b}_fp8_full.json > gpurun_out/e2e_${b}_fp8_full.log 2>&1; echo "[e2e $b fp8 full] exit=$? $(grep '^{"metric"' gpurun_out/e2e_${b}_fp8_full.log | tail -1 | cut -c1-330)"; grep -n "Error" gpurun_out/e2e_${b}_fp8_full.log | tail -3 | cut -c1-300
done
$E --compare gpurun_out/e2e_b200_fp8_full.json gpurun_out/e2e_triton_fp8_full.json
# long-prompt variant: prefill-heavy (8 x 6144-token prompts, 32 out), bf16 and fp8 caches
for kv in auto fp8; do for b in b200 triton; do
timeout 900 $E --backend $b --kv-cache-dtype $kv --num-prompts 8 --input-len 6144 --output-len 32 --max-model-len 8192 --save gpurun_out/e2e_${b}_${kv}_long.json > gpurun_out/e2e_${b}_${kv}_long.log 2>&1; echo "[e2e $b $kv long-prompt] exit=$? $(grep '^{"metric"' gpurun_out/e2e_${b}_${kv}_long.log | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["value"], d["total_tokens_per_s"], d["seconds"])')"
done; $E --compare gpurun_out/e2e_b200_${kv}_long.json gpurun_out/e2e_triton_${kv}_long.json | cut -c1-260; done

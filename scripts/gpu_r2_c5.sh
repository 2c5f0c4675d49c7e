#!/bin/bash
# C5: end-to-end vLLM offline inference through the plugin (granite-3.1-8b shape, dummy weights, 64 x 512 in / 256 out).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TP=${1:-1}
KV=${3:-auto}   # kv_cache_dtype: auto | fp8 | fp8_per_token_head | int8_per_token_head
SUF=""; [ "$KV" != auto ] && SUF="_$KV"
OUT=gpurun_out/r2_c5_tp${TP}${SUF}.jsonl
: > $OUT
export MASTER_ADDR=127.0.0.1 VLLM_HOST_IP=127.0.0.1
for b in $2; do
  timeout 900 python scripts/e2e_vllm.py --backend $b --tp $TP --kv-cache-dtype $KV --save /tmp/c5_${b}.json > gpurun_out/r2_c5_${b}_tp${TP}${SUF}.log 2>&1
  echo "[c5 $b tp$TP] exit=$?"
  grep '^{"metric"' gpurun_out/r2_c5_${b}_tp${TP}${SUF}.log | tail -1 | tee -a $OUT | cut -c1-400
  tail -3 gpurun_out/r2_c5_${b}_tp${TP}${SUF}.log | cut -c1-300
done
for b in $2; do
  [ "$b" != "b200" ] && [ -f /tmp/c5_b200.json ] && [ -f /tmp/c5_${b}.json ] && python scripts/e2e_vllm.py --compare /tmp/c5_b200.json /tmp/c5_${b}.json | tee -a $OUT
done

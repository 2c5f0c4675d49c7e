#!/bin/bash
# C5 variant: continuous batching with ragged lengths (batch grows and shrinks), 8-layer model, b200 vs TRITON_ATTN tokens.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
E="python scripts/e2e_vllm.py"
A="--layers 8 --num-prompts 96 --max-num-seqs 24 --input-len 700 --output-len 160 --max-model-len 1024 --max-batched-tokens 2048 --ragged --repeats 1"
for b in b200 triton; do
  timeout 900 $E --backend $b $A --save gpurun_out/e2e_ragged_$b.json > gpurun_out/e2e_ragged_$b.log 2>&1; echo "[e2e ragged $b] exit=$? $(grep '^{"metric"' gpurun_out/e2e_ragged_$b.log | tail -1 | cut -c1-300)"
done
$E --compare gpurun_out/e2e_ragged_b200.json gpurun_out/e2e_ragged_triton.json | tail -1

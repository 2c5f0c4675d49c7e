#!/bin/bash
# C5: vLLM offline inference through the plugin; progressive (small eager -> small graphs -> full model, three backends).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
E="python scripts/e2e_vllm.py"
step() { name=$1; shift; timeout ${TMO:-900} $E "$@" --save gpurun_out/e2e_$name.json > gpurun_out/e2e_$name.log 2>&1; rc=$?; echo "[e2e $name] exit=$rc $(grep '^{"metric"' gpurun_out/e2e_$name.log | tail -1 | cut -c1-420)"; [ $rc -ne 0 ] && { grep -n "Error\|error:" gpurun_out/e2e_$name.log | tail -8 | cut -c1-300; }; return $rc; }
if [ -z "$SKIP_TRITON_TESTS" ]; then
timeout 900 python -m pytest tests/test_gpu_vs_triton.py -q -m gpu -p no:cacheprovider -k fp8 > gpurun_out/test_triton_fp8.log 2>&1; echo "[triton fp8 parity] exit=$? $(tail -1 gpurun_out/test_triton_fp8.log | cut -c1-200)"
fi
if [ -z "$SKIP_SMALL" ]; then
step b200_small_eager --backend b200 --layers 4 --eager --num-prompts 8 --output-len 32 || exit 0
step b200_small_graphs --backend b200 --layers 4 --num-prompts 16 --output-len 64 || exit 0
step triton_small_graphs --backend triton --layers 4 --num-prompts 16 --output-len 64
$E --compare gpurun_out/e2e_b200_small_graphs.json gpurun_out/e2e_triton_small_graphs.json
fi
[ -n "$SMALL_ONLY" ] && exit 0
step b200_full --backend b200 || exit 0
step triton_full --backend triton
step flash_full --backend flash
$E --compare gpurun_out/e2e_b200_full.json gpurun_out/e2e_triton_full.json
$E --compare gpurun_out/e2e_b200_full.json gpurun_out/e2e_flash_full.json

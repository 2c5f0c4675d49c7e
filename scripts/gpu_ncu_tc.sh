#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
Pf="python scripts/kbench.py prefill --shape full8k --tune ${TUNE:-16}"
$Pf > gpurun_out/plain_prefill.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attn_prefill_tc_kernel -s 3 -c 1 -o gpurun_out/prof_prefill_8k -f $Pf > gpurun_out/ncu_prefill.log 2>&1
echo "[ncu prefill] exit=$? $(ls -la gpurun_out/*.ncu-rep 2>/dev/null)"
cat gpurun_out/plain_prefill.log | cut -c1-300

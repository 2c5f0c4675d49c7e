#!/bin/bash
# Round-1 step s: bench after the event-sampling change; ncu capture of the mma.sync prefill kernel (K3b).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python bench.py --steps 1000 --cpu-seconds 3 > gpurun_out/bench_r1s.log 2>&1; echo "[bench] exit=$? $(tail -1 gpurun_out/bench_r1s.log | cut -c1-200)"
K="python scripts/kbench.py"
for a in "prefill --shape full8k --force 22" "prefill --shape full8k --hq 64 --hkv 8 --force 22" "prefill --shape full2k --force 22"; do
  timeout 300 $K $a 2>&1 | tail -1 | tee -a gpurun_out/kbench_r1s.jsonl | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["kind"], d.get("shape"), d.get("hq"), d["ms"], d.get("tflops"), d["kernel"])'
done
Pf="$K prefill --shape full8k --force 22"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_prefill_kernel -s 3 -c 1 -o gpurun_out/prof_prefill_mma_8k -f $Pf > gpurun_out/ncu_prefill_mma.log 2>&1
echo "[ncu K3b] exit=$? $(ls -la gpurun_out/prof_prefill_mma_8k.ncu-rep 2>/dev/null)"

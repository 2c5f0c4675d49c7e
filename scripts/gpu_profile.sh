#!/bin/bash
# One gpurun call: full default bench line, then the ncu passes of B200_PROFILING.md
# (launch list of the bench command; --set full captures of the decode and the
# tcgen05 prefill kernel).  Each ncu pass runs only after the same command exited 0
# without ncu.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
K="python scripts/kbench.py"
python bench.py > gpurun_out/bench_full.log 2>&1; echo "[bench] exit=$? $(tail -c 2500 gpurun_out/bench_full.log)"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "[bench ref] exit=$? $(tail -c 600 gpurun_out/bench_ref.log)"
: > gpurun_out/kbench.jsonl
for a in "decode --workload c4 --hq 8 --hkv 1 --bs 128" "decode --workload c4 --hq 8 --hkv 1" "decode --workload c4" "decode --workload c2" \
         "decode --workload c2 --flush" "cache --tokens 8192" "cache --tokens 8192 --fp8"; do
  timeout 300 $K $a >> gpurun_out/kbench.jsonl 2>> gpurun_out/kbench.err || echo "[kbench $a] failed"
done
cut -c1-330 gpurun_out/kbench.jsonl
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-prefill"
$B > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_bench.log 2>&1
echo "[ncu launches] exit=$?"
D="$K decode --workload c2"
$D > gpurun_out/plain_decode.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attn_decode_kernel -s 3 -c 2 -o gpurun_out/prof_decode_c2 -f $D > gpurun_out/ncu_decode.log 2>&1
echo "[ncu decode] exit=$?"
Pf="$K prefill --shape full8k"
$Pf > gpurun_out/plain_prefill.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attn_prefill_tc_kernel -s 3 -c 2 -o gpurun_out/prof_prefill_8k -f $Pf > gpurun_out/ncu_prefill.log 2>&1
echo "[ncu prefill] exit=$? $(ls -la gpurun_out/*.ncu-rep 2>/dev/null)"

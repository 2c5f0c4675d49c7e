#!/bin/bash
# Round-2 evidence pass (one GPU): smoke, bench (both arms), ncu launch list of the bench command, ncu --set full
# captures of the decode kernel (C2), the tcgen05 prefill kernels (D = 128, 64, 256) and the cache writes.
# Each ncu pass runs only after the same command exited 0 without ncu.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
K="python scripts/kbench.py"
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2f_smoke.log 2>&1; echo "[smoke] exit=$? $(tail -1 gpurun_out/r2f_smoke.log | cut -c1-200)"
timeout 900 python bench.py > gpurun_out/r2f_bench.log 2>&1; echo "[bench] exit=$?"; grep '^{"metric"' gpurun_out/r2f_bench.log | tail -1 > gpurun_out/r2f_bench.json; cut -c1-900 gpurun_out/r2f_bench.json
timeout 900 python bench.py --impl reference > gpurun_out/r2f_bench_ref.log 2>&1; echo "[bench ref] exit=$?"; grep '^{' gpurun_out/r2f_bench_ref.log | tail -1 > gpurun_out/r2f_bench_ref.json; cut -c1-600 gpurun_out/r2f_bench_ref.json
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-prefill --no-comparators --no-extras"
timeout 600 $B > gpurun_out/r2f_plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2f_launches.csv $B > gpurun_out/r2f_ncu_bench.log 2>&1
echo "[ncu launches] exit=$?"
cap() {  # name, kernel regex, kbench args
  local name=$1 re=$2; shift 2
  timeout 300 $K "$@" > gpurun_out/r2f_plain_$name.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$re -s 3 -c 1 -f -o gpurun_out/r2f_$name $K "$@" > gpurun_out/r2f_ncu_$name.log 2>&1
  echo "[ncu $name] exit=$? $(grep '^{' gpurun_out/r2f_plain_$name.log | tail -1 | cut -c1-260)"
}
cap decode_c2 attn_decode_kernel decode --workload c2 --plan
cap prefill_d128 attn_prefill_tc_kernel prefill --shape full8k
cap prefill_d64 attn_prefill_tc_kernel prefill --shape full8k --d 64
cap prefill_d256 attn_prefill_tc256_kernel prefill --shape full8k --d 256
cap cache_pth cache_quant_per_token_head cache --tokens 8192 --pth fp8
ls -la gpurun_out/*.ncu-rep

#!/bin/bash
# Round-2 closing pass (one GPU): whole GPU suite, smoke, bench (both arms), C4 on one GPU through the same bench
# (the like-for-like base of the N = 8 strong-scaling figure), ncu launch list + --set full capture of the decode kernel.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
K="python scripts/kbench.py"
timeout 2400 python -m pytest tests -x -q -m gpu -p no:cacheprovider 2>&1 | tail -4 > gpurun_out/r2g_tests.txt; echo "[tests] $(tail -1 gpurun_out/r2g_tests.txt)"
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2g_smoke.log 2>&1; echo "[smoke] exit=$? $(tail -1 gpurun_out/r2g_smoke.log | cut -c1-200)"
timeout 900 python bench.py > gpurun_out/r2g_bench.log 2>&1; echo "[bench] exit=$?"; grep '^{"metric"' gpurun_out/r2g_bench.log | tail -1 > gpurun_out/r2g_bench.json; cut -c1-300 gpurun_out/r2g_bench.json
timeout 900 python bench.py --impl reference > gpurun_out/r2g_bench_ref.log 2>&1; echo "[bench ref] exit=$?"; grep '^{' gpurun_out/r2g_bench_ref.log | tail -1 > gpurun_out/r2g_bench_ref.json; cut -c1-300 gpurun_out/r2g_bench_ref.json
for i in 1 2; do timeout 600 python bench.py --workload c4 --steps 300 --warmup 5 --no-prefill --no-cpu-baseline --no-comparators --no-extras 2>/dev/null | grep '^{"metric"' >> gpurun_out/r2g_bench_c4_n1.jsonl; done
cut -c1-200 gpurun_out/r2g_bench_c4_n1.jsonl
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-prefill --no-comparators --no-extras"
timeout 600 $B > gpurun_out/r2g_plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2g_launches.csv $B > gpurun_out/r2g_ncu_bench.log 2>&1
echo "[ncu launches] exit=$?"
timeout 300 $K decode --workload c2 --plan > gpurun_out/r2g_plain_decode.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode_kernel -s 3 -c 1 -f -o gpurun_out/r2g_decode_c2 $K decode --workload c2 --plan > gpurun_out/r2g_ncu_decode.log 2>&1
echo "[ncu decode] exit=$? $(grep '^{' gpurun_out/r2g_plain_decode.log | tail -1 | cut -c1-300)"

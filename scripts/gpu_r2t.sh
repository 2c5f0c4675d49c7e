#!/bin/bash
# round 2: per-(token, head) caches - tests, timings, and one ncu capture of the decode kernel in this mode beside the per-tensor one
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_per_token_head.py -x -q -m gpu -s 2>&1 | tail -40 > gpurun_out/r2t_tests.txt
tail -5 gpurun_out/r2t_tests.txt
{
for pth in fp8 int8; do
  timeout 300 python scripts/kbench.py decode --workload c4 --plan --pth $pth
  timeout 300 python scripts/kbench.py decode --workload c4 --plan --pth $pth --hq 8 --hkv 1
  timeout 300 python scripts/kbench.py decode --workload c2 --plan --pth $pth
done
} 2>&1 | grep '^{' > gpurun_out/r2t_kbench.jsonl
cat gpurun_out/r2t_kbench.jsonl | cut -c1-330
for v in fp8 int8 none; do
  extra="--pth $v"; [ $v = none ] && extra=""
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_decode_kernel -s 3 -c 1 -f -o gpurun_out/r2t_decode_c4_$v \
    python scripts/kbench.py decode --workload c4 --plan $extra > gpurun_out/r2t_ncu_$v.log 2>&1
done
ls -la gpurun_out/*.ncu-rep

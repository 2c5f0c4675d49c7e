#!/bin/bash
# round 2: per-token-head decode after slimming the per-step prefetch arithmetic; traces of short-context decode (the C5 regime)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_per_token_head.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -5 > gpurun_out/r2v_tests.txt
tail -3 gpurun_out/r2v_tests.txt
{
for pth in fp8 int8; do
  timeout 300 python scripts/kbench.py decode --workload c4 --plan --pth $pth
  timeout 300 python scripts/kbench.py decode --workload c4 --plan --pth $pth --hq 8 --hkv 1
  timeout 300 python scripts/kbench.py decode --workload c2 --plan --pth $pth
done
timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 640 --trace
timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 640 --graph
timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 768 --trace
timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 640 --hpc 2
timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 640 --hpc 4
timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 640 --hpc 8
timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 640 --hpc 1
timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 640 --stages 2
timeout 300 python scripts/kbench.py decode --workload c2 --ctx 640 --impl triton
} 2>&1 | grep '^{' > gpurun_out/r2v_kbench.jsonl
cut -c1-1200 gpurun_out/r2v_kbench.jsonl

#!/bin/bash
# round 2: per-token-head cache write with 32-bit index arithmetic; then C5 on one GPU with the final kernels
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_per_token_head.py -x -q -m gpu -k "cache_write or backend" 2>&1 | tail -3 > gpurun_out/r2z_tests.txt
tail -2 gpurun_out/r2z_tests.txt
{
for pth in fp8 int8; do
  timeout 300 python scripts/kbench.py cache --tokens 8192 --pth $pth
done
timeout 300 python scripts/kbench.py cache --tokens 65536 --pth fp8
timeout 300 python scripts/kbench.py cache --tokens 64 --pth fp8
} 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('cache', d['tokens'], d['per_token_head'], d['ms'], d['gbs'])"
bash scripts/gpu_r2_c5.sh 1 "b200 flash triton"

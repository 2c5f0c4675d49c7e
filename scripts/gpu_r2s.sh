#!/bin/bash
# round 2: per-(token, head) quantised caches - parity tests, then kernel timings beside vLLM's Triton kernels
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_per_token_head.py -x -q -m gpu -s 2>&1 | tail -40 > gpurun_out/r2s_tests.txt
tail -5 gpurun_out/r2s_tests.txt
{
for pth in fp8 int8; do
  timeout 300 python scripts/kbench.py decode --workload c4 --plan --pth $pth
  timeout 300 python scripts/kbench.py decode --workload c4 --plan --pth $pth --hq 8 --hkv 1
  timeout 300 python scripts/kbench.py decode --workload c2 --plan --pth $pth
  timeout 300 python scripts/kbench.py decode --workload c2 --pth $pth --impl triton
  timeout 300 python scripts/kbench.py cache --tokens 8192 --pth $pth
  timeout 300 python scripts/kbench.py cache --tokens 64 --pth $pth
  timeout 300 python scripts/kbench.py cache --tokens 8192 --pth $pth --impl triton
done
timeout 300 python scripts/kbench.py decode --workload c4 --plan
timeout 300 python scripts/kbench.py decode --workload c4 --plan --hq 8 --hkv 1
timeout 300 python scripts/kbench.py decode --workload c2 --plan --fp8
} 2>&1 | grep '^{' > gpurun_out/r2s_kbench.jsonl
cat gpurun_out/r2s_kbench.jsonl | cut -c1-400

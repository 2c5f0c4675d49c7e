#!/bin/bash
# round 2: vectorised run epilogue of the decode kernel - decode-related tests, then timings and traces
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 2000 python -m pytest tests/test_gpu_attention.py tests/test_gpu_decode_plan.py tests/test_gpu_lse.py tests/test_gpu_fuzz.py tests/test_gpu_per_token_head.py tests/test_gpu_guard_bands.py tests/test_gpu_full_size_numeric.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -5 > gpurun_out/r2w_tests.txt
tail -3 gpurun_out/r2w_tests.txt
{
timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 640 --trace
timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 640 --hpc 2
timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 768
timeout 300 python scripts/kbench.py decode --workload c2 --plan --trace
timeout 300 python scripts/kbench.py decode --workload c4 --plan --hq 8 --hkv 1 --trace
timeout 300 python scripts/kbench.py decode --workload c4 --plan
timeout 300 python scripts/kbench.py decode --workload c2var --plan
timeout 300 python scripts/kbench.py decode --workload c2 --plan --batch 8
timeout 300 python scripts/kbench.py decode --workload c4 --plan --pth fp8
timeout 300 python scripts/kbench.py decode --workload c4 --plan --pth int8
} 2>&1 | grep '^{' > gpurun_out/r2w_kbench.jsonl
cut -c1-1500 gpurun_out/r2w_kbench.jsonl

#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_attention.py tests/test_gpu_vs_triton.py tests/test_gpu_fuzz.py -x -q -p no:cacheprovider > gpurun_out/r2p_tests.log 2>&1; echo "[tests] exit=$? $(tail -6 gpurun_out/r2p_tests.log | tr '\n' ' ' | cut -c1-1200)"

#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider -x > gpurun_out/r2r_tests.log 2>&1; echo "[tests] exit=$? $(tail -4 gpurun_out/r2r_tests.log | tr '\n' ' ' | cut -c1-900)"

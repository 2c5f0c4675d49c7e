#!/bin/bash
# round 2: the no-allocation entry point alone in a fresh process, then the whole GPU suite (which collects it last)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -15 > gpurun_out/r2u_alone.txt
cat gpurun_out/r2u_alone.txt
timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 > gpurun_out/r2u_all.txt
cat gpurun_out/r2u_all.txt

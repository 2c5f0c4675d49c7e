#!/bin/bash
# Multi-GPU: the driver's launch line for N ranks (C4, KV heads sharded, strong scaling) + shard verification.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}
nvidia-smi -L | head -8
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $T bench.py --gpus $N --steps 300 --warmup 5 --no-prefill > gpurun_out/r2_n${N}_bench.log 2>&1; echo "[bench n$N] exit=$?"; grep "^{" gpurun_out/r2_n${N}_bench.log | tail -1 | cut -c1-3000
timeout 300 $T scripts/verify_head_shards.py > gpurun_out/r2_n${N}_verify.log 2>&1; echo "[verify n$N] exit=$?"; grep "^{" gpurun_out/r2_n${N}_verify.log | tail -1 | cut -c1-600
timeout 300 $T bench.py --gpus $N --impl reference --steps 2 --warmup 1 > gpurun_out/r2_n${N}_ref.log 2>&1; echo "[ref n$N] exit=$?"; grep "^{" gpurun_out/r2_n${N}_ref.log | tail -1 | cut -c1-900

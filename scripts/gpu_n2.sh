#!/bin/bash
# run with: gpurun --gpus N -- 'NG=N bash scripts/gpu_n2.sh'
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
NG=${NG:-2}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $T scripts/verify_head_shards.py > gpurun_out/verify_n$NG.log 2>&1; echo "[verify n=$NG] exit=$? $(grep verify_head_shards gpurun_out/verify_n$NG.log | cut -c1-600)"
timeout 900 $T bench.py --gpus $NG --no-prefill --no-cpu-baseline > gpurun_out/bench_n$NG.log 2>&1; echo "[bench c2 n=$NG] exit=$? $(tail -1 gpurun_out/bench_n$NG.log | cut -c1-330)"
timeout 900 $T bench.py --gpus $NG --workload c4 --no-prefill --no-cpu-baseline > gpurun_out/bench_c4_n$NG.log 2>&1; echo "[bench c4 n=$NG] exit=$? $(tail -1 gpurun_out/bench_c4_n$NG.log | cut -c1-330)"
timeout 600 $T bench.py --impl reference --gpus $NG --steps 2 --warmup 1 > gpurun_out/bench_ref_n$NG.log 2>&1; echo "[bench ref n=$NG] exit=$? $(tail -1 gpurun_out/bench_ref_n$NG.log | cut -c1-200)"

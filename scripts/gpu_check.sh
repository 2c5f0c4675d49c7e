#!/bin/bash
# One gpurun call: GPU parity tests by group (separate processes so a faulting
# kernel cannot poison the rest), smoke, short bench.  Logs land in gpurun_out/.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
run() { # name, pytest args...
  name=$1; shift
  timeout 900 python -m pytest "$@" -q -m gpu -p no:cacheprovider > gpurun_out/test_$name.log 2>&1
  echo "[$name] exit=$? $(tail -1 gpurun_out/test_$name.log)"
}
run cache tests/test_gpu_cache_write.py
run generic tests/test_gpu_attention.py -k "generic or odd_head or argument_checks"
run prefill tests/test_gpu_attention.py -k "prefill and not through_prefill and not prefill_tc"
run tc tests/test_gpu_attention.py -k "prefill_tc or mixed_batch or workspace_counters"
run decode tests/test_gpu_attention.py -k "decode"
run golden_tc tests/test_gpu_attention.py -k "golden_tensor_core"
run surface tests/test_backend_surface.py
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "[smoke] exit=$? $(tail -1 gpurun_out/smoke.log)"
timeout 900 python bench.py --steps ${BENCH_STEPS:-300} --warmup 10 ${BENCH_ARGS:-} > gpurun_out/bench.log 2>&1; echo "[bench] exit=$? $(tail -c 1500 gpurun_out/bench.log)"

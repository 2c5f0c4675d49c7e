#!/bin/bash
# Round-1 step p: fused FP8 output quantisation + chunked local attention tests, full GPU suite, perf sanity.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_output_quant.py tests/test_gpu_local_attention.py -q -m gpu -p no:cacheprovider > gpurun_out/test_r1p_new.log 2>&1; echo "[new tests] exit=$? $(tail -1 gpurun_out/test_r1p_new.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)|Error|assert " gpurun_out/test_r1p_new.log | head -12 | cut -c1-300
timeout 1200 python -m pytest tests -q -m gpu -p no:cacheprovider --deselect tests/test_gpu_output_quant.py --deselect tests/test_gpu_local_attention.py > gpurun_out/test_r1p_all.log 2>&1; echo "[suite] exit=$? $(tail -1 gpurun_out/test_r1p_all.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)" gpurun_out/test_r1p_all.log | head -8 | cut -c1-250
for a in "decode --workload c2" "decode --workload c4" "prefill --shape full8k" "prefill --shape mixed"; do
  timeout 300 python scripts/kbench.py $a 2>&1 | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["kind"], d.get("workload"), d.get("shape"), d["ms"], d.get("gbps"), d.get("tflops"), d["kernel"])'
done
timeout 600 python bench.py --steps 300 --cpu-seconds 3 > gpurun_out/bench_r1p.log 2>&1; echo "[bench] exit=$? $(tail -1 gpurun_out/bench_r1p.log | cut -c1-300)"

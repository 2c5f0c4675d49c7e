#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_vs_triton.py tests/test_gpu_cuda_graph.py -q -m gpu -p no:cacheprovider -x > gpurun_out/test_r1o.log 2>&1; echo "[tests] exit=$? $(tail -1 gpurun_out/test_r1o.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)|Error|assert " gpurun_out/test_r1o.log | head -8 | cut -c1-250
for a in "prefill --shape full8k --fp8" "prefill --shape full8k --fp8 --no-staging" "prefill --shape mixed --fp8" "prefill --shape chunk --fp8" "prefill --shape full2k --fp8" "prefill --shape full8k" "prefill --shape mixed"; do
  timeout 300 python scripts/kbench.py $a 2>&1 | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["kind"], d.get("shape"), "fp8",d.get("fp8"), d["ms"], d.get("tflops"), d["kernel"])'
done
timeout 600 python bench.py --steps 200 --cpu-seconds 3 > gpurun_out/bench_r1o.log 2>&1; echo "[bench] exit=$? $(tail -1 gpurun_out/bench_r1o.log | cut -c1-200)"

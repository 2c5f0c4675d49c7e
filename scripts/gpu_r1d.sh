#!/bin/bash
# Triton-reference parity + timings beside ours, and a first C5 plumbing check (few layers).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_vs_triton.py -q -m gpu -p no:cacheprovider > gpurun_out/test_triton.log 2>&1; echo "[triton parity] exit=$? $(tail -3 gpurun_out/test_triton.log | cut -c1-300)"
: > gpurun_out/kbench_triton.jsonl
for a in "decode --workload c2" "decode --workload c4" "decode --workload c4 --hq 8 --hkv 1" "prefill --shape full8k" "prefill --shape mixed" "prefill --shape chunk"; do
  for impl in ours triton; do
    timeout 300 python scripts/kbench.py $a --impl $impl >> gpurun_out/kbench_triton.jsonl 2>> gpurun_out/kbench_triton.err || echo "[kbench $a $impl] failed: $(tail -2 gpurun_out/kbench_triton.err | cut -c1-300)"
  done
done
cut -c1-120 gpurun_out/kbench_triton.jsonl | sed 's/"B".*"fp8"/"fp8"/' ; python - <<'PY'
import json
for l in open("gpurun_out/kbench_triton.jsonl"):
    d=json.loads(l); print(d["kind"], d.get("workload",d.get("shape")), d.get("hkv"), d["impl"], d["ms"], d.get("gbs",d.get("tflops")))
PY
L=${E2E_LAYERS:-4}
timeout 900 python scripts/e2e_vllm.py --backend b200 --layers $L --eager --num-prompts 8 --output-len 32 --save gpurun_out/e2e_b200_small.json > gpurun_out/e2e_b200_small.log 2>&1; echo "[e2e b200 small eager] exit=$? $(tail -1 gpurun_out/e2e_b200_small.log | cut -c1-700)"
grep -n "Error\|error\|Traceback" gpurun_out/e2e_b200_small.log | head -20

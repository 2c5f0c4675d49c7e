#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_cuda_graph.py tests/test_gpu_fuzz.py tests/test_gpu_full_size_numeric.py tests/test_gpu_full_size_properties.py tests/test_gpu_local_attention.py tests/test_gpu_output_quant.py -x -q -p no:cacheprovider > gpurun_out/r2l_tests.log 2>&1; echo "[tests] exit=$? $(tail -4 gpurun_out/r2l_tests.log | tr '\n' ' ' | cut -c1-600)"
K="timeout 300 python scripts/kbench.py"
: > gpurun_out/r2l_kbench.jsonl
for h in 0 2 4 8; do for c in 0 24 32 40 48; do
  $K prefill --shape mixed --hpc $h --tune $((c<<13)) >> gpurun_out/r2l_kbench.jsonl 2>> gpurun_out/r2l_kbench.err || echo "[kbench mixed] failed"
done; done
$K prefill --shape mixed --tune 4096 >> gpurun_out/r2l_kbench.jsonl 2>> gpurun_out/r2l_kbench.err
$K prefill --shape mixed --fp8 >> gpurun_out/r2l_kbench.jsonl 2>> gpurun_out/r2l_kbench.err
for h in 1 2 4 8; do $K decode --workload c2 --plan --hpc $h --batch 32 --ctx 4096 >> gpurun_out/r2l_kbench.jsonl 2>> gpurun_out/r2l_kbench.err; done
python - <<'PY'
import json
for l in open("gpurun_out/r2l_kbench.jsonl"):
    d = json.loads(l)
    if d["kind"] == "prefill": print(d["shape"], "fp8", d["fp8"], "hpc", d.get("hpc"), "ctas", (d["tune"] >> 13) & 255, "tune", d["tune"] & 8191, d["ms"], d["tflops"])
    else: print("decode B", d["B"], "hpc", d["hpc"], d["ms"], d["gbs"])
PY
tail -3 gpurun_out/r2l_kbench.err

#!/bin/bash
# Round 2, call K: mixed batch (C3): single-token rows beside the prompts on a second stream.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_cuda_graph.py tests/test_gpu_fuzz.py tests/test_gpu_full_size_numeric.py tests/test_gpu_full_size_properties.py tests/test_gpu_local_attention.py tests/test_gpu_output_quant.py -x -q -p no:cacheprovider > gpurun_out/r2k_tests.log 2>&1; echo "[tests] exit=$? $(tail -4 gpurun_out/r2k_tests.log | tr '\n' ' ' | cut -c1-600)"
K="timeout 300 python scripts/kbench.py"
: > gpurun_out/r2k_kbench.jsonl
for t in 4096 0 $((16<<13)) $((24<<13)) $((32<<13)) $((40<<13)) $((56<<13)); do
  $K prefill --shape mixed --tune $t >> gpurun_out/r2k_kbench.jsonl 2>> gpurun_out/r2k_kbench.err || echo "[kbench mixed $t] failed"
done
$K prefill --shape mixed --fp8 --tune 4096 >> gpurun_out/r2k_kbench.jsonl 2>> gpurun_out/r2k_kbench.err
$K prefill --shape mixed --fp8 >> gpurun_out/r2k_kbench.jsonl 2>> gpurun_out/r2k_kbench.err
$K prefill --shape full8k >> gpurun_out/r2k_kbench.jsonl 2>> gpurun_out/r2k_kbench.err
python - <<'PY'
import json
for l in open("gpurun_out/r2k_kbench.jsonl"):
    d = json.loads(l)
    print(d["shape"], "fp8", d["fp8"], "tune", d["tune"], "ctas", (d["tune"] >> 13) & 255, d["ms"], d["tflops"], d["kernel"])
PY
tail -3 gpurun_out/r2k_kbench.err

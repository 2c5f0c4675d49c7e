#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_attention.py tests/test_gpu_fuzz.py tests/test_gpu_vs_triton.py tests/test_gpu_output_quant.py -x -q -p no:cacheprovider > gpurun_out/r2n_tests.log 2>&1; echo "[tests] exit=$? $(tail -4 gpurun_out/r2n_tests.log | tr '\n' ' ' | cut -c1-800)"
K="timeout 300 python scripts/kbench.py"
: > gpurun_out/r2n_kbench.jsonl
for a in "prefill --shape full8k --d 64 --hq 64 --hkv 8" "prefill --shape full8k --d 64 --hq 64 --hkv 8 --tune 16" "prefill --shape full8k --d 64 --hq 64 --hkv 8 --tune 32" "prefill --shape full8k --d 64 --hq 64 --hkv 8 --tune 48" "prefill --shape full8k --d 64 --hq 64 --hkv 8 --tune 64" \
         "prefill --shape full8k --d 64 --hq 64 --hkv 8 --force 22" "prefill --shape full8k --hq 40 --hkv 8" "prefill --shape full8k --hq 40 --hkv 8 --force 22" "prefill --shape full8k" "prefill --shape mixed --d 64 --hq 64 --hkv 8"; do
  $K $a >> gpurun_out/r2n_kbench.jsonl 2>> gpurun_out/r2n_kbench.err || echo "[kbench $a] failed"
done
python - <<'PY'
import json
for l in open("gpurun_out/r2n_kbench.jsonl"):
    d = json.loads(l)
    print(d["shape"], "D", d["D"], "hq", d["hq"], "hkv", d["hkv"], "tune", d["tune"], "force", d["force"], d["ms"], d["tflops"], d["kernel"])
PY
tail -3 gpurun_out/r2n_kbench.err

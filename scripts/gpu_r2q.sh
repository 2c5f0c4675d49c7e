#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_attention.py -x -q -p no:cacheprovider -k "prefill_tc and 256" > gpurun_out/r2q_tests256.log 2>&1; echo "[tests 256] exit=$? $(tail -6 gpurun_out/r2q_tests256.log | tr '\n' ' ' | cut -c1-900)"
timeout 1200 python -m pytest tests/test_gpu_attention.py tests/test_gpu_lse.py tests/test_gpu_fuzz.py tests/test_gpu_output_quant.py -x -q -p no:cacheprovider > gpurun_out/r2q_tests.log 2>&1; echo "[tests] exit=$? $(tail -6 gpurun_out/r2q_tests.log | tr '\n' ' ' | cut -c1-900)"
K="timeout 300 python scripts/kbench.py"
: > gpurun_out/r2q_kbench.jsonl
for a in "prefill --shape full8k --d 256 --hq 16 --hkv 8" "prefill --shape full8k --d 256 --hq 16 --hkv 8 --force 22" "prefill --shape full8k --d 256 --hq 32 --hkv 16 --softcap 50" "prefill --shape mixed --d 256 --hq 16 --hkv 8" "prefill --shape full2k --d 256 --hq 16 --hkv 8"; do
  $K $a >> gpurun_out/r2q_kbench.jsonl 2>> gpurun_out/r2q_kbench.err || echo "[kbench $a] failed"
done
python - <<'PY'
import json
for l in open("gpurun_out/r2q_kbench.jsonl"):
    d = json.loads(l)
    print(d["shape"], "D", d["D"], "hq", d["hq"], "hkv", d["hkv"], "cap", d["softcap"], "force", d["force"], d["ms"], d["tflops"], d["kernel"])
PY
tail -3 gpurun_out/r2q_kbench.err

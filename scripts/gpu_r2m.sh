#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider -x > gpurun_out/r2m_tests.log 2>&1; echo "[tests] exit=$? $(tail -4 gpurun_out/r2m_tests.log | tr '\n' ' ' | cut -c1-600)"
K="timeout 300 python scripts/kbench.py"
: > gpurun_out/r2m_kbench.jsonl
for a in "prefill --shape mixed" "prefill --shape mixed --tune 4096" "prefill --shape mixed_prompts" "prefill --shape mixed --fp8" "prefill --shape mixed --fp8 --tune 4096" "prefill --shape mixed_prompts --fp8" "prefill --shape full8k" "prefill --shape chunk" "prefill --shape full2k"; do
  $K $a >> gpurun_out/r2m_kbench.jsonl 2>> gpurun_out/r2m_kbench.err || echo "[kbench $a] failed"
done
python - <<'PY'
import json
for l in open("gpurun_out/r2m_kbench.jsonl"):
    d = json.loads(l)
    print(d["shape"], "fp8", d["fp8"], "tune", d["tune"], d["ms"], d["tflops"], d["kernel"])
PY
python bench.py --steps 300 --warmup 5 > gpurun_out/r2m_bench.log 2>&1; echo "[bench] exit=$?"; tail -c 5000 gpurun_out/r2m_bench.log

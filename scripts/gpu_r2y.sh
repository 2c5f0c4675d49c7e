#!/bin/bash
# round 2: per-token-head cache write with 4 rows per warp; C3 sweep of the decode rows' CTA share after the epilogue work
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_per_token_head.py -x -q -m gpu 2>&1 | tail -3 > gpurun_out/r2y_tests.txt
tail -2 gpurun_out/r2y_tests.txt
{
for pth in fp8 int8; do
  timeout 300 python scripts/kbench.py cache --tokens 8192 --pth $pth
  timeout 300 python scripts/kbench.py cache --tokens 65536 --pth $pth
done
timeout 300 python scripts/kbench.py cache --tokens 512 --pth fp8
timeout 300 python scripts/kbench.py cache --tokens 8192 --fp8
timeout 300 python scripts/kbench.py prefill --shape mixed
for k in 20 24 28 32 36 40; do
  timeout 300 python scripts/kbench.py prefill --shape mixed --tune $((k << 13))
done
timeout 300 python scripts/kbench.py prefill --shape mixed_prompts
} 2>&1 | grep '^{' > gpurun_out/r2y_kbench.jsonl
python - <<'PY'
import json
for l in open("gpurun_out/r2y_kbench.jsonl"):
    d=json.loads(l)
    if d["kind"]=="cache_write": print("cache", d["tokens"], d["per_token_head"] or ("fp8" if d["fp8"] else "bf16"), d["ms"], d["gbs"], d["kernel"])
    else: print("prefill", d["shape"], "tune>>13", d.get("tune",0)>>13, d["ms"], d["tflops"], d["kernel"])
PY

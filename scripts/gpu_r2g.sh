#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
K="timeout 300 python scripts/kbench.py"
: > gpurun_out/r2g_kbench.jsonl
for a in "decode --workload c4 --hq 8 --hkv 1 --plan --trace" "decode --workload c4 --hq 8 --hkv 1 --plan --pool 100 --trace" "decode --workload c2 --plan --trace" "decode --workload c2 --plan --pool 100 --trace"; do
  $K $a >> gpurun_out/r2g_kbench.jsonl 2>> gpurun_out/r2g_kbench.err || echo "[kbench $a] failed"
done
python - <<'PY'
import json
for l in open("gpurun_out/r2g_kbench.jsonl"):
    d = json.loads(l)
    if d["kind"] == "decode_trace":
        print({k: (v["mean"], v["max"]) if isinstance(v, dict) else v for k, v in d.items() if k.endswith("_us") or k in ("runs_per_cta",)})
    else:
        print(d["workload"], "hkv", d["hkv"], "pool", d["ctas"], d["ms"], d["gbs"])
PY

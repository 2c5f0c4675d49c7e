#!/bin/bash
# round 2: work stealing re-measured now that a run epilogue costs a third of what it did
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
{
for pool in 0 2 4 6; do
  timeout 300 python scripts/kbench.py decode --workload c4 --plan --hq 8 --hkv 1 --pool $pool
done
timeout 300 python scripts/kbench.py decode --workload c4 --plan --hq 8 --hkv 1 --pool 4 --trace
for pool in 0 3; do
  timeout 300 python scripts/kbench.py decode --workload c2 --plan --pool $pool
  timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 640 --pool $pool
  timeout 300 python scripts/kbench.py decode --workload c2var --plan --pool $pool
done
} 2>&1 | grep '^{' > gpurun_out/r2aa_kbench.jsonl
python - <<'PY'
import json
for l in open("gpurun_out/r2aa_kbench.jsonl"):
    d=json.loads(l)
    if d["kind"]=="decode_trace": print("TRACE", {k:(v["mean"] if isinstance(v,dict) else v) for k,v in d.items() if k in ("kernel_span_us","streaming_us","tail_us","end_skew_us","runs_per_cta","finalize_us","producer_switch_us")})
    else: print(d["workload"], "ctx", d["ctx"], "hkv", d["hkv"], "pool", d["ctas"], d["ms"], d["gbs"])
PY

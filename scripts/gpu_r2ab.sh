#!/bin/bash
# round 2: range boundaries snapped onto item boundaries they are one step from - decode tests, A/B timings
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 2000 python -m pytest tests/test_gpu_attention.py tests/test_gpu_decode_plan.py tests/test_gpu_lse.py tests/test_gpu_fuzz.py tests/test_gpu_per_token_head.py tests/test_gpu_guard_bands.py tests/test_gpu_full_size_numeric.py tests/test_gpu_cuda_graph.py tests/test_gpu_decode_out_alignment.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -4 > gpurun_out/r2ab_tests.txt
tail -2 gpurun_out/r2ab_tests.txt
{
for t in 0 4194304; do
  timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 640 --split $t
  timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 768 --split $t
  timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 1024 --split $t
  timeout 300 python scripts/kbench.py decode --workload c2var --plan --split $t
  timeout 300 python scripts/kbench.py decode --workload c2 --plan --split $t
  timeout 300 python scripts/kbench.py decode --workload c4 --plan --hq 8 --hkv 1 --split $t
done
timeout 300 python scripts/kbench.py decode --workload c2 --plan --ctx 640 --trace
} 2>&1 | grep '^{' > gpurun_out/r2ab_kbench.jsonl
python - <<'PY'
import json
for l in open("gpurun_out/r2ab_kbench.jsonl"):
    d=json.loads(l)
    if d["kind"]=="decode_trace": print("TRACE", {k:(v["mean"] if isinstance(v,dict) else v) for k,v in d.items() if k in ("kernel_span_us","streaming_us","tail_us","end_skew_us","runs_per_cta","finalize_us")})
    else: print(d["workload"], "ctx", d["ctx"], "hkv", d["hkv"], "no_snap" if d["split"] else "snap", d["ms"], d["gbs"])
PY

#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_cuda_graph.py tests/test_gpu_vs_triton.py -q -m gpu -p no:cacheprovider -x > gpurun_out/test_r1h.log 2>&1; echo "[tests] exit=$? $(tail -1 gpurun_out/test_r1h.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)|Error" gpurun_out/test_r1h.log | head -5
K="timeout 300 python scripts/kbench.py decode"
: > gpurun_out/kbench_r1h.jsonl
for a in "--workload c4 --trace" "--workload c4 --hq 32 --hkv 4" "--workload c4 --hq 16 --hkv 2" "--workload c4 --hq 8 --hkv 1 --trace" "--workload c2 --trace" "--workload c2 --ctx 640 --trace" "--workload c2 --ctx 256" "--workload c2var" "--workload c4 --ctx 4096" "--workload c2 --batch 8" "--workload c2 --batch 1 --ctx 32768"; do
  $K $a >> gpurun_out/kbench_r1h.jsonl 2>> gpurun_out/kbench_r1h.err || echo "[kbench $a] failed: $(tail -1 gpurun_out/kbench_r1h.err | cut -c1-200)"
done
python - <<'PY'
import json
for l in open("gpurun_out/kbench_r1h.jsonl"):
    d=json.loads(l)
    if d["kind"]=="decode_trace":
        print("TRACE", d["workload"], "hkv",d["hkv"],"ctx",d["ctx"], {k:(v["mean"] if isinstance(v,dict) else v) for k,v in d.items() if k.endswith("_us") or k in("items_per_cta",)})
    else:
        print(d["workload"], "B",d["B"],"hkv",d["hkv"],"bs",d["bs"],"ctx",d["ctx"],"hpc",d["hpc"], d["ms"], d["gbs"])
PY

#!/bin/bash
# decode prologue rewrite: parity, trace of one launch, timings; CUDA-graph test; bench e2e with plans + staging
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_cuda_graph.py tests/test_gpu_cache_write.py -q -m gpu -p no:cacheprovider -x > gpurun_out/test_r1f.log 2>&1; echo "[tests] exit=$? $(tail -1 gpurun_out/test_r1f.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)|Error" gpurun_out/test_r1f.log | head -5
K="timeout 300 python scripts/kbench.py decode"
: > gpurun_out/kbench_r1f.jsonl
for a in "--workload c2 --trace" "--workload c2 --ctx 256 --trace" "--workload c2 --ctx 640 --trace" "--workload c2 --ctx 1024" "--workload c2 --ctx 16384" "--workload c4 --trace" "--workload c4 --hq 8 --hkv 1 --trace" "--workload c2var" "--workload c2 --batch 300 --ctx 1024"; do
  $K $a >> gpurun_out/kbench_r1f.jsonl 2>> gpurun_out/kbench_r1f.err || echo "[kbench $a] failed: $(tail -1 gpurun_out/kbench_r1f.err | cut -c1-200)"
done
python - <<'PY'
import json
for l in open("gpurun_out/kbench_r1f.jsonl"):
    d=json.loads(l)
    if d["kind"]=="decode_trace":
        print("TRACE", d["workload"], "B",d["B"],"ctx",d["ctx"],"hkv",d["hkv"], {k:v for k,v in d.items() if k.endswith("_us") or k in("items_per_cta","sm_ghz")})
    else:
        print(d["workload"], "B",d["B"],"hkv",d["hkv"],"ctx",d["ctx"], d["kernel"], d["ms"], d["gbs"])
PY
python bench.py --no-prefill --cpu-seconds 3 > gpurun_out/bench_r1f.log 2>&1; echo "[bench] exit=$? $(tail -1 gpurun_out/bench_r1f.log | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["e2e"], d["roofline"]["achieved"])')"

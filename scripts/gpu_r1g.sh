#!/bin/bash
# FP8 decode with 32-token warp tiles + split accumulation chains; pipelined e2e in bench.py
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_cuda_graph.py tests/test_gpu_vs_triton.py -q -m gpu -p no:cacheprovider -x > gpurun_out/test_r1g.log 2>&1; echo "[tests] exit=$? $(tail -1 gpurun_out/test_r1g.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)|Error" gpurun_out/test_r1g.log | head -5
K="timeout 300 python scripts/kbench.py decode"
: > gpurun_out/kbench_r1g.jsonl
for a in "--workload c4 --trace" "--workload c4 --hpc 2" "--workload c4 --hpc 8" "--workload c4 --hq 32 --hkv 4" "--workload c4 --hq 16 --hkv 2" "--workload c4 --hq 8 --hkv 1 --trace" "--workload c4 --hq 8 --hkv 1 --bs 32" "--workload c4 --hq 8 --hkv 1 --bs 128" "--workload c2" "--workload c2 --ctx 640" "--workload c2var" "--workload c4 --ctx 4096" ; do
  $K $a >> gpurun_out/kbench_r1g.jsonl 2>> gpurun_out/kbench_r1g.err || echo "[kbench $a] failed: $(tail -1 gpurun_out/kbench_r1g.err | cut -c1-200)"
done
python - <<'PY'
import json
for l in open("gpurun_out/kbench_r1g.jsonl"):
    d=json.loads(l)
    if d["kind"]=="decode_trace":
        print("TRACE", d["workload"], "hkv",d["hkv"], {k:(v["mean"] if isinstance(v,dict) else v) for k,v in d.items() if k.endswith("_us") or k in("items_per_cta",)})
    else:
        print(d["workload"], "B",d["B"],"hkv",d["hkv"],"bs",d["bs"],"ctx",d["ctx"],"hpc",d["hpc"], d["ms"], d["gbs"])
PY
python bench.py --no-prefill --cpu-seconds 3 > gpurun_out/bench_r1g.log 2>&1; echo "[bench] exit=$? $(tail -1 gpurun_out/bench_r1g.log | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["e2e"], d["roofline"]["achieved"])')"
python bench.py --no-prefill --no-cpu-baseline --e2e-serial > gpurun_out/bench_r1g_serial.log 2>&1; echo "[bench serial e2e] exit=$? $(tail -1 gpurun_out/bench_r1g_serial.log | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["value"], d["e2e"]["value"], d["e2e"]["ms_per_step"])')"

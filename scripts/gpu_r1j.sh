#!/bin/bash
# paired K/V boxes (fp8, D=128): parity + timings against unpaired (tune bit 3), both cache layouts
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_cuda_graph.py tests/test_gpu_vs_triton.py -q -m gpu -p no:cacheprovider -x > gpurun_out/test_r1j.log 2>&1; echo "[tests] exit=$? $(tail -1 gpurun_out/test_r1j.log | cut -c1-200)"; grep -E "^(FAILED|ERROR)|Error" gpurun_out/test_r1j.log | head -5
K="timeout 300 python scripts/kbench.py decode"
: > gpurun_out/kbench_r1j.jsonl
for lay in planes interleaved; do for tune in 0 8; do
for a in "--workload c4" "--workload c4 --hq 32 --hkv 4" "--workload c4 --hq 16 --hkv 2" "--workload c4 --hq 8 --hkv 1" "--workload c4 --ctx 4096"; do
  $K $a --layout $lay --split $tune >> gpurun_out/kbench_r1j.jsonl 2>> gpurun_out/kbench_r1j.err || echo "[kbench $a] failed: $(tail -1 gpurun_out/kbench_r1j.err | cut -c1-200)"
done; done; done
for a in "--workload c4 --hpc 1" "--workload c4 --hpc 4" "--workload c4 --hq 16 --hkv 2 --hpc 1" "--workload c4 --hq 8 --hkv 1 --layout interleaved --trace"; do $K $a >> gpurun_out/kbench_r1j.jsonl 2>> gpurun_out/kbench_r1j.err; done
python - <<'PY'
import json
for l in open("gpurun_out/kbench_r1j.jsonl"):
    d=json.loads(l)
    if d["kind"]=="decode_trace":
        print("TRACE", d["workload"], "hkv",d["hkv"],"ctx",d["ctx"], {k:(v["mean"] if isinstance(v,dict) else v) for k,v in d.items() if k.endswith("_us") or k in("items_per_cta",)})
    else:
        print(d["layout"], "tune",d["split"], "hkv",d["hkv"],"ctx",d["ctx"],"hpc",d["hpc"], d["ms"], d["gbs"])
PY

#!/bin/bash
# cp.async producer mode of the decode kernel: parity, then forced-mode timings; decode fixed-cost fit; then C5 full.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_vs_triton.py -q -m gpu -p no:cacheprovider -x > gpurun_out/test_decode2.log 2>&1; echo "[tests] exit=$? $(tail -1 gpurun_out/test_decode2.log | cut -c1-200)"
K="timeout 300 python scripts/kbench.py decode"
: > gpurun_out/kbench_cpasync.jsonl
for tune in 4 2; do   # 4 = TMA boxes only, 2 = cp.async producers forced
  for a in "--workload c4 --hq 8 --hkv 1" "--workload c4 --hq 16 --hkv 2" "--workload c4 --hq 16 --hkv 2 --hpc 1" "--workload c4 --hq 8 --hkv 1 --bs 32" "--workload c2 --hq 4 --hkv 1" "--workload c2" "--workload c4" "--workload c4 --hpc 1" "--workload c4 --hpc 2"; do
    $K $a --split $tune >> gpurun_out/kbench_cpasync.jsonl 2>> gpurun_out/kbench_cpasync.err || echo "[kbench $a tune=$tune] failed: $(tail -1 gpurun_out/kbench_cpasync.err | cut -c1-200)"
  done
done
# fixed cost fit: time vs bytes on C2 shape
for ctx in 256 512 1024 2048 4096 8192 16384; do $K --workload c2 --ctx $ctx >> gpurun_out/kbench_cpasync.jsonl 2>> gpurun_out/kbench_cpasync.err; done
python - <<'PY'
import json
for l in open("gpurun_out/kbench_cpasync.jsonl"):
    d=json.loads(l); print(d["workload"], "hkv",d["hkv"],"bs",d["bs"],"ctx",d["ctx"],"hpc",d["hpc"],"tune",d["split"], d["kernel"], d["ms"], d["gbs"])
PY
SKIP_TRITON_TESTS=1 SKIP_SMALL=1 bash scripts/gpu_e2e.sh

#!/bin/bash
# Round-end evidence pass: full GPU suite, smoke, bench (both arms), kernel benches, ncu launch list + full captures.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/test_final.log 2>&1; echo "[tests] exit=$? $(tail -1 gpurun_out/test_final.log)"
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "[smoke] exit=$? $(tail -1 gpurun_out/smoke_final.log | cut -c1-200)"
bash scripts/gpu_profile.sh
K="timeout 300 python scripts/kbench.py"
for a in "decode --workload c4 --hq 32 --hkv 4" "decode --workload c4 --hq 16 --hkv 2" "decode --workload c2var" "decode --workload c2 --ctx 640" "decode --workload c2 --ctx 16384" "prefill --shape full8k" "prefill --shape full2k" "prefill --shape chunk" "prefill --shape mixed" "prefill --shape full8k --fp8" "prefill --shape mixed --fp8" "prefill --shape full8k --softcap 50" "prefill --shape full8k --force 22" "prefill --shape full8k --force 22 --softcap 50" "prefill --shape full8k --fp8 --no-staging" "decode --workload c2 --graph" "decode --workload c2 --split 512" "decode --workload c2 --batch 4 --ctx 1024"; do
  $K $a >> gpurun_out/kbench.jsonl 2>> gpurun_out/kbench.err || echo "[kbench $a] failed"
done
for a in "decode --workload c2" "decode --workload c4" "prefill --shape full8k" "prefill --shape mixed" "prefill --shape full8k --softcap 50"; do $K $a --impl triton >> gpurun_out/kbench.jsonl 2>> gpurun_out/kbench.err || echo "[kbench triton $a] failed"; done
python - <<'PY'
import json
for l in open("gpurun_out/kbench.jsonl"):
    d=json.loads(l); print(d["kind"], d.get("workload",d.get("shape","")), "hkv",d.get("hkv"), "fp8",d.get("fp8"), "cap",d.get("softcap"), "force",d.get("force"), "graph",d.get("graph"), "split",d.get("split"), d.get("impl","ours"), d["ms"], d.get("gbs",d.get("tflops")), d.get("kernel"))
PY

#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_rope_cache.py -q -m gpu -p no:cacheprovider -rs -k vllm_unfused > gpurun_out/test_rope2.log 2>&1; echo "[rope vs vllm] exit=$? $(tail -1 gpurun_out/test_rope2.log | cut -c1-200)"; grep -E "SKIPPED|FAILED|Error" gpurun_out/test_rope2.log | head -5 | cut -c1-300
# C3 breakdown: per-kernel durations of the mixed batch
P="python scripts/kbench.py prefill --shape mixed"
$P > gpurun_out/plain_mixed.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:attn_ -c 60 --csv --log-file gpurun_out/launches_mixed.csv $P > gpurun_out/ncu_mixed.log 2>&1
python - <<'PY'
import csv, collections
rows=[l for l in open("gpurun_out/launches_mixed.csv") if not l.startswith("==")]
agg=collections.defaultdict(list)
for r in csv.DictReader(rows):
    try: agg[r["Kernel Name"][:60]].append(float(r["Metric Value"].replace(",","")))
    except Exception: pass
for k,v in agg.items(): print("C3 kernel", k, "n",len(v), "mean_us", round(sum(v)/len(v)/1e3,2))
PY
cut -c1-200 gpurun_out/plain_mixed.log | tail -1
# chunked prefill + prefix caching inside vLLM (mixed batches in situ), full model
E="python scripts/e2e_vllm.py"
for b in b200 triton; do
timeout 900 $E --backend $b --prefix-caching --shared-prefix 256 --max-batched-tokens 2048 --save gpurun_out/e2e_${b}_chunked.json > gpurun_out/e2e_${b}_chunked.log 2>&1; echo "[e2e $b chunked+prefix] exit=$? $(grep '^{"metric"' gpurun_out/e2e_${b}_chunked.log | tail -1 | cut -c1-330)"; grep -n "Error" gpurun_out/e2e_${b}_chunked.log | tail -3 | cut -c1-300
done
$E --compare gpurun_out/e2e_b200_chunked.json gpurun_out/e2e_triton_chunked.json

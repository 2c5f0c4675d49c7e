#!/bin/bash
# Round 2, call J: full GPU suite on the decode rework (stealing off by default) + ncu capture of the 1-kv-head C4 shard.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/r2j_tests.log 2>&1; echo "[tests] exit=$? $(tail -3 gpurun_out/r2j_tests.log | tr '\n' ' ' | cut -c1-400)"
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2j_smoke.log 2>&1; echo "[smoke] exit=$? $(tail -1 gpurun_out/r2j_smoke.log | cut -c1-200)"
D="python scripts/kbench.py decode --workload c4 --hq 8 --hkv 1 --plan"
$D > gpurun_out/r2j_plain_decode.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attn_decode_kernel -s 3 -c 1 -o gpurun_out/r2j_prof_decode_shard -f $D > gpurun_out/r2j_ncu_decode.log 2>&1
echo "[ncu decode shard] exit=$? $(tail -2 gpurun_out/r2j_plain_decode.log | cut -c1-300)"
ls -la gpurun_out/*.ncu-rep

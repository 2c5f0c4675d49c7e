#!/bin/bash
# Round 2, call A: baseline decode traces (1-kv-head C4 shard), compute-sanitizer over a reduced GPU subset,
# ncu captures of the cache-write kernels at prompt sizes.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
K="timeout 300 python scripts/kbench.py"
: > gpurun_out/r2a_kbench.jsonl
for a in "decode --workload c4 --hq 8 --hkv 1 --trace" "decode --workload c4 --hq 8 --hkv 1 --graph" "decode --workload c4 --trace" \
         "decode --workload c2 --trace" "decode --workload c2 --batch 8 --trace" "cache --tokens 8192" "cache --tokens 8192 --fp8" "cache --tokens 64"; do
  $K $a >> gpurun_out/r2a_kbench.jsonl 2>> gpurun_out/r2a_kbench.err || echo "[kbench $a] failed"
done
cut -c1-700 gpurun_out/r2a_kbench.jsonl
# ---- compute-sanitizer (SURVEY.md section 5 "race detection"): split/merge decode with changing batch sizes, tcgen05
# with CLC stealing, FP8 staging, CUDA-graph replay
S="compute-sanitizer --error-exitcode 17 --launch-timeout 0"
T1="tests/test_gpu_attention.py::test_one_workspace_serves_batches_of_changing_size tests/test_gpu_attention.py::test_workspace_counters_are_left_clean tests/test_gpu_attention.py::test_mixed_batch_is_routed_across_three_kernels tests/test_gpu_attention.py::test_fp8_prefill_is_staged_for_the_tcgen05_kernel tests/test_gpu_attention.py::test_decode_empty_and_single_token_sequences"
T2="tests/test_gpu_cuda_graph.py tests/test_gpu_attention.py::test_prefill_tc_large_logits_exercise_the_lazy_rescale tests/test_gpu_cache_write.py::test_golden_reference_vector_bit_exact tests/test_gpu_rope_cache.py"
for tool in memcheck racecheck synccheck initcheck; do
  timeout 900 $S --tool $tool python -m pytest $T1 -x -q -p no:cacheprovider > gpurun_out/r2a_san_${tool}_1.log 2>&1
  echo "[sanitizer $tool set1] exit=$? $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|passed|failed' gpurun_out/r2a_san_${tool}_1.log | tail -3 | tr '\n' ' ')"
done
for tool in memcheck racecheck; do
  timeout 900 $S --tool $tool python -m pytest $T2 -x -q -p no:cacheprovider > gpurun_out/r2a_san_${tool}_2.log 2>&1
  echo "[sanitizer $tool set2] exit=$? $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY|passed|failed' gpurun_out/r2a_san_${tool}_2.log | tail -3 | tr '\n' ' ')"
done
# ---- ncu: K1 / K5 at 8192 tokens
C="python scripts/kbench.py cache --tokens 8192"
$C > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:cache -s 3 -c 1 -o gpurun_out/r2a_prof_k1_8k -f $C > gpurun_out/r2a_ncu_k1.log 2>&1
echo "[ncu k1] exit=$?"
ls -la gpurun_out | tail -20

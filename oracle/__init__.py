"""CPU oracle for the paged-KV attention path.  TEST INFRASTRUCTURE ONLY.

This package is a CPU restatement of the reference algorithm
(foundation-model-stack/vllm-triton-backend, `ibm-triton-lib`), used as the
checker for the CUDA product in `vllm_triton_backend_b200/`.  Only `tests/`,
`__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of
`bench.py` may import it.  The product package never imports `oracle`; it fails
loudly when its CUDA library is missing.

Parity status: PINNED.  `oracle/make_golden.py` (run in the build container,
where `/root/reference` is mounted) executes the reference's own Triton kernels
(`kernel_unified_attention_2d/3d`, `reduce_segments`) under
`TRITON_INTERPRET=1` and the reference's torch functions (`ref_paged_attn`,
`ref_reshape_and_cache_flash`) on seeded inputs and commits their outputs to
`tests/golden/*.npz`.  `tests/test_oracle_golden.py` holds this oracle to those
vectors.
"""

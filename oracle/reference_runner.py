"""CPU oracle helper (test infrastructure): run the REFERENCE's own code.

Only usable where `/root/reference` is mounted (the build container).  It loads
`ibm-triton-lib/ibm_triton_lib/kernels/triton_unified_attention.py` by file path
— nothing is copied into this repo — and executes `kernel_unified_attention_2d`,
`kernel_unified_attention_3d` and `reduce_segments` with Triton's CPU
interpreter (`TRITON_INTERPRET=1`).  `triton_dejavu` (the reference's autotuner,
missing from the archive and contributing no arithmetic) is
replaced by a stub that pins BLOCK_M/BLOCK_N.

The Triton 3.6 interpreter mishandles bf16 inputs, so this runner is used with
fp32 / fp16 tensors only.  Outputs feed `oracle/make_golden.py`.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("B200ATTN_REFERENCE_ROOT", "/root/reference")
_REPO_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# The GPU box has no /root/reference.  __graft_entry__.build() puts a copy of this ONE file into the git-ignored
# baseline/_ref/ (which travels with gpurun), so the reference's own kernels - not a sibling copy - run there too.
REF_COPY = os.path.join(_REPO_ROOT, "baseline", "_ref", "triton_unified_attention.py")
_KERNEL_FILE = os.path.join(
    REFERENCE_ROOT, "ibm-triton-lib", "ibm_triton_lib", "kernels", "triton_unified_attention.py"
)
if not os.path.isfile(_KERNEL_FILE) and os.path.isfile(REF_COPY):
    _KERNEL_FILE = REF_COPY


def reference_available() -> bool:
    return os.path.isfile(_KERNEL_FILE)


def _install_dejavu_stub(block_m: int, block_n: int) -> None:
    stub = types.ModuleType("triton_dejavu")

    def jitcache(*_a, **_k):
        return lambda fn: fn

    class ConfigSpace:  # noqa: D401 - placeholder for the tuner's search space
        def __init__(self, *a, **k):
            pass

    class _Pinned:
        """Stands in for an autotuned kernel: kernel[grid](**kw) with fixed tiles."""

        def __init__(self, fn):
            self.fn = fn

        def __getitem__(self, grid):
            def launch(**kw):
                meta = dict(kw)
                meta.update(BLOCK_M=stub.BLOCK_M, BLOCK_N=stub.BLOCK_N)
                g = grid(meta) if callable(grid) else grid
                return self.fn[g](**kw, BLOCK_M=stub.BLOCK_M, BLOCK_N=stub.BLOCK_N)

            return launch

    def autotune(*_a, **_k):
        return lambda fn: _Pinned(fn)

    stub.jitcache = jitcache
    stub.ConfigSpace = ConfigSpace
    stub.autotune = autotune
    stub.global_cache_lock = types.SimpleNamespace(lock=lambda: None, unlock=lambda: None, is_locked=False)
    stub.BLOCK_M = block_m
    stub.BLOCK_N = block_n
    sys.modules["triton_dejavu"] = stub


_MOD = None


def load_reference_kernels(block_m: int = 16, block_n: int = 32, interpret: bool = True):
    """Returns the reference module (with `.unified_attention`).  `interpret=False`: compiled for the GPU in hand."""
    global _MOD
    if not reference_available():
        raise FileNotFoundError(_KERNEL_FILE)
    if interpret and os.environ.get("TRITON_INTERPRET") != "1":
        raise RuntimeError("set TRITON_INTERPRET=1 before importing triton (CPU interpreter)")
    _install_dejavu_stub(block_m, block_n)
    if _MOD is None:
        spec = importlib.util.spec_from_file_location("_ref_triton_unified_attention", _KERNEL_FILE)
        _MOD = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_MOD)
    sys.modules["triton_dejavu"].BLOCK_M = block_m
    sys.modules["triton_dejavu"].BLOCK_N = block_n
    return _MOD


def load_reference_torch_utils():
    """`scripts/vllm_utils.py` (ref_paged_attn, ref_reshape_and_cache_flash, ...)."""
    path = os.path.join(REFERENCE_ROOT, "scripts", "vllm_utils.py")
    # the file imports two vLLM helpers at module scope that moved in vLLM 0.22;
    # neither is used by the functions we call.
    import vllm.utils as vu

    if not hasattr(vu, "get_kv_cache_torch_dtype"):
        vu.get_kv_cache_torch_dtype = lambda *a, **k: None
    spec = importlib.util.spec_from_file_location("_ref_vllm_utils", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod

"""b200attn — B200-native (sm_100a) paged-KV attention path for vLLM.

Drop-in for the hot path `ibm-triton-lib` exposes to vLLM: KV-cache write,
split-KV paged decode attention, varlen / prefix / chunked-prefill attention,
behind the same operator and AttentionBackend/AttentionImpl surface.  The
kernels live in `libb200attn.so` (hand-written CUDA, C ABI in include/b200attn.h);
this package is the thin Python host side.
"""
from ._cabi import LIB_PATH, B200AttnError, load_library  # noqa: F401

__version__ = "0.1.0"


def __getattr__(name):
    if name in ("reshape_and_cache_flash", "unified_attention", "scaled_fp8_quant", "workspace_bytes"):
        from . import ops

        return getattr(ops, name)
    raise AttributeError(name)

"""ctypes binding of libb200attn.so (the C ABI declared in include/b200attn.h).

The library is the product: if it is missing this module raises — there is no
CPU or PyTorch fallback for the kernels.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int32, c_int64, c_uint64, c_void_p

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG_DIR, "libb200attn.so")

# enum b200attn_dtype
BF16, F16, F32, FP8_E4M3, INT8 = 0, 1, 2, 3, 4
# enum b200attn_path
PATH_AUTO, PATH_GENERIC, PATH_PREFILL, PATH_DECODE, PATH_PREFILL_TC = 0, 1, 2, 3, 4

STATUS_NAMES = {
    0: "OK", 1: "E_BAD_DTYPE", 2: "E_BAD_SHAPE", 3: "E_ALIGNMENT", 4: "E_WORKSPACE", 5: "E_UNSUPPORTED",
    6: "E_CUDA", 7: "E_NULL",
}

ABI_VERSION = 5  # include/b200attn.h B200ATTN_ABI_VERSION

EXPORTED_SYMBOLS = [
    "b200attn_reshape_and_cache_flash",
    "b200attn_reshape_and_cache_flash_per_token_head",
    "b200attn_workspace_bytes",
    "b200attn_unified_attention",
    "b200attn_decode_plan_bytes",
    "b200attn_decode_plan",
    "b200attn_decode_attention_static_scratch",
    "b200attn_scaled_fp8_quant",
    "b200attn_rope_and_cache_flash",
    "b200attn_abi_version",
    "b200attn_params_size",
    "b200attn_last_error",
    "b200attn_launch_count",
    "b200attn_reset_launch_count",
    "b200attn_last_kernel",
    "b200attn_set_tuning",
    "b200attn_debug_set_decode_trace",
]


class AttentionParams(Structure):
    """Mirror of `b200attn_attention_params` (include/b200attn.h)."""

    _fields_ = [
        ("out", c_void_p), ("q", c_void_p), ("k_cache", c_void_p), ("v_cache", c_void_p),
        ("block_tables", c_void_p), ("seq_lens", c_void_p), ("cu_seqlens_q", c_void_p),
        ("alibi_slopes", c_void_p), ("k_scale", c_void_p), ("v_scale", c_void_p), ("sinks", c_void_p),
        ("softmax_scale", c_float), ("softcap", c_float), ("sliding_window", c_int32),
        ("num_tokens", c_int32), ("num_seqs", c_int32), ("num_q_heads", c_int32), ("num_kv_heads", c_int32),
        ("head_size", c_int32), ("block_size", c_int32), ("max_seqlen_q", c_int32), ("max_seqlen_k", c_int32),
        ("q_dtype", c_int32), ("kv_dtype", c_int32), ("path", c_int32), ("reserved0", c_int32),
        ("num_blocks", c_int64), ("q_stride0", c_int64), ("q_stride1", c_int64),
        ("out_stride0", c_int64), ("out_stride1", c_int64), ("bt_stride0", c_int64),
        ("kc_block_stride", c_int64), ("kc_slot_stride", c_int64), ("kc_head_stride", c_int64),
        ("vc_block_stride", c_int64), ("vc_slot_stride", c_int64), ("vc_head_stride", c_int64),
        ("workspace", c_void_p), ("workspace_bytes", c_uint64),
        ("out_scale", c_void_p),
        ("qq_bias", c_void_p), ("qq_bias_stride0", c_int64),
        ("lse", c_void_p), ("lse_stride0", c_int64),
        ("decode_plan", c_void_p), ("decode_plan_key", c_uint64),
        ("k_scale_cache", c_void_p), ("v_scale_cache", c_void_p),
        ("ks_block_stride", c_int64), ("ks_slot_stride", c_int64), ("ks_head_stride", c_int64),
        ("vs_block_stride", c_int64), ("vs_slot_stride", c_int64), ("vs_head_stride", c_int64),
    ]


class RopeCacheParams(Structure):
    """Mirror of `b200attn_rope_cache_params` (include/b200attn.h)."""

    _fields_ = [
        ("q", c_void_p), ("k", c_void_p), ("v", c_void_p), ("positions", c_void_p), ("cos_sin_cache", c_void_p),
        ("k_cache", c_void_p), ("v_cache", c_void_p), ("slot_mapping", c_void_p), ("k_scale", c_void_p),
        ("v_scale", c_void_p), ("rope_tokens", c_int64), ("cache_tokens", c_int64),
        ("q_stride0", c_int64), ("q_stride1", c_int64), ("k_stride0", c_int64), ("k_stride1", c_int64),
        ("v_stride0", c_int64), ("v_stride1", c_int64),
        ("kc_block_stride", c_int64), ("kc_slot_stride", c_int64), ("kc_head_stride", c_int64),
        ("vc_block_stride", c_int64), ("vc_slot_stride", c_int64), ("vc_head_stride", c_int64),
        ("num_q_heads", c_int32), ("num_kv_heads", c_int32), ("head_size", c_int32), ("rot_dim", c_int32),
        ("block_size", c_int32), ("is_neox", c_int32), ("src_dtype", c_int32), ("kv_dtype", c_int32),
    ]


class B200AttnError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"b200attn {STATUS_NAMES.get(code, code)}: {message}")
        self.code = code


_lib = None


def load_library() -> ctypes.CDLL:
    """Loads the in-tree library; raises if it was not built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m vllm_triton_backend_b200.build` "
            "(nvcc, sm_100a). There is no CPU/PyTorch fallback for these kernels."
        )
    lib = ctypes.CDLL(LIB_PATH)
    lib.b200attn_reshape_and_cache_flash.restype = c_int32
    lib.b200attn_reshape_and_cache_flash.argtypes = (
        [c_void_p] * 5 + [c_int64, c_int32, c_int32, c_int32] + [c_int64] * 8 + [c_int32, c_int32, c_void_p, c_void_p, c_void_p]
    )
    lib.b200attn_reshape_and_cache_flash_per_token_head.restype = c_int32
    lib.b200attn_reshape_and_cache_flash_per_token_head.argtypes = (
        [c_void_p] * 7 + [c_int64, c_int32, c_int32, c_int32] + [c_int64] * 16 + [c_int32, c_int32, c_void_p]
    )
    lib.b200attn_workspace_bytes.restype = c_uint64
    lib.b200attn_workspace_bytes.argtypes = [c_int32] * 5
    lib.b200attn_unified_attention.restype = c_int32
    lib.b200attn_unified_attention.argtypes = [POINTER(AttentionParams), c_void_p]
    lib.b200attn_decode_plan_bytes.restype = c_uint64
    lib.b200attn_decode_plan_bytes.argtypes = [c_int32]
    lib.b200attn_decode_plan.restype = c_int32
    lib.b200attn_decode_plan.argtypes = [POINTER(AttentionParams), c_void_p, c_uint64, POINTER(c_uint64), c_void_p]
    lib.b200attn_decode_attention_static_scratch.restype = c_int32
    lib.b200attn_decode_attention_static_scratch.argtypes = [POINTER(AttentionParams), c_void_p]
    lib.b200attn_rope_and_cache_flash.restype = c_int32
    lib.b200attn_rope_and_cache_flash.argtypes = [POINTER(RopeCacheParams), c_void_p]
    lib.b200attn_scaled_fp8_quant.restype = c_int32
    lib.b200attn_scaled_fp8_quant.argtypes = [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int32, c_void_p, c_void_p]
    lib.b200attn_abi_version.restype = c_int32
    lib.b200attn_params_size.restype = c_uint64
    lib.b200attn_params_size.argtypes = [c_int32]
    if lib.b200attn_abi_version() != ABI_VERSION or lib.b200attn_params_size(0) != ctypes.sizeof(AttentionParams) \
            or lib.b200attn_params_size(1) != ctypes.sizeof(RopeCacheParams):
        raise ImportError(
            f"{LIB_PATH} was built from other sources (ABI {lib.b200attn_abi_version()} vs {ABI_VERSION}, params "
            f"{lib.b200attn_params_size(0)}/{lib.b200attn_params_size(1)} bytes vs {ctypes.sizeof(AttentionParams)}/"
            f"{ctypes.sizeof(RopeCacheParams)}): rebuild with `python -m vllm_triton_backend_b200.build`"
        )
    lib.b200attn_last_error.restype = c_char_p
    lib.b200attn_launch_count.restype = c_uint64
    lib.b200attn_reset_launch_count.restype = None
    lib.b200attn_last_kernel.restype = c_char_p
    lib.b200attn_set_tuning.restype = None
    lib.b200attn_set_tuning.argtypes = [c_int32] * 4
    lib.b200attn_debug_set_decode_trace.restype = None
    lib.b200attn_debug_set_decode_trace.argtypes = [c_void_p]
    _lib = lib
    return lib


def check(code: int) -> None:
    if code != 0:
        raise B200AttnError(code, load_library().b200attn_last_error().decode())


def launch_count() -> int:
    return int(load_library().b200attn_launch_count())


def reset_launch_count() -> None:
    load_library().b200attn_reset_launch_count()


def last_kernel() -> str:
    return load_library().b200attn_last_kernel().decode()


def set_tuning(decode_heads_per_cta: int = 0, decode_pool_sixteenths: int = 0, decode_stages: int = 0, reserved: int = 0):
    load_library().b200attn_set_tuning(decode_heads_per_cta, decode_pool_sixteenths, decode_stages, reserved)

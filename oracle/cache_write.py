"""CPU oracle (test infrastructure, not product): KV-cache scatter write.

Restates `reshape_and_cache_flash_kernel` (vllm/csrc/cache_kernels.cu:286-358,
host wrapper :700-738) and the reference's torch version
`ref_reshape_and_cache_flash` (scripts/vllm_utils.py:377-401):

    for i < slot_mapping.size(0), slot = slot_mapping[i], slot >= 0:
        cache[slot // bs, slot % bs, h, :] = f(src[i, h, :])

with f = copy (`auto`) or, for fp8, `__nv_cvt_float_to_fp8(float(x) / scale,
__NV_SATFINITE, __NV_E4M3)` (quant_utils.cuh:479-502) — a true fp32 division,
round-to-nearest-even, saturating to +-448, NaN -> 0x7f.

Two implementations that must agree byte for byte (tests/test_oracle_cache.py):
torch (`.to(float8_e4m3fn)` after clamping) and plain C (`oracle/e4m3_ref.c`,
loaded with ctypes when built).
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import numpy as np
import torch

_FP8 = torch.float8_e4m3fn
_HERE = os.path.dirname(os.path.abspath(__file__))
_C_LIB_PATH = os.path.join(_HERE, "_build", "liboracle_e4m3.so")


def quantize_e4m3_torch(x: torch.Tensor, scale: float) -> torch.Tensor:
    """fp32 divide, saturate, RNE -> uint8 bit patterns."""
    y = x.to(torch.float32) / torch.tensor(scale, dtype=torch.float32)
    nan = torch.isnan(y)
    y = y.clamp(-448.0, 448.0)
    bits = y.to(_FP8).view(torch.uint8)
    # __NV_SATFINITE keeps NaN as NaN; the canonical pattern hardware emits is 0x7f.
    return torch.where(nan, torch.full_like(bits, 0x7F), bits)


def scaled_fp8_quant_ref(x: torch.Tensor, scale: float) -> torch.Tensor:
    """Static per-tensor quantisation of the query (`ops.scaled_fp8_quant(query, layer._q_scale)`, reference
    ibm-triton-lib/ibm_triton_lib/backend/triton_attn.py:417): vllm/csrc/quantization/w8a8/fp8/common.cu:11-29 calls
    `scaled_fp8_conversion<true>(val, 1 / *scale)` (common.cuh:40-60) - a MULTIPLICATION by the fp32 reciprocal (not the
    division of the cache write), `fmaxf(-448, fminf(x, 448))` (so NaN becomes +448: fminf returns the number), then
    the round-to-nearest-even e4m3 conversion.  Returns uint8 bit patterns."""
    inv = torch.tensor(1.0, dtype=torch.float32) / torch.tensor(scale, dtype=torch.float32)
    y = x.to(torch.float32) * inv
    hi, lo = torch.tensor(448.0), torch.tensor(-448.0)
    y = torch.fmax(lo, torch.fmin(y, hi))  # fmin/fmax ignore a NaN operand, like fminf/fmaxf
    return y.to(_FP8).view(torch.uint8)


def load_c_lib() -> Optional[ctypes.CDLL]:
    if not os.path.exists(_C_LIB_PATH):
        return None
    lib = ctypes.CDLL(_C_LIB_PATH)
    lib.oracle_f32_to_e4m3.restype = ctypes.c_uint8
    lib.oracle_f32_to_e4m3.argtypes = [ctypes.c_float]
    lib.oracle_quantize_e4m3.restype = None
    lib.oracle_quantize_e4m3.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_float]
    lib.oracle_e4m3_to_f32.restype = ctypes.c_float
    lib.oracle_e4m3_to_f32.argtypes = [ctypes.c_uint8]
    lib.oracle_reshape_and_cache_flash_u16.restype = None
    lib.oracle_reshape_and_cache_flash_u16.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int64] * 9
    return lib


def quantize_e4m3_c(x: torch.Tensor, scale: float) -> torch.Tensor:
    lib = load_c_lib()
    if lib is None:
        raise RuntimeError("oracle C library not built; run `make -C oracle`")
    src = np.ascontiguousarray(x.to(torch.float32).numpy())
    dst = np.empty(src.shape, dtype=np.uint8)
    lib.oracle_quantize_e4m3(src.ctypes.data, dst.ctypes.data, src.size, ctypes.c_float(scale))
    return torch.from_numpy(dst)


def reshape_and_cache_flash_ref(
    key: torch.Tensor,  # [>=T, Hkv, D]
    value: torch.Tensor,
    key_cache: torch.Tensor,  # [NB, bs, Hkv, D]  (any block stride), model dtype or uint8/float8
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,  # int64 [T]; T comes from here (cache_kernels.cu:709-719)
    kv_cache_dtype: str = "auto",
    k_scale: float = 1.0,
    v_scale: float = 1.0,
) -> None:
    bs = key_cache.shape[1]
    fp8 = kv_cache_dtype.startswith("fp8")
    kc = key_cache.view(torch.uint8) if key_cache.dtype == _FP8 else key_cache
    vc = value_cache.view(torch.uint8) if value_cache.dtype == _FP8 else value_cache
    for i, slot in enumerate(slot_mapping.tolist()):
        if slot < 0:  # padded token (cache_kernels.cu:300-303)
            continue
        b, o = slot // bs, slot % bs
        if fp8:
            kc[b, o] = quantize_e4m3_torch(key[i], k_scale)
            vc[b, o] = quantize_e4m3_torch(value[i], v_scale)
        else:
            kc[b, o] = key[i]
            vc[b, o] = value[i]


def per_token_head_quant_ref(x: torch.Tensor, kv_dtype: torch.dtype):
    """Dynamic per-(token, head) quantisation of one [T, H, D] tensor.  The algorithm lives in a third-party dependency
    of the reference, vLLM (pinned here: the installed vllm 0.22.0, `_reshape_cache_per_token_head`,
    vllm/v1/attention/ops/triton_reshape_and_cache_flash.py:138-236; `ibm-triton-lib` is a plugin for that vLLM and its
    backend sits where `TritonAttentionImpl.do_kv_cache_update`, vllm/v1/attention/backends/triton_attn.py:733-748,
    calls it):
        scale = max(absmax(x[t, h, :]) / QMAX, 1e-6)                    (:190, :218)
        q     = clamp(x * (1.0 / scale), QMIN, QMAX) stored through the cache pointer   (:199-208)
    QMAX, QMIN = 448, -448 for e4m3 (round to nearest even) and 127, -128 for int8 (the store of an fp32 value through
    an int8 pointer converts with truncation toward zero).  Returns (q as `kv_dtype`, scale fp32 [T, H])."""
    qmax, qmin = (448.0, -448.0) if kv_dtype == _FP8 else (127.0, -128.0)
    xf = x.to(torch.float32)
    scale = torch.clamp(xf.abs().amax(dim=-1) / torch.tensor(qmax, dtype=torch.float32), min=1e-6)
    inv = torch.tensor(1.0, dtype=torch.float32) / scale
    y = (xf * inv.unsqueeze(-1)).clamp(qmin, qmax)
    q = y.to(_FP8) if kv_dtype == _FP8 else y.trunc().to(torch.int8)
    return q, scale


def reshape_and_cache_flash_per_token_head_ref(key, value, key_cache, value_cache, k_scale_cache, v_scale_cache,
                                               slot_mapping) -> None:
    """`triton_reshape_and_cache_flash_per_token_head_quant` (vLLM 0.22
    vllm/v1/attention/ops/triton_reshape_and_cache_flash.py:247-320): quantise every (token, head) row of key and
    value with its own scale, scatter the rows into the paged caches and the scales into the scale caches
    ([num_blocks, block_size, num_kv_heads] fp32, any strides).  In place."""
    bs = key_cache.shape[1]
    kq, ks = per_token_head_quant_ref(key, key_cache.dtype)
    vq, vs = per_token_head_quant_ref(value, value_cache.dtype)
    for i, slot in enumerate(int(x) for x in slot_mapping.tolist()):
        if slot < 0:
            continue
        b, o = slot // bs, slot % bs
        key_cache[b, o] = kq[i]
        value_cache[b, o] = vq[i]
        k_scale_cache[b, o] = ks[i]
        v_scale_cache[b, o] = vs[i]


def reshape_and_cache_flash_c(key, value, key_cache, value_cache, slot_mapping) -> None:
    """Plain-C scatter for 16-bit element types (copy path), contiguous NHD."""
    lib = load_c_lib()
    if lib is None:
        raise RuntimeError("oracle C library not built; run `make -C oracle`")
    assert key.element_size() == 2 and key_cache.element_size() == 2
    T = slot_mapping.numel()
    _, bs, H, D = key_cache.shape
    sm = slot_mapping.to(torch.int64).contiguous()
    lib.oracle_reshape_and_cache_flash_u16(
        key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), sm.data_ptr(),
        T, H, D, bs, key.stride(0), value.stride(0),
        key_cache.stride(0), key_cache.stride(1), value_cache.stride(0),
    )


def cache_write_bytes(num_tokens: int, num_kv_heads: int, head_size: int, src_bytes: int, kv_bytes: int) -> int:
    """T*Hkv*D*(2*src + K + V) + 8T."""
    return num_tokens * num_kv_heads * head_size * (2 * src_bytes + 2 * kv_bytes) + 8 * num_tokens

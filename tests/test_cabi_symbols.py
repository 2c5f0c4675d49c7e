"""The C-ABI library loads without a GPU/driver and exports every symbol that
include/b200attn.h declares (no compute calls here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from vllm_triton_backend_b200 import _cabi

    if not os.path.exists(_cabi.LIB_PATH):
        from vllm_triton_backend_b200.build import build

        build()
    return _cabi.load_library()


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200attn.h")).read()
    return sorted(set(re.findall(r"B200ATTN_API[^;(]*?\b(b200attn_\w+)\s*\(", text)))


def test_header_declares_the_expected_entry_points():
    from vllm_triton_backend_b200 import _cabi

    assert _declared_symbols() == sorted(_cabi.EXPORTED_SYMBOLS)


def test_every_declared_symbol_is_exported(lib):
    for sym in _declared_symbols():
        assert getattr(lib, sym) is not None, sym


def test_abi_version_and_host_only_calls(lib):
    from vllm_triton_backend_b200 import _cabi as cabi

    assert lib.b200attn_abi_version() == cabi.ABI_VERSION == 5
    # workspace size is a function of host ints only
    a = lib.b200attn_workspace_bytes(64, 64, 32, 8, 128)
    b = lib.b200attn_workspace_bytes(128, 128, 32, 8, 128)
    assert 0 < a < b
    lib.b200attn_reset_launch_count()
    assert lib.b200attn_launch_count() == 0


def test_params_struct_matches_header_layout():
    """ctypes mirror and the C struct must agree field by field."""
    from vllm_triton_backend_b200._cabi import AttentionParams

    text = open(os.path.join(ROOT, "include", "b200attn.h")).read()
    body = text.split("typedef struct b200attn_attention_params {")[1].split("} b200attn_attention_params;")[0]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        for part in decl.split(","):
            names.append(re.findall(r"(\w+)\s*$", part.strip())[0])
    assert names == [f[0] for f in AttentionParams._fields_]


def test_null_params_is_an_error_code_not_a_crash(lib):
    assert lib.b200attn_unified_attention(None, None) != 0
    assert b"null" in lib.b200attn_last_error()


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from vllm_triton_backend_b200 import _cabi

    monkeypatch.setattr(_cabi, "_lib", None)
    monkeypatch.setattr(_cabi, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ImportError, match="no CPU/PyTorch fallback"):
        _cabi.load_library()


def test_struct_sizes_agree_with_the_compiled_library(lib):
    """load_library() refuses a stale .so; the same check spelled out."""
    from vllm_triton_backend_b200._cabi import AttentionParams, RopeCacheParams

    assert lib.b200attn_params_size(0) == ctypes.sizeof(AttentionParams)
    assert lib.b200attn_params_size(1) == ctypes.sizeof(RopeCacheParams)
    assert lib.b200attn_params_size(7) == 0


def test_static_scratch_limit_matches_what_the_header_promises(lib):
    """include/b200attn.h: b200attn_decode_attention_static_scratch takes shapes whose workspace fits 24 MiB, 'e.g. up to 32
    query heads at head size 128 with 8 kv heads, or a 1-2 kv-head shard of a 64-head model' (host-int arithmetic only)."""
    text = open(os.path.join(ROOT, "include", "b200attn.h")).read()
    m = re.search(r"#define\s+B200ATTN_STATIC_SCRATCH_BYTES\s+\((\d+)u\s*<<\s*(\d+)\)", text)
    limit = int(m.group(1)) << int(m.group(2))
    assert limit == 24 << 20
    fits = [(64, 32, 8, 128), (256, 32, 8, 128), (64, 8, 1, 128), (64, 16, 2, 128), (64, 32, 4, 64), (1024, 8, 8, 128)]
    too_big = [(64, 64, 8, 128), (4, 64, 8, 128), (64, 32, 8, 256)]
    for b, hq, hkv, d in fits:
        assert lib.b200attn_workspace_bytes(b, b, hq, hkv, d) <= limit, (b, hq, hkv, d)
    for b, hq, hkv, d in too_big:
        assert lib.b200attn_workspace_bytes(b, b, hq, hkv, d) > limit, (b, hq, hkv, d)

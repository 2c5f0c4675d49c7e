"""vLLM plugin entry points (reference: ibm-triton-lib/ibm_triton_lib/backend/__init__.py:20-22,
registered under `vllm.platform_plugins` in ibm-triton-lib/setup.py:70-72).

Two drop-in routes:
  * `register()`          — platform plugin: returns the platform class path; the
                            platform then picks this backend for every layer, as
                            the reference plugin does.
  * `register_backend()`  — general plugin: registers this backend as
                            AttentionBackendEnum.CUSTOM; select it with
                            `attention_config={"backend": "CUSTOM"}`.
"""
from __future__ import annotations


def register() -> str:
    """`vllm.platform_plugins` entry point."""
    return "vllm_triton_backend_b200.backend.platform.B200Platform"


def register_backend() -> None:
    """`vllm.general_plugins` entry point."""
    from vllm.v1.attention.backends.registry import AttentionBackendEnum, register_backend as _register

    from .attn import BACKEND_PATH

    _register(AttentionBackendEnum.CUSTOM, BACKEND_PATH)

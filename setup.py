"""Packaging for the vLLM plugin entry points (mirrors ibm-triton-lib/setup.py:70-72).
The CUDA library is built in-tree by `python -m vllm_triton_backend_b200.build`."""
from setuptools import setup

setup(
    name="vllm-triton-backend-b200",
    version="0.1.0",
    packages=["vllm_triton_backend_b200", "vllm_triton_backend_b200.backend"],
    package_data={"vllm_triton_backend_b200": ["libb200attn.so", "csrc/*"]},
    entry_points={
        "vllm.platform_plugins": ["b200_attn = vllm_triton_backend_b200.backend:register"],
        "vllm.general_plugins": ["b200_attn_backend = vllm_triton_backend_b200.backend:register_backend"],
    },
)

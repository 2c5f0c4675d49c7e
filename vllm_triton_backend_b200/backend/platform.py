"""Out-of-tree vLLM platform that always selects the b200attn backend.

Mirror of the reference's `TritonPlatform`
(ibm-triton-lib/ibm_triton_lib/backend/platform.py:57-72): a CUDA platform whose
`get_attn_backend_cls` returns this package's backend unconditionally.  The
signature is the one of the installed vLLM (0.22:
`get_attn_backend_cls(selected_backend, attn_selector_config, num_heads)`).
"""
from __future__ import annotations

from vllm.platforms.cuda import CudaPlatform

from .attn import BACKEND_PATH


class B200Platform(CudaPlatform):
    @classmethod
    def get_attn_backend_cls(cls, selected_backend, attn_selector_config=None, num_heads=None, *args, **kwargs) -> str:
        use_mla = getattr(attn_selector_config, "use_mla", False)
        if use_mla:
            raise NotImplementedError("b200attn implements standard (non-MLA) decoder attention only")
        return BACKEND_PATH

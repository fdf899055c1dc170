"""vq-vae_b200: the FoR-EST vector-quantisation hot path on B200 (sm_100a) kernels.

Importable as ``vqvae_b200`` (the directory name has a hyphen; ``vqvae_b200/__init__.py`` at the repo
root forwards here).  Only what the hot path needs lives in this package: ``csrc/`` (CUDA + C ABI),
``_lib`` (ctypes binding), ``ops`` (tensor -> pointer wrappers) and ``quantizer`` (the nn.Module the
reference scripts call).
"""
from .quantizer import CodebookManager, SplitQuantizer, VectorQuantizer, extract_code_usage_from_state  # noqa: F401

__all__ = ["VectorQuantizer", "SplitQuantizer", "CodebookManager", "extract_code_usage_from_state"]

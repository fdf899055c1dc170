"""vqvae_b200: the FoR-EST vector-quantisation hot path on B200 (sm_100a) kernels.

Only what the hot path needs lives in this package: ``csrc/`` (CUDA + C ABI), ``_lib`` (ctypes binding),
``ops`` (tensor -> pointer wrappers), ``quantizer`` (the nn.Module the reference scripts call), ``parallel``
(data-parallel statistics exchange) and ``knn`` (the mutual-kNN pair mining of the live training step).
(``vq-vae_b200`` at the repo root is a symlink to this directory: a hyphen cannot appear in a module name.)
"""
from .quantizer import CodebookManager, SplitQuantizer, VectorQuantizer, extract_code_usage_from_state  # noqa: F401

__all__ = ["VectorQuantizer", "SplitQuantizer", "CodebookManager", "extract_code_usage_from_state"]

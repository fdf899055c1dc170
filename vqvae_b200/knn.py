"""Mutual-kNN pair mining on the library's top-k kernels.

Same signature and return convention as the reference's `pairs_mutual_knn_chunked`
(frl/losses/pairs.py:531-610): `[P, 2]` int64 (anchor, target) pairs, both (i, j) and (j, i), ordered by anchor and
then by the neighbour's rank.  `chunk_size` is accepted for compatibility and ignored (the kernel never materialises a
[chunk, N] block).  CUDA only, no fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from . import _lib
from ._lib import check
from .ops import _On, _ptr


def _workspace(N: int, Cdim: int, device) -> torch.Tensor:
    out = C.c_size_t(0)
    check(_lib.load().vq_knn_workspace_bytes(int(N), int(Cdim), C.byref(out)), "vq_knn_workspace_bytes")
    return torch.empty(max(int(out.value), 256), dtype=torch.uint8, device=device)


def knn_topk(features: torch.Tensor, k: int, coords: Optional[torch.Tensor] = None,
             patch: Optional[torch.Tensor] = None, pos_min_spatial: float = 4.0, return_d2: bool = False):
    """[N, C] fp32 -> knn_idx [N, k] int64 (-1 = no valid neighbour) (+ squared distances)."""
    with _On(features, coords, patch) as on:
        X = features.detach().to(torch.float32).contiguous()
        N, Cdim = X.shape
        knn = torch.empty((N, k), dtype=torch.int64, device=X.device)
        d2 = torch.empty((N, k), dtype=torch.float32, device=X.device) if return_d2 else None
        ws = _workspace(N, Cdim, X.device)
        if coords is not None:
            coords = coords.detach().to(torch.float32).contiguous()
            patch = patch.to(torch.int32).contiguous()
        check(_lib.load().vq_knn_topk(_ptr(X), N, Cdim, _ptr(coords), _ptr(patch), float(pos_min_spatial), int(k),
                                      _ptr(knn), _ptr(d2), _ptr(ws), on.stream), "vq_knn_topk")
    return (knn, d2) if return_d2 else knn


def mutual_pairs(knn: torch.Tensor) -> torch.Tensor:
    with _On(knn) as on:
        N, k = knn.shape
        pairs = torch.empty((max(N * k, 1), 2), dtype=torch.int64, device=knn.device)
        n = torch.zeros(1, dtype=torch.int64, device=knn.device)
        ws = _workspace(N, 4, knn.device)
        check(_lib.load().vq_knn_mutual_pairs(_ptr(knn), N, k, _ptr(pairs), _ptr(n), _ptr(ws), on.stream),
              "vq_knn_mutual_pairs")
    return pairs[:int(n)]                     # the size is data dependent: one device -> host read, as in the reference


def pairs_mutual_knn_chunked(features: torch.Tensor, coord_list: Sequence[torch.Tensor], offsets: List[int], k: int,
                             pos_min_spatial: float = 4.0, chunk_size: int = 128) -> torch.Tensor:
    N = features.shape[0]
    if N == 0:
        return torch.empty((0, 2), dtype=torch.long, device=features.device)
    dev = features.device
    coords = torch.cat([c.to(dev).float() for c in coord_list], dim=0)
    sizes = torch.tensor([offsets[p + 1] - offsets[p] for p in range(len(coord_list))], device=dev)
    patch = torch.repeat_interleave(torch.arange(len(coord_list), device=dev), sizes)
    actual_k = min(k, N - 1)
    if actual_k <= 0:
        return torch.empty((0, 2), dtype=torch.long, device=dev)
    knn = knn_topk(features, actual_k, coords, patch, pos_min_spatial)
    return mutual_pairs(knn)

"""CPU oracle for the mutual-kNN pair mining row  --  TEST INFRASTRUCTURE ONLY.

PARITY PINNED: unlike the quantiser, this path exists in the reference (frl/losses/pairs.py:531-610
`pairs_mutual_knn_chunked`).  tests/golden/knn_*.npz were produced by importing that function from /root/reference
(tests/golden/make_knn_golden.py); tests/test_oracle_knn.py checks this numpy restatement against them, and the GPU
tests check the kernels against both.  Only tests/ may import this module.
"""
from __future__ import annotations

import numpy as np


def knn_topk(X: np.ndarray, k: int, coords=None, patch=None, pos_min_spatial: float = 4.0):
    """Follows pairs.py:566-594: d = cdist(X, X) (matmul form, fp32), self and within-patch close pairs -> inf,
    k smallest per row ascending (ties: lower index), -1 where the distance is inf."""
    X = np.asarray(X, np.float32)
    N = X.shape[0]
    n2 = (X * X).sum(1, dtype=np.float32)
    d2 = np.maximum(n2[:, None] + n2[None, :] - 2.0 * (X @ X.T), 0.0).astype(np.float32)
    d2[np.arange(N), np.arange(N)] = np.inf
    if coords is not None:
        c = np.asarray(coords, np.float32)
        sp = np.sqrt(((c[:, None, :] - c[None, :, :]) ** 2).sum(-1))
        same = np.asarray(patch)[:, None] == np.asarray(patch)[None, :]
        d2[same & (sp < pos_min_spatial)] = np.inf
    order = np.argsort(d2, axis=1, kind="stable")[:, :k]
    vals = np.take_along_axis(d2, order, 1)
    knn = np.where(np.isinf(vals), -1, order).astype(np.int64)
    return knn, vals


def mutual_pairs(knn: np.ndarray) -> np.ndarray:
    """pairs.py:596-610: (i, j) with j in knn[i] and i in knn[j]; ordered by i, then rank of j."""
    out = []
    for i in range(knn.shape[0]):
        for j in knn[i]:
            if j >= 0 and (knn[j] == i).any():
                out.append((i, int(j)))
    return np.asarray(out, np.int64).reshape(-1, 2)


def pairs_mutual_knn_chunked(features, coord_list, offsets, k, pos_min_spatial=4.0, chunk_size=128):
    N = features.shape[0]
    coords = np.concatenate([np.asarray(c, np.float32) for c in coord_list], 0)
    patch = np.concatenate([np.full(offsets[p + 1] - offsets[p], p) for p in range(len(coord_list))])
    knn, _ = knn_topk(features, min(k, N - 1), coords, patch, pos_min_spatial)
    return mutual_pairs(knn)

"""Golden vectors for the mutual-kNN row, produced by the REFERENCE's own function
(frl/losses/pairs.py:531-610) imported from /root/reference.  Run in the authoring container only:
    python tests/golden/make_knn_golden.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference/frl")
from losses.pairs import pairs_mutual_knn_chunked  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def case(name, sizes, C, k, min_sp, grid, seed, clustered=False):
    g = torch.Generator().manual_seed(seed)
    coord_list, feats = [], []
    for n in sizes:
        coord_list.append(torch.stack([torch.randint(0, grid, (n,), generator=g), torch.randint(0, grid, (n,), generator=g)], 1))
        f = torch.randn(n, C, generator=g)
        if clustered:                                   # a few tight clusters: many near neighbours inside a patch
            f = torch.randn(8, C, generator=g)[torch.randint(0, 8, (n,), generator=g)] + 0.05 * f
        feats.append(f)
    features = torch.cat(feats, 0)
    offsets = [0]
    for n in sizes:
        offsets.append(offsets[-1] + n)
    pairs = pairs_mutual_knn_chunked(features, coord_list, offsets, k, pos_min_spatial=min_sp, chunk_size=128)
    np.savez_compressed(os.path.join(HERE, f"knn_{name}.npz"), features=features.numpy(),
                        coords=torch.cat(coord_list, 0).numpy().astype(np.int32), offsets=np.asarray(offsets, np.int64),
                        k=np.int64(k), pos_min_spatial=np.float32(min_sp), pairs=pairs.numpy())
    print(name, "N", features.shape[0], "pairs", tuple(pairs.shape))


case("3x200_c16_k8", [200, 200, 200], 16, 8, 4.0, 64, 1)
case("5x300_c64_k16", [300] * 5, 64, 16, 4.0, 128, 2)
case("ragged_c32_k16", [17, 250, 1, 90], 32, 16, 6.0, 48, 3)
case("tiny_k_gt_n", [3, 2], 8, 16, 1.0, 32, 4)
case("all_masked", [10], 8, 4, 100.0, 3, 5)
case("clustered_c24_k12", [256, 256], 24, 12, 4.0, 40, 6, clustered=True)

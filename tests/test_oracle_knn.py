"""Mutual-kNN row: the numpy oracle against golden vectors produced by the REFERENCE's own
`pairs_mutual_knn_chunked` (tests/golden/make_knn_golden.py), and the CUDA kernels against both (-m gpu)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import knn_reference as KR

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "knn_*.npz")))


def _load(path):
    z = np.load(path)
    offsets = z["offsets"].tolist()
    coords = z["coords"].astype(np.float32)
    coord_list = [coords[offsets[p]:offsets[p + 1]] for p in range(len(offsets) - 1)]
    return z["features"], coord_list, offsets, int(z["k"]), float(z["pos_min_spatial"]), z["pairs"]


def _explain_difference(features, a, b):
    """Pairs that differ between two results must come from distance near-ties (relative gap < 1e-5)."""
    sa, sb = set(map(tuple, a.tolist())), set(map(tuple, b.tolist()))
    diff = sa ^ sb
    if not diff:
        return 0
    X = features.astype(np.float64)
    worst = 0.0
    for i, j in diff:
        d = ((X[i] - X) ** 2).sum(1)
        d[i] = np.inf
        dij = d[j]
        gaps = np.abs(np.sort(d) - dij) / max(dij, 1e-30)
        worst = max(worst, np.sort(gaps)[1])          # nearest OTHER distance to d[i, j]
    assert worst < 1e-5, f"{len(diff)} differing pairs not explained by near-ties (gap {worst:.2e})"
    return len(diff)


def test_golden_files_present():
    assert len(GOLDEN) >= 6


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_numpy_oracle_matches_reference_output(path):
    features, coord_list, offsets, k, min_sp, ref_pairs = _load(path)
    got = KR.pairs_mutual_knn_chunked(features, coord_list, offsets, k, min_sp)
    if got.shape == ref_pairs.shape and (got == ref_pairs).all():
        return                                            # identical, order included
    _explain_difference(features, got, ref_pairs)


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_cuda_matches_reference_output(path):
    from vqvae_b200 import knn
    features, coord_list, offsets, k, min_sp, ref_pairs = _load(path)
    dev = torch.device("cuda:0")
    got = knn.pairs_mutual_knn_chunked(torch.from_numpy(features).to(dev), [torch.from_numpy(c).to(dev) for c in coord_list],
                                       offsets, k, pos_min_spatial=min_sp).cpu().numpy()
    assert got.dtype == np.int64 and got.ndim == 2 and got.shape[1] == 2
    if got.shape == ref_pairs.shape and (got == ref_pairs).all():
        return
    _explain_difference(features, got, ref_pairs)
    # order rule: by anchor, then by neighbour rank (ascending distance)
    assert (np.diff(got[:, 0]) >= 0).all()


@pytest.mark.gpu
def test_cuda_topk_against_numpy_at_training_scale():
    """16 patches x 256 anchors, C = 64, k = 16 (frl/training/representation/step.py:734-741 scale)."""
    from vqvae_b200 import knn
    g = torch.Generator().manual_seed(11)
    P, n, C, k = 16, 256, 64, 16
    X = torch.randn(P * n, C, generator=g)
    coords = torch.stack([torch.randint(0, 128, (P * n,), generator=g), torch.randint(0, 128, (P * n,), generator=g)], 1)
    patch = torch.arange(P).repeat_interleave(n)
    got, d2 = knn.knn_topk(X.cuda(), k, coords.cuda(), patch.cuda(), 4.0, return_d2=True)
    ref, vals = KR.knn_topk(X.numpy(), k, coords.numpy(), patch.numpy(), 4.0)
    got, d2 = got.cpu().numpy(), d2.cpu().numpy()
    assert (np.diff(d2, axis=1) >= 0).all()                                  # ascending
    np.testing.assert_allclose(d2, vals, rtol=2e-5, atol=1e-4)
    bad = (got != ref)
    if bad.any():                                                            # only at near-ties
        rows = np.where(bad.any(1))[0]
        for r in rows:
            assert np.allclose(np.sort(d2[r]), np.sort(vals[r]), rtol=2e-5, atol=1e-4)
        assert bad.sum() <= 0.001 * bad.size
    pairs = knn.mutual_pairs(torch.from_numpy(got).cuda()).cpu().numpy()
    ref_pairs = KR.mutual_pairs(got)
    assert pairs.shape == ref_pairs.shape and (pairs == ref_pairs).all()


@pytest.mark.gpu
@pytest.mark.parametrize("N,C,k,masked", [(1000, 64, 16, True), (777, 36, 5, False), (4096, 128, 32, True), (300, 8, 16, True)])
def test_cuda_tiled_path_equals_streaming_path(N, C, k, masked, monkeypatch):
    """The tiled two-pass selection and the streaming sorted-list kernel share their arithmetic: results are bit-equal."""
    from vqvae_b200 import knn
    g = torch.Generator().manual_seed(N + C)
    X = torch.randn(N, C, generator=g).cuda()
    coords = torch.randint(0, 32, (N, 2), generator=g).float().cuda() if masked else None
    patch = torch.randint(0, 4, (N,), generator=g).sort().values.int().cuda() if masked else None
    a_idx, a_d = knn.knn_topk(X, k, coords, patch, 6.0, return_d2=True)
    monkeypatch.setenv("VQ_KNN_PATH", "stream")
    b_idx, b_d = knn.knn_topk(X, k, coords, patch, 6.0, return_d2=True)
    assert torch.equal(a_idx, b_idx) and torch.equal(a_d, b_d)
    ref, _ = KR.knn_topk(X.cpu().numpy(), k, None if coords is None else coords.cpu().numpy(),
                         None if patch is None else patch.cpu().numpy(), 6.0)
    assert (a_idx.cpu().numpy() != ref).mean() < 2e-3          # fp32 summation order: near-ties only


@pytest.mark.gpu
def test_cuda_masses_of_exact_ties_take_the_redo_path():
    """Duplicated feature rows: hundreds of candidates at exactly the k-th distance overflow the survivor buffer; the
    flagged rows are redone by the streaming kernel and the lowest indices win (torch.topk / stable argsort order)."""
    from vqvae_b200 import knn
    g = torch.Generator().manual_seed(5)
    base = torch.randint(-3, 4, (3, 16), generator=g).float()   # small integers: every sum is exact in fp32, ties are exact
    X = base[torch.arange(900) % 3].contiguous()                # 300 exact copies of each of 3 points
    idx, d2 = knn.knn_topk(X.cuda(), 8, return_d2=True)
    ref, vals = KR.knn_topk(X.numpy(), 8)
    assert np.array_equal(idx.cpu().numpy(), ref)
    assert float(d2.max()) == 0.0
    pairs = knn.mutual_pairs(idx).cpu().numpy()
    assert np.array_equal(pairs, KR.mutual_pairs(ref))

#!/usr/bin/env python
"""Mutual-kNN pair mining at the live training step's scale: fused top-k kernels vs an eager torch
restatement of the reference's chunked loop (frl/losses/pairs.py:531-610: cdist + masks + topk per 128-row chunk, then
the mutual filter) on the same B200, and the numpy oracle on the host.  Prints one JSON line.

    python tools/bench_knn.py [--patches 16] [--anchors 256] [--channels 64] [--k 16]
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def eager_pairs(features, coords, patch, k, min_sp, chunk=128):
    """What stock torch does for this function on a GPU: O(N/chunk) rounds of ~10 launches and a [chunk, N] matrix."""
    N = features.shape[0]
    k = min(k, N - 1)
    knn = torch.empty(N, k, dtype=torch.long, device=features.device)
    for s in range(0, N, chunk):
        e = min(N, s + chunk)
        d = torch.cdist(features[s:e], features)
        ar = torch.arange(s, e, device=features.device)
        d[ar - s, ar] = float("inf")
        sp = torch.cdist(coords[s:e], coords)
        d[(patch[s:e, None] == patch[None, :]) & (sp < min_sp)] = float("inf")
        v, i = d.topk(k, dim=1, largest=False)
        knn[s:e] = torch.where(torch.isinf(v), torch.full_like(i, -1), i)
    valid = knn >= 0
    safe = knn.clamp_min(0)
    me = torch.arange(N, device=features.device)[:, None, None]
    mutual = (knn[safe] == me).any(-1) & valid
    rows = torch.arange(N, device=features.device)[:, None].expand_as(knn)
    return torch.stack([rows[mutual], knn[mutual]], 1)


def timeit(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--patches", type=int, default=16)
    ap.add_argument("--anchors", type=int, default=256)
    ap.add_argument("--channels", type=int, default=64)
    ap.add_argument("--k", type=int, default=16)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    from vqvae_b200 import knn
    from oracle import knn_reference as KR
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(0)
    P, n, C = a.patches, a.anchors, a.channels
    N = P * n
    X = torch.randn(N, C, generator=g)
    coords = torch.stack([torch.randint(0, 128, (N,), generator=g), torch.randint(0, 128, (N,), generator=g)], 1).float()
    patch = torch.arange(P).repeat_interleave(n)
    offsets = [i * n for i in range(P + 1)]
    Xd, cd, pd = X.to(dev), coords.to(dev), patch.to(dev)
    clist = [cd[offsets[p]:offsets[p + 1]] for p in range(P)]

    ours = knn.pairs_mutual_knn_chunked(Xd, clist, offsets, a.k, 4.0)
    eager = eager_pairs(Xd, cd, pd, a.k, 4.0)
    same = ours.shape == eager.shape and bool((ours == eager).all())
    t_ours = timeit(lambda: knn.pairs_mutual_knn_chunked(Xd, clist, offsets, a.k, 4.0), a.iters)
    t_topk = timeit(lambda: knn.knn_topk(Xd, a.k, cd, pd, 4.0), a.iters)
    t_eager = timeit(lambda: eager_pairs(Xd, cd, pd, a.k, 4.0), max(3, a.iters // 4))
    t0 = time.perf_counter()
    ref = KR.pairs_mutual_knn_chunked(X.numpy(), [c.numpy() for c in (coords[offsets[p]:offsets[p + 1]] for p in range(P))],
                                      offsets, a.k, 4.0)
    t_cpu = (time.perf_counter() - t0) * 1e3
    diff = len(set(map(tuple, ours.cpu().numpy().tolist())) ^ set(map(tuple, ref.tolist())))
    print(json.dumps({"metric": "mutual-kNN pair mining", "N": N, "C": C, "k": a.k, "pairs": int(ours.shape[0]),
                      "ms_fused_total": t_ours, "ms_fused_topk_kernels": t_topk, "ms_torch_eager_same_gpu": t_eager,
                      "ms_numpy_oracle_cpu": t_cpu, "speedup_vs_eager": t_eager / t_ours,
                      "identical_to_eager": same, "pairs_differing_from_oracle": diff,
                      "flops_2N2C": 2.0 * N * N * C, "gflops_fused_topk": 2.0 * N * N * C / (t_topk * 1e-3) / 1e9}))


if __name__ == "__main__":
    main()

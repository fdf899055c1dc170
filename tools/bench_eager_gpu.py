#!/usr/bin/env python
"""GPU eager baseline (tier 2): the same VQ forward + EMA step written with stock PyTorch ops on the
same B200 -- chunked matmul + argmin (the [chunk, K] score matrix goes through HBM), index_select, elementwise
straight-through / loss, index_add_ statistics, elementwise EMA update.  This is what a user gets without this
library (the reference ships no GPU kernel of its own); it is a baseline, not a checker, and imports nothing from oracle/.

    python tools/bench_eager_gpu.py c2 c4
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


@torch.no_grad()
def eager_step(z, E, ema_cnt, ema_sum, decay=0.99, eps=1e-5, beta=0.25, chunk=65536):
    N, D = z.shape
    K = E.shape[0]
    Eb = E.to(torch.bfloat16)
    e2 = Eb.float().pow(2).sum(1)
    idx = torch.empty(N, dtype=torch.int64, device=z.device)
    for s in range(0, N, chunk):                                   # [chunk, K] fp32 scores: 256 MiB at K = 1024
        zc = z[s:s + chunk]
        sc = e2[None, :] - 2.0 * (zc @ Eb.t()).float()
        idx[s:s + chunk] = sc.argmin(1)
    q = E.index_select(0, idx).to(z.dtype)
    loss = beta * (z.float() - E.index_select(0, idx)).pow(2).mean()
    cnt = torch.bincount(idx, minlength=K).float()
    sums = torch.zeros(K, D, device=z.device).index_add_(0, idx, z.float())
    ema_cnt.mul_(decay).add_(cnt, alpha=1 - decay)
    ema_sum.mul_(decay).add_(sums, alpha=1 - decay)
    n = ema_cnt.sum()
    E.copy_(ema_sum / ((ema_cnt + eps) / (n + K * eps) * n)[:, None])
    p = cnt / cnt.sum().clamp_min(1)
    perplexity = torch.exp(-(p * torch.log(p + 1e-10)).sum())
    return q, loss, idx, perplexity


def run(name, wl, make_inputs, iters=10):
    """bench.py's `gpu_eager` record for one workload (current CUDA device)."""
    dev = torch.device("cuda", torch.cuda.current_device())
    z, E = make_inputs(wl, 0, dev)
    E = E.to(dev)
    ema_cnt, ema_sum = torch.ones(wl["K"], device=dev), E.clone()
    for _ in range(3):
        eager_step(z, E, ema_cnt, ema_sum)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        eager_step(z, E, ema_cnt, ema_sum)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    return {"impl": "torch eager ops on the same B200 (chunked bf16 matmul + argmin, index_select, index_add_)",
            "workload": wl["desc"], "ms_per_step": ms, "vectors_per_s": wl["N"] / (ms * 1e-3)}


def main():
    from bench import WORKLOADS, make_inputs
    for name in sys.argv[1:] or ["c2", "c4"]:
        print(json.dumps(run(name, WORKLOADS[name], make_inputs)), flush=True)


if __name__ == "__main__":
    main()

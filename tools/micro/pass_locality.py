"""Is the gather pass bound by DRAM locality?  Same kernel, same bytes: idx random (rows gathered in random order after the
code sort) vs idx = row // (N / K) (the sorted order IS the natural row order: sequential reads and writes).
    python tools/micro/pass_locality.py c2 c4"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from bench import WORKLOADS, make_inputs, time_kernel  # noqa: E402
from vqvae_b200 import ops  # noqa: E402
from vqvae_b200._lib import VQ_STAT_SUM_Z  # noqa: E402

dev = torch.device("cuda:0")
flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for name in sys.argv[1:] or ["c2", "c4"]:
    wl = WORKLOADS[name]
    N, K, D = wl["N"], wl["K"], wl["D"]
    z, E = make_inputs(wl, 0, dev)
    E = E.to(dev)
    ws = torch.empty(ops.workspace_bytes(N, K, D, z.dtype), dtype=torch.uint8, device=dev)
    q = torch.empty_like(z); sq = torch.empty(1, device=dev)
    cnt = torch.empty(K, device=dev); sums = torch.empty((K, D), device=dev)
    Eb = torch.empty((K, D), dtype=torch.bfloat16, device=dev)
    nrm = torch.empty(ops.norm_elems(K), device=dev)
    ops.prepare_codebook(E, "bf16", Eb, nrm)
    real = torch.empty(N, dtype=torch.int32, device=dev)
    ops.assign(z, E, Eb, nrm, "bf16", real, None, ws)
    p = torch.bincount(real.long(), minlength=K).float() / N
    print(f"{name}: nearest-code usage: perplexity {float(torch.exp(-(p * torch.log(p + 1e-10)).sum())):.1f} of {K}, "
          f"largest share {float(p.max()):.4f}, codes used {int((p > 0).sum())}", flush=True)
    for label, idx in (("nearest codes", real), ("nearest, shuffled", real[torch.randperm(N, device=dev)]),
                       ("uniform random codes", torch.randint(0, K, (N,), device=dev, dtype=torch.int32)),
                       ("codes in row order", (torch.arange(N, device=dev) // ((N + K - 1) // K)).to(torch.int32))):
        t = time_kernel(lambda: ops.gather_st_loss_scatter(z, idx, E, q, sq, cnt, sums, VQ_STAT_SUM_Z, False, ws), 10,
                        lambda: flush_buf.fill_(1))
        gb = N * (D * 4 + 4) / (t * 1e-3) / 1e9
        print(f"{name}: {label:20s} pass group {t * 1e3:7.1f} us  {gb:7.1f} GB/s", flush=True)

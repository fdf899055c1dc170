"""Launch the two hot kernels a few times on a named workload (for ncu captures).
    python tools/prof_kernels.py c4 [iters]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import WORKLOADS, make_inputs  # noqa: E402
from vqvae_b200 import ops  # noqa: E402
from vqvae_b200._lib import VQ_STAT_SUM_Z  # noqa: E402

wl = WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "c4"]
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda:0")
N, K, D = wl["N"], wl["K"], wl["D"]
z, E = make_inputs(wl, 0, dev)
E = E.to(dev)
Eb = ops.new_shadow(K, D, "bf16", dev)
nrm = torch.empty(ops.norm_elems(K), device=dev)
ops.prepare_codebook(E, "bf16", Eb, nrm)
ws = torch.zeros(ops.workspace_bytes(N, K, D, z.dtype), dtype=torch.uint8, device=dev)
idx = torch.empty(N, dtype=torch.int32, device=dev)
q = torch.empty_like(z)
sq = torch.empty(1, device=dev)
cnt = torch.empty(K, device=dev)
sums = torch.empty((K, D), device=dev)
for _ in range(iters):
    ops.assign(z, E, Eb, nrm, "bf16", idx, None, ws, presort=True, ws_clean=True)
    ops.gather_st_loss_scatter(z, idx, E, q, sq, cnt, sums, VQ_STAT_SUM_Z, False, ws, presorted=True)
torch.cuda.synchronize()
print("ok", wl["desc"], float(sq), int(cnt.sum()))

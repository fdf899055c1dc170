"""Time the two hot launch groups alone on named workloads (CUDA events, L2 flushed between launches).
    python tools/bench_kernels.py c2 c4 ..."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import WORKLOADS, make_inputs, time_kernel, load_peaks  # noqa: E402
from vqvae_b200 import ops  # noqa: E402
from vqvae_b200._lib import VQ_STAT_SUM_Z  # noqa: E402

dev = torch.device("cuda:0")
peaks = load_peaks()
flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
flush = lambda: flush_buf.fill_(1)
for name in sys.argv[1:] or ["c2", "c4"]:
    if name in WORKLOADS:
        wl = WORKLOADS[name]
    else:                                   # ad-hoc shape "N,K,D"
        n_, k_, d_ = (int(x) for x in name.split(","))
        wl = dict(N=n_, K=k_, D=d_, desc=name)
    N, K, D = wl["N"], wl["K"], wl["D"]
    z, E = make_inputs(wl, 0, dev)
    E = E.to(dev)
    Eb = torch.empty((K, D), dtype=torch.bfloat16, device=dev)
    nrm = torch.empty(ops.norm_elems(K), device=dev)
    ops.prepare_codebook(E, "bf16", Eb, nrm)
    ws = torch.empty(ops.workspace_bytes(N, K, D, z.dtype), dtype=torch.uint8, device=dev)
    idx = torch.empty(N, dtype=torch.int32, device=dev)
    q = torch.empty_like(z)
    sq = torch.empty(1, device=dev)
    cnt = torch.empty(K, device=dev)
    sums = torch.empty((K, D), device=dev)
    for _ in range(3):
        ops.assign(z, E, Eb, nrm, "bf16", idx, None, ws)
        ops.gather_st_loss_scatter(z, idx, E, q, sq, cnt, sums, VQ_STAT_SUM_Z, False, ws)
    torch.cuda.synchronize()
    ta = time_kernel(lambda: ops.assign(z, E, Eb, nrm, "bf16", idx, None, ws), 10, flush)
    tp = time_kernel(lambda: ops.gather_st_loss_scatter(z, idx, E, q, sq, cnt, sums, VQ_STAT_SUM_Z, False, ws), 10, flush)
    tf = 2.0 * N * K * D / (ta * 1e-3) / 1e12
    gb = (N * (D * 4 + 4) + K * D * 8 + K * 4) / (tp * 1e-3) / 1e9
    print(f"{name}: assign {ta*1e3:8.1f} us  {tf:7.1f} TF/s ({tf/peaks['tf_burst']:.3f} of burst)   "
          f"pass group {tp*1e3:8.1f} us  {gb:7.1f} GB/s ({gb/peaks['hbm']:.3f} of hbm)  env={os.environ.get('VQ_ASSIGN_EPI_WARPS','-')}",
          flush=True)

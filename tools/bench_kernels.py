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
    Eb = ops.new_shadow(K, D, "bf16", dev)
    nrm = torch.empty(ops.norm_elems(K), device=dev)
    ops.prepare_codebook(E, "bf16", Eb, nrm)
    ws = torch.empty(ops.workspace_bytes(N, K, D, z.dtype), dtype=torch.uint8, device=dev)
    wsc = torch.zeros_like(ws)              # VQ_SEARCH_WS_CLEAN contract: only presort / presorted calls touch it
    idx = torch.empty(N, dtype=torch.int32, device=dev)
    q = torch.empty_like(z)
    sq = torch.empty(1, device=dev)
    cnt = torch.empty(K, device=dev)
    sums = torch.empty((K, D), device=dev)
    for _ in range(3):
        ops.assign(z, E, Eb, nrm, "bf16", idx, None, ws)
        ops.gather_st_loss_scatter(z, idx, E, q, sq, cnt, sums, VQ_STAT_SUM_Z, False, ws)
    torch.cuda.synchronize()
    fns = [lambda: ops.assign(z, E, Eb, nrm, "bf16", idx, None, ws),
           lambda: ops.assign(z, E, Eb, nrm, "bf16", idx, None, wsc, presort=True, ws_clean=True)]
    evs = [[], []]
    for _ in range(20):                      # interleaved A/B: both see the same clocks
        for j in (0, 1):
            flush()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fns[j](); e1.record()
            evs[j].append((e0, e1))
    torch.cuda.synchronize()
    ta0, ta = (sum(a.elapsed_time(b) for a, b in ev) / len(ev) for ev in evs)
    if os.environ.get("VQ_BENCH_SKIP_PASS"):      # timing variants whose by-products are garbage
        print(f"{name}: assign {ta0*1e3:8.1f} us, search+presort {ta*1e3:8.1f} us", flush=True)
        continue
    tp0 = time_kernel(lambda: ops.gather_st_loss_scatter(z, idx, E, q, sq, cnt, sums, VQ_STAT_SUM_Z, False, ws), 10, flush)
    ops.assign(z, E, Eb, nrm, "bf16", idx, None, wsc, presort=True, ws_clean=True)
    tp = time_kernel(lambda: ops.gather_st_loss_scatter(z, idx, E, q, sq, cnt, sums, VQ_STAT_SUM_Z, False, wsc,
                                                        presorted=True), 10, flush)
    print(f"{name}: stand-alone calls: assign {ta0*1e3:8.1f} us, pass group {tp0*1e3:8.1f} us; "
          f"search+presort {ta*1e3:8.1f} us, presorted pass group {tp*1e3:8.1f} us", flush=True)
    tf = 2.0 * N * K * D / (ta * 1e-3) / 1e12
    gb = (N * (D * 2 * z.element_size() + 4) + K * D * 8 + K * 4) / (tp * 1e-3) / 1e9
    print(f"{name}: assign {ta*1e3:8.1f} us  {tf:7.1f} TF/s ({tf/peaks['tf_burst']:.3f} of burst)   "
          f"pass group {tp*1e3:8.1f} us  {gb:7.1f} GB/s ({gb/peaks['hbm']:.3f} of hbm)  env={os.environ.get('VQ_ASSIGN_EPI_WARPS','-')}",
          flush=True)

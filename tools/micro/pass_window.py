"""Experiment: does the gather pass get faster when rows are code-sorted inside ROW WINDOWS instead of globally?
The pass visits rows in code-sorted order (random 256/512-byte rows over the whole latent tensor).  Here the same pass is
called per window of T consecutive rows (each window sorted on its own, VQ_STAT_ACCUMULATE), and the pass_kernel device time
is summed from CUPTI records -- launch overheads of the extra calls are NOT in the figure, only the kernels' own time.
    python tools/micro/pass_window.py c2 c4 [--data trained]"""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from bench import WORKLOADS, load_peaks, make_inputs  # noqa: E402
from vqvae_b200 import ops  # noqa: E402
from vqvae_b200._lib import VQ_STAT_SUM_Z  # noqa: E402

dev = torch.device("cuda:0")
peaks = load_peaks()
data = "trained" if "--data" in sys.argv and sys.argv[sys.argv.index("--data") + 1] == "trained" else "gaussian"
for name in [a for a in sys.argv[1:] if a in WORKLOADS]:
    wl = WORKLOADS[name]
    N, K, D = wl["N"], wl["K"], wl["D"]
    z, E = make_inputs(wl, 0, dev, data=data)
    E = E.to(dev)
    Eb = ops.new_shadow(K, D, "bf16", dev)
    nrm = torch.empty(ops.norm_elems(K), device=dev)
    ops.prepare_codebook(E, "bf16", Eb, nrm)
    ws = torch.empty(ops.workspace_bytes(N, K, D, z.dtype), dtype=torch.uint8, device=dev)
    idx = torch.empty(N, dtype=torch.int32, device=dev)
    ops.assign(z, E, Eb, nrm, "bf16", idx, None, ws)
    q = torch.empty_like(z)
    sq = torch.empty(1, device=dev)
    cnt = torch.empty(K, device=dev)
    sums = torch.empty((K, D), device=dev)
    ref = None
    algo = N * (D * 4 + 4) + K * D * 8 + K * 4
    for T in [N, 1 << 19, 1 << 18, 1 << 17, 1 << 16, 1 << 15, 1 << 14, 1 << 13]:
        if T > N:
            continue
        bounds = [(a, min(N, a + T)) for a in range(0, N, T)]

        def run():
            for c, (a, b) in enumerate(bounds):
                ops.gather_st_loss_scatter(z[a:b], idx[a:b], E, q[a:b], sq, cnt, sums, VQ_STAT_SUM_Z, False, ws, accumulate=c > 0)
        run(); run()
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(5):
                run()
            torch.cuda.synchronize()
        t_pass = sum(e.time_range.end - e.time_range.start for e in prof.events()
                     if e.device_type == torch.autograd.DeviceType.CUDA and "pass_kernel" in e.name) / 5
        t_all = sum(e.time_range.end - e.time_range.start for e in prof.events()
                    if e.device_type == torch.autograd.DeviceType.CUDA and "vq::" in e.name) / 5
        if ref is None:
            ref = (cnt.clone(), sums.clone())
        ok = torch.equal(cnt, ref[0]) and float((sums - ref[1]).abs().max() / ref[1].abs().max()) < 1e-4
        print(f"{name} {data}: window {T:8d} rows ({len(bounds):3d} calls): pass_kernel {t_pass:8.1f} us = {algo / t_pass / 1e3:7.1f} GB/s "
              f"({algo / t_pass / 1e3 / peaks['hbm']:.3f} of hbm); all kernels of the group {t_all:8.1f} us; stats equal {ok}", flush=True)

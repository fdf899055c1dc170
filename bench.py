#!/usr/bin/env python
"""bench.py -- quantised latent vectors / second for the VQ hot path (forward + EMA), BASELINE.json's metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c1|c2|c3|c4|c5] [--impl native|reference]

One "step" = one pass of the hot path over one batch of synthetic latents: nearest-code search (a-1),
fused gather / straight-through / loss / code statistics pass (a-2..a-5), code-stat exchange when N > 1
(a-6), EMA codebook update + usage + perplexity (a-7..a-9) -- i.e. VectorQuantizer.forward in training mode,
quantizer="ema".  Workloads:
    c4 (default, BASELINE configs[3]) : N = 524 288 vectors/GPU (4 194 304 per 8-GPU step), K=8192, D=256, bf16 --
                                        the configuration north_star's >= 70 % tensor / >= 75 % HBM / near-linear
                                        8-GPU targets are quoted on
    c2 (BASELINE configs[1])          : N = 256*5*32*32 = 1 310 720 vectors/GPU, K=1024, D=128, bf16 (carried as the
                                        second record `c2` of the default line)
    c1 (BASELINE configs[0])          : N = 40 960, K=512, D=64 (the reference-scale case)
    c3 (BASELINE configs[2])          : type/phase split, two codebooks K=2048 D=128 (262 144 + 1 310 720 vectors/GPU)
    c5 (BASELINE configs[4])          : the batch-2048 step's quantiser load, 10 485 760 vectors/GPU, K=8192, D=256
Weak scaling: every rank runs the same per-GPU workload on its own shard; `value` is the whole-job rate.
Prints ONE JSON line on rank 0.  Keys beyond the base contract: `roofline` (search vs the measured bf16 burst peak),
`roofline_pass` (scatter + pass vs the measured HBM copy peak), `cpu_baseline`, `e2e` (training-step contract: latents
host -> device, loss + perplexity back), `e2e_full` (latents in, quantised rows + indices back to pinned host memory),
`c2` (the same measurements on BASELINE configs[1]), `sustained` (ms/step over a >= 2 s loop with its own clock
record), `gpu_eager` (the same step in stock PyTorch ops on the same GPU), `fp32_tier` (the 1e-5 search tier at this
workload), `dp_check` (N > 1: replicas bit-identical, rows counted).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "c1": dict(N=8 * 5 * 32 * 32, K=512, D=64, desc="C1: 5x32x32 patch, batch 8, K=512 D=64"),
    "c2": dict(N=256 * 5 * 32 * 32, K=1024, D=128, desc="C2: batch 256 of 5x32x32 latents, K=1024 D=128, bf16 fwd+EMA"),
    "c3": dict(N=256 * 5 * 32 * 32, N_type=256 * 32 * 32, K=2048, D=128,
               desc="C3: type/phase split, two codebooks K=2048 D=128; type 262144 + phase 1310720 vectors per step"),
    "c4": dict(N=524288, K=8192, D=256, desc="C4: K=8192 D=256, 524288 vectors per GPU (4M per 8-GPU step), bf16 fwd+EMA"),
    "c5": dict(N=2048 * 5120, K=8192, D=256, desc="C5: quantiser load of the batch-2048 step, 10485760 vectors, K=8192 D=256"),
}
METRIC = "quantised latent vectors/sec (VQ fwd+EMA)"
UNIT = "vectors/s"
CPU_SAMPLE = 32768          # vectors per step of the CPU legs (cpu_baseline AND --impl reference: one sample size)
TRAFFIC_SOURCE = "profiles/traffic.json (dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full capture " \
                 "committed with the profiles; not re-measured in this run)"


def load_traffic(workload_key, kernel):
    """Measured DRAM bytes per launch of `kernel` (ncu --set full capture, committed under profiles/), or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(p) as fh:
            return json.load(fh).get(workload_key, {}).get(kernel)
    except Exception:
        return None


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            d = json.load(fh)
        return dict(hbm=float(d["hbm_gbs"]), tf_burst=float(d["bf16_tflops"]),
                    tf_sust=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")   # B200_PROFILING.md fallback


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.samples, self.proc, self.thr, self.idx = [], None, None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            return self
        self.thr = threading.Thread(target=self._read, daemon=True)
        self.thr.start()
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def make_inputs(wl, rank, device, dtype=torch.bfloat16, data="gaussian"):
    """z ~ N(0,1) seeded 1234+rank; codebook rows from the same distribution seeded 4321.
    data="trained": the "trained-like" set, z = E[c] + 0.1 N(0,1) with c ~ Zipf(1.1) over the K codes (skewed usage:
    scatter contention on the hot codes, dead codes in the tail)."""
    g = torch.Generator(device="cpu").manual_seed(4321)
    E = torch.randn(wl["K"], wl["D"], generator=g)
    gz = torch.Generator(device=device).manual_seed(1234 + rank) if device != "cpu" else torch.Generator().manual_seed(1234 + rank)
    z = torch.randn(wl["N"], wl["D"], generator=gz, device=device, dtype=torch.float32)
    if data == "trained":
        p = 1.0 / torch.arange(1, wl["K"] + 1, dtype=torch.float64).pow(1.1)
        c = torch.multinomial((p / p.sum()).float().to(device), wl["N"], replacement=True, generator=gz)
        z = E.to(device)[c] + 0.1 * z
    return z.to(dtype), E


# ------------------------------------------------------------------------------------------------ CPU legs (oracle)
def oracle_step(R, z, E, state, K):
    """The same step on the CPU oracle (fp32 eager torch, chunked so the [chunk, K] matrix stays small)."""
    idx, _ = R.assign(z, E, "fp32", chunk=8192)
    q = R.gather(E, idx, z.dtype)
    sq = R.sq_err_sum(z, E, idx)
    cnt, sums = R.code_stats(z, idx, K)
    En, n, m = R.ema_update(E, state["n"], state["m"], cnt, sums, 0.99, 1e-5)
    state["n"], state["m"] = n, m
    R.perplexity(cnt)
    return En, q, sq


def cpu_inputs(wl):
    sample = min(wl["N"], CPU_SAMPLE)
    z = torch.randn(sample, wl["D"], generator=torch.Generator().manual_seed(1234))
    E = torch.randn(wl["K"], wl["D"], generator=torch.Generator().manual_seed(4321))
    return sample, z, E, {"n": torch.ones(wl["K"]), "m": E.clone()}


def cpu_baseline(wl, target_seconds=12.0):
    """Oracle ("port") on the host cores, on a bounded sample of the same workload."""
    from oracle import vq_reference as R
    torch.set_num_threads(os.cpu_count() or 1)
    sample, z, E, state = cpu_inputs(wl)
    oracle_step(R, z[:4096], E, dict(state), wl["K"])          # warm-up
    t0 = time.perf_counter(); reps = 0
    while True:
        oracle_step(R, z, E, state, wl["K"])
        reps += 1
        if time.perf_counter() - t0 > target_seconds or reps >= 50:
            break
    dt = time.perf_counter() - t0
    return {"value": sample * reps / dt, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{reps} x {sample} vectors of the workload (K={wl['K']}, D={wl['D']}), fp32 eager torch oracle, "
                      f"{dt:.1f} s"}


def run_reference(args, wl):
    """--impl reference: the reference tree has NO quantiser source, so the reference arm is the
    CPU oracle port, all host threads, each step a bounded sample of the workload (the same sample cpu_baseline takes)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import vq_reference as R
    torch.set_num_threads(os.cpu_count() or 1)
    sample, z, E, state = cpu_inputs(wl)
    for _ in range(args.warmup):
        oracle_step(R, z, E, state, wl["K"])
    t0 = time.perf_counter()
    for _ in range(args.steps):
        oracle_step(R, z, E, state, wl["K"])
    dt = time.perf_counter() - t0
    val = sample * args.steps / dt
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["desc"], "N_per_step_sampled": sample, "K": wl["K"], "D": wl["D"]},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                             "sample": f"{sample} of {wl['N']} vectors per step; reference has no quantiser source, "
                                       "oracle port timed instead"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ native arm
def time_kernel(fn, iters, flush, back_to_back=False):
    """Average device time of one launch group, CUDA events on the launching (current) stream.
    Default: every group has its own event pair with an L2 flush in front (right for a millisecond-long kernel).
    back_to_back: `iters` groups between ONE pair of events, after a flush and two untimed groups -- for groups of
    100-160 us, where the host's launch latency and the event records were 6-9 us of every sample.  fn(i) must alternate
    between two sets of latent / output buffers (i & 1): a set comes round again only after the other set's > 500 MB
    have gone through the 126 MB L2, so no group finds its inputs cached."""
    if back_to_back:
        flush()
        fn(0); fn(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters
    evs = []
    for _ in range(iters):
        flush()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        evs.append((e0, e1))
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in evs) / len(evs)


class Ctx:
    """Process-wide state of one bench run (rank, device, peaks, L2 flush buffer, barrier)."""

    def __init__(self):
        import torch.distributed as dist
        self.dist = dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py (native arm) needs a CUDA device; there is no CPU fallback")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.numa = None
        if self.world > 1:
            from vqvae_b200 import parallel
            self.numa = parallel.pin_to_gpu_numa(self.local)      # before any pinned allocation (first touch)
            dist.init_process_group("nccl", device_id=self.dev)
        self.peaks = load_peaks()
        self.flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)      # > 126 MB L2

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, ms: float) -> float:
        t = torch.tensor([ms], device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t)

    def flush(self):
        self.flush_buf.fill_(1)


def build_module(ctx, wl, E, args):
    from vqvae_b200 import VectorQuantizer
    mod = VectorQuantizer(wl["D"], wl["K"], beta=0.25, quantizer="ema", ema_decay=0.99, ema_eps=1e-5).to(ctx.dev)
    with torch.no_grad():
        mod.codebook.copy_(E); mod.ema_sum.copy_(E)
    return mod.train()


def measure(ctx, key, args, steps, warmup, full=True, clocks=True):
    """One workload: `value` (latents resident in HBM), per-kernel rooflines (rank 0) and -- full -- the host-buffer
    numbers.  Returns (record, module, rows this rank pushed through the module)."""
    from vqvae_b200 import ops
    from vqvae_b200._lib import VQ_STAT_SUM_Z
    wl = WORKLOADS[key]
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    N, K, D = wl["N"], wl["K"], wl["D"]
    z, E = make_inputs(wl, rank, dev, data=args.data)
    mod = build_module(ctx, wl, E, args)
    rows_done = 0
    n_type = wl.get("N_type", 0)           # C3: a second codebook for the type stream
    split = None
    if n_type:
        from vqvae_b200 import SplitQuantizer
        split = SplitQuantizer(D, D, K, channel_first=False, beta=0.25, quantizer="ema", ema_decay=0.99, ema_eps=1e-5,
                               overlap=os.environ.get("VQ_SPLIT_OVERLAP", "") == "1").to(dev)
        split.quant_phase = mod
        with torch.no_grad():
            split.quant_type.codebook.copy_(E.flip(0)); split.quant_type.ema_sum.copy_(E.flip(0))
        split.train()
        z_type = z[:n_type].clone()

    def step_resident():
        with torch.no_grad():
            return split(z_type, z) if split is not None else mod(z)

    # ---------------- value: inputs resident in HBM, larger than L2 (268 MB of latents per step at c4, 336 MB at c2)
    for _ in range(warmup):
        step_resident()
    rows_done += warmup * N
    ctx.barrier()
    sampler = ClockSampler(ctx.local).start() if clocks else None
    l0 = ops.launch_count() + mod.graph_launches        # eager launches + launches replayed inside CUDA graphs
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step_resident()
    e1.record()
    ctx.barrier()
    rows_done += steps * N
    launches = ops.launch_count() + mod.graph_launches - l0
    ms_total = ctx.max_over_ranks(e0.elapsed_time(e1))
    n_total = (N + n_type) * world
    rec = {"value": n_total * steps / (ms_total * 1e-3), "ms_per_step": ms_total / steps, "gpu_launches": int(launches)}

    # ---------------- e2e: HOST buffers through the public module call, H2D + D2H inside the timed region
    if full:
        z_host = z.cpu().pin_memory()
        res_host = torch.empty(2, dtype=torch.float32).pin_memory()
        if split is not None:
            zt_host = z_type.cpu().pin_memory()

        def step_e2e():
            with torch.no_grad():
                if split is not None:
                    split.forward_host(zt_host, z_host)
                    loss = split.loss
                else:                       # public host entry: row chunks cross PCIe while the previous chunk is quantised
                    _, loss, _ = mod.forward_host(z_host)
            res_host[0:1].copy_(loss.reshape(1), non_blocking=True)
            res_host[1:2].copy_(mod.last_perplexity, non_blocking=True)
            torch.cuda.current_stream().synchronize()             # the caller reads the step's result

        def timed(fn, n):
            ctx.barrier()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record()
            for _ in range(n):
                fn()
            f1.record()
            ctx.barrier()
            return ctx.max_over_ranks(f0.elapsed_time(f1))

        for _ in range(max(1, warmup // 2)):
            step_e2e()
        e2e_ms = timed(step_e2e, steps)
        rows_done += (max(1, warmup // 2) + steps) * N
        rec["e2e"] = {"value": n_total * steps / (e2e_ms * 1e-3), "unit": UNIT,
                      "h2d_bytes_per_step": (N + n_type) * D * 2, "d2h_bytes_per_step": 8, "ms_per_step": e2e_ms / steps,
                      "what": "training-step contract (scripts/train_vqvae.py:275-347): latents host -> device, loss + "
                              "perplexity back; quantised rows and indices stay on the device for the decoder"}
        if split is None:
            q_host = torch.empty(z_host.shape, dtype=z_host.dtype).pin_memory()
            i_host = torch.empty(N, dtype=torch.int64).pin_memory()

            def step_e2e_full():
                with torch.no_grad():
                    _, loss, _ = mod.forward_host(z_host, q_host=q_host, idx_host=i_host)
                res_host[0:1].copy_(loss.reshape(1), non_blocking=True)
                res_host[1:2].copy_(mod.last_perplexity, non_blocking=True)
                torch.cuda.current_stream().synchronize()

            step_e2e_full()
            nfull = max(5, steps // 2)
            full_ms = timed(step_e2e_full, nfull)
            rows_done += (1 + nfull) * N
            rec["e2e_full"] = {"value": N * world * nfull / (full_ms * 1e-3), "unit": UNIT,
                               "h2d_bytes_per_step": N * D * 2, "d2h_bytes_per_step": N * D * 2 + N * 8 + 8,
                               "ms_per_step": full_ms / nfull,
                               "what": "latents in, quantised rows + int64 indices + loss + perplexity back to pinned host "
                                       "memory (results stream back chunk by chunk under the remaining chunks)"}
            del q_host, i_host
        del z_host
    if sampler is not None:
        rec["clocks"] = sampler.stop()

    # ---------------- per-kernel timing for the roofline (rank 0, each kernel group timed alone; see time_kernel)
    if rank == 0:
        peaks = ctx.peaks
        Eb, nrm = mod._E_bf16, mod._e_norm2
        # private zero-filled workspace under the same contract the module uses (VQ_SEARCH_WS_CLEAN)
        ws = torch.zeros(ops.workspace_bytes(N, K, D, z.dtype), dtype=torch.uint8, device=dev)
        idx = torch.empty(N, dtype=torch.int32, device=dev)
        q = torch.empty_like(z)
        sq = torch.empty(1, device=dev)
        stats = torch.empty(K * D + K, dtype=torch.float32, device=dev)   # private: never the module's exchange buffer
        csum, ccnt = stats[:K * D].view(K, D), stats[K * D:]
        Ec = mod.codebook.detach()
        it = max(5, min(20, steps))
        # the pairing the module runs: the search leaves the counting-sort inputs in ws, the pass group picks them up
        t_assign = time_kernel(lambda: ops.assign(z, Ec, Eb, nrm, "bf16", idx, None, ws, presort=True, ws_clean=True),
                               it, ctx.flush)
        idx64 = torch.empty(N, dtype=torch.int64, device=dev)
        b2b = N * D * 2 > (128 << 20)
        if b2b:                                                   # second buffer set for the alternation (see time_kernel)
            zq = [(z, q), (z.clone(), torch.empty_like(z))]
            t_pass = time_kernel(lambda i: ops.gather_st_loss_scatter(zq[i & 1][0], idx, Ec, zq[i & 1][1], sq, ccnt, csum,
                                                                      VQ_STAT_SUM_Z, False, ws, idx64_out=idx64,
                                                                      presorted=True), 2 * it, ctx.flush, back_to_back=True)
            del zq
        else:
            t_pass = time_kernel(lambda: ops.gather_st_loss_scatter(z, idx, Ec, q, sq, ccnt, csum, VQ_STAT_SUM_Z, False, ws,
                                                                    idx64_out=idx64, presorted=True), it, ctx.flush)
        flops = 2.0 * N * K * D                                   # 2*K*D per vector
        pass_bytes = N * (D * 2 + 4 + D * 2) + K * D * 4 * 2 + K * 4   # read z + idx, write q; codebook + stats once
        tf = flops / (t_assign * 1e-3) / 1e12
        gbs = pass_bytes / (t_pass * 1e-3) / 1e9
        rec["roofline"] = {"kernel": "assign_tc_kernel (a-1, tcgen05; incl. the histogram / rank / scan by-product)",
                           "bound": "tensor", "achieved": tf, "peak": peaks["tf_burst"], "unit": "TFLOP/s",
                           "frac": tf / peaks["tf_burst"], "frac_of_sustained_peak": tf / peaks["tf_sust"],
                           "traffic": load_traffic(key, "assign_tc_kernel"), "traffic_source": TRAFFIC_SOURCE,
                           "algorithmic_bytes": N * (D * 2 + 4) + K * D * 2 + K * 4,
                           "peak_source": peaks["src"] + " burst (kernel timed alone)", "ms_per_launch": t_assign,
                           "frac_of_nominal_dense_bf16": tf / 2250.0,
                           "peak_note": "peak = the pool's measured torch.matmul bf16 throughput (MEASURED_PEAKS.json), not a "
                                        "hardware limit: a CTA-pair tcgen05 kernel that keeps its operands in shared memory "
                                        "can exceed it (frac > 1), the nominal dense bf16 figure is 2250 TFLOP/s"}
        rec["roofline_pass"] = {"kernel": "scatter_ranked_kernel + pass_kernel (a-2..a-5; histogram / ranks / scan come "
                                          "from the search kernel; the scatter also widens the indices to int64)", "bound": "hbm", "achieved": gbs, "peak": peaks["hbm"],
                                "unit": "GB/s", "frac": gbs / peaks["hbm"], "traffic": load_traffic(key, "pass_group"),
                                "traffic_source": TRAFFIC_SOURCE, "algorithmic_bytes": pass_bytes,
                                "peak_source": peaks["src"], "ms_per_launch": t_pass,
                                "timing": "groups back to back between one event pair, alternating between two latent / output buffer sets" if b2b
                                else "one event pair per group, L2 flushed in front"}
        del ws, idx, q, stats
    return rec, mod, z, rows_done


def sustained(ctx, mod, z, seconds=2.0):
    """The resident step back to back for >= `seconds` (burst clocks decay to the power-capped steady state)."""
    N = z.shape[0]
    with torch.no_grad():
        for _ in range(3):
            mod(z)
    ctx.barrier()
    sampler = ClockSampler(ctx.local).start()
    # the loop length is fixed up front (same on every rank): a first short run sizes it
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    with torch.no_grad():
        for _ in range(20):
            mod(z)
    e1.record()
    torch.cuda.synchronize()
    est = ctx.max_over_ranks(e0.elapsed_time(e1)) / 20
    steps = int(min(20000, max(50, seconds * 1e3 / max(est, 1e-3))))
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    with torch.no_grad():
        for _ in range(steps):
            mod(z)
    e1.record()
    ctx.barrier()
    ms = ctx.max_over_ranks(e0.elapsed_time(e1))
    clk = sampler.stop()
    return {"ms_per_step": ms / steps, "value": N * ctx.world * steps / (ms * 1e-3), "unit": UNIT, "steps": steps,
            "seconds": ms * 1e-3, "clocks": clk}, (23 + steps) * N


def gpu_eager(key):
    """The same step in stock PyTorch ops on this GPU (tools/bench_eager_gpu.py): what a user gets without the library."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import bench_eager_gpu
    return bench_eager_gpu.run(key, WORKLOADS[key], make_inputs)


def fp32_tier(ctx, key, mod, z, iters=5):
    """The 1e-5 search tier (search="fp32") on this workload's shape: fp32 latents, search kernel timed alone."""
    from vqvae_b200 import ops
    wl = WORKLOADS[key]
    N, K, D = wl["N"], wl["K"], wl["D"]
    dev = ctx.dev
    zf = z.float()
    E = mod.codebook.detach()
    nrm = torch.empty(ops.norm_elems(K), dtype=torch.float32, device=dev)
    out = {}
    try:
        Eb = ops.new_shadow(K, D, "fp32", dev)
        ops.prepare_codebook(E, "fp32", Eb, nrm)
        ws = torch.zeros(ops.workspace_bytes(N, K, D, zf.dtype, "fp32"), dtype=torch.uint8, device=dev)
        idx = torch.empty(N, dtype=torch.int32, device=dev)
        idx2 = torch.empty(N, dtype=torch.int32, device=dev)
        fn = lambda: ops.assign(zf, E, Eb, nrm, "fp32", idx, None, ws, presort=True, ws_clean=True)
        fn(); torch.cuda.synchronize()
        t = time_kernel(fn, iters, ctx.flush)
        fn2 = lambda: ops.assign(zf, E, None, nrm, "fp32_simt", idx2, None, ws, presort=True, ws_clean=True)
        fn2(); torch.cuda.synchronize()
        t2 = time_kernel(fn2, 2, ctx.flush)
        tf = 2.0 * N * K * D / (t * 1e-3) / 1e12
        out = {"ms_per_launch": t, "algorithmic_tflops": tf, "latents": "fp32",
               "kernel": "split_bf16 staging + assign_tc_kernel, split operands (z = hi + lo, E = hi + lo, 3 bf16 MMAs per "
                         "product, fp32 accumulate in TMEM)",
               "cuda_core_kernel_ms": t2, "index_disagreements_vs_cuda_core_kernel": int((idx != idx2).sum()),
               "note": "search='fp32' (1e-5 tier) timed alone incl. the staging pass, 2*N*K*D algorithmic flops"}
    except Exception as exc:                       # never fail the headline because of the side record
        out = {"error": str(exc)[:300]}
    return out


def run_native(args):
    ctx = Ctx()
    key = args.workload
    wl = WORKLOADS[key]
    rec, mod, z, rows = measure(ctx, key, args, args.steps, args.warmup, full=True)
    extras = {}
    if not args.no_extras:
        if key in ("c2", "c4") :
            srec, n = sustained(ctx, mod, z)
            rows += n
            extras["sustained"] = srec
    # ---------------- multi-GPU correctness the driver can see
    if ctx.world > 1:
        from vqvae_b200 import parallel
        dp = {"replicas_identical": False, "rows_counted": None, "rows_expected": None, "exchange": mod.peer_exchange
              if mod._peers is None else "peer (" + mod._peers.kind + ")", "exchange_timing_us": mod.peer_exchange_timing()}
        try:
            parallel.assert_replicas_identical(mod.codebook)
            dp["replicas_identical"] = True
        except RuntimeError as exc:
            dp["error"] = str(exc)
        tot = torch.tensor([rows], dtype=torch.int64, device=ctx.dev)
        ctx.dist.all_reduce(tot)
        dp["rows_expected"] = int(tot)
        dp["rows_counted"] = int(mod.usage_total.sum())
        dp["ok"] = dp["replicas_identical"] and dp["rows_counted"] == dp["rows_expected"]
        extras["dp_check"] = dp
    else:
        extras["dp_check"] = "skipped: 1 GPU"
        assert int(mod.usage_total.sum()) == rows, "usage_total does not add up to the rows quantised"
    if mod._peers is not None:
        mod._peers.close()
    if not args.no_extras and key == "c4":
        if ctx.rank == 0:
            extras["fp32_tier"] = fp32_tier(ctx, key, mod, z)
        del mod, z
        torch.cuda.empty_cache()
        c2rec, c2mod, c2z, _ = measure(ctx, "c2", args, args.steps, args.warmup, full=(ctx.world == 1), clocks=False)
        c2rec["config"] = {"workload": WORKLOADS["c2"]["desc"], "N_per_gpu": WORKLOADS["c2"]["N"], "K": 1024, "D": 128}
        c2rec["unit"] = UNIT
        if ctx.world == 1:
            s2, _ = sustained(ctx, c2mod, c2z)
            c2rec["sustained"] = s2
        if c2mod._peers is not None:
            c2mod._peers.close()
        extras["c2"] = c2rec
        del c2mod, c2z
        torch.cuda.empty_cache()
    if ctx.rank == 0 and ctx.world == 1 and not args.no_extras and key in ("c2", "c4"):
        try:
            extras["gpu_eager"] = {key: gpu_eager(key)}
            if key == "c4":
                extras["gpu_eager"]["c2"] = gpu_eager("c2")
        except Exception as exc:
            extras["gpu_eager"] = {"error": str(exc)[:300]}
    if ctx.rank == 0:
        extras["cpu_baseline"] = cpu_baseline(wl) if (ctx.world == 1 and not args.skip_cpu) else None
    dp_ok = True
    if ctx.world > 1:
        dp_ok = bool(extras["dp_check"]["ok"])
        ctx.dist.barrier()
        ctx.dist.destroy_process_group()
    if ctx.rank != 0:
        sys.exit(0 if dp_ok else 3)
    N, K, D = wl["N"], wl["K"], wl["D"]
    line = {"metric": METRIC, "value": rec.pop("value"), "unit": UNIT, "n_gpus": ctx.world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": rec.pop("ms_per_step"), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic" if args.data == "gaussian" else "synthetic (trained-like: Zipf(1.1) code usage, sigma 0.1)",
            "config": {"workload": wl["desc"], "N_per_gpu": N + wl.get("N_type", 0), "K": K, "D": D, "quantizer": "ema",
                       "parallelism": f"dp{ctx.world}", "numa": ctx.numa,
                       "l2": "inputs larger than L2 (%.0f MB of latents per step)" % (N * D * 2 / 1e6)},
            "headline_e2e": "e2e (training-step contract); e2e_full also returns quantised rows and indices to the host"}
    line.update(rec)
    line.update(extras)
    print(json.dumps(line), flush=True)
    if not dp_ok:
        sys.exit(3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="c4")
    ap.add_argument("--impl", choices=["native", "reference"], default="native")
    ap.add_argument("--data", choices=["gaussian", "trained"], default="gaussian",
                    help="synthetic latents: N(0,1) (default) or trained-like (Zipf-skewed codes + noise)")
    ap.add_argument("--skip-cpu", action="store_true", help="omit the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-extras", action="store_true",
                    help="only the base line of the chosen workload (no c2 record, sustained, gpu_eager, fp32_tier)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup
    if args.impl == "reference":
        run_reference(args, WORKLOADS[args.workload])
    else:
        run_native(args)


if __name__ == "__main__":
    main()

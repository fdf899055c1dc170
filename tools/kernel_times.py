#!/usr/bin/env python
"""Per-kernel device times (and, under torchrun, a per-rank TIMELINE) of the quantiser step, from CUPTI activity records
(torch.profiler) -- no ncu serialisation, the kernels run back to back as in production.

    python tools/kernel_times.py c4 [--data trained] [--steps 20] [--mode sort|det]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 \
        tools/kernel_times.py c2 --exchange peer --json gpurun_out/timeline8_c2.json

Single process: a table "kernel, launches, mean / min / max us" for the step's kernels.
Multi process: additionally one step's timeline per rank -- start and end of every kernel relative to the earliest
kernel start of that step over all ranks (CUPTI timestamps are host-clock aligned, good to a few microseconds) -- so
that straggle (when does each rank's pass end), the wait inside the exchange kernels (their duration minus the
single-GPU duration) and the tail (apply end) can be read apart.
--mode sort: stand-alone ops (vq_assign, then the pass with its own counting sort) instead of the module;
--mode det:  the module in deterministic mode (stable sort: rows of a code in row order)."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import WORKLOADS, make_inputs  # noqa: E402
from vqvae_b200 import VectorQuantizer, ops  # noqa: E402
from vqvae_b200._lib import VQ_STAT_SUM_Z  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("workload")
ap.add_argument("--data", default="gaussian")
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--mode", default="module", choices=["module", "sort", "det", "presort"])
ap.add_argument("--pdl", type=int, default=31, help="programmatic-dependent-launch mask; 0 = plain stream order, so that "
                "a kernel's duration holds no wait for its predecessor")
ap.add_argument("--exchange", default="auto")
ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp16", "fp32"])
ap.add_argument("--search", default="auto")
ap.add_argument("--json", default=None)
ap.add_argument("--cf", type=int, default=0, help="channel-first call: latents as [N / S, D, S] with S = this value (e.g. 1024)")
args = ap.parse_args()

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
if args.workload in WORKLOADS:
    wl = WORKLOADS[args.workload]
else:
    n_, k_, d_ = (int(x) for x in args.workload.split(","))
    wl = dict(N=n_, K=k_, D=d_, desc=args.workload)
N, K, D = wl["N"], wl["K"], wl["D"]
dt = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.float32}[args.dtype]
z, E = make_inputs(wl, rank, dev, dtype=dt, data=args.data)
mod = VectorQuantizer(D, K, quantizer="ema", peer_exchange=args.exchange, deterministic=args.mode == "det",
                      search=args.search, channel_first=args.cf > 0).to(dev).train()
if args.cf > 0:
    z = z.view(N // args.cf, args.cf, D).transpose(1, 2).contiguous()          # [B, D, S]
with torch.no_grad():
    mod.codebook.copy_(E); mod.ema_sum.copy_(E)

ops.set_pdl_mask(args.pdl)
if args.mode in ("sort", "presort"):
    Ed = E.to(dev)
    Eb = ops.new_shadow(K, D, "bf16", dev)
    nrm = torch.empty(ops.norm_elems(K), device=dev)
    ops.prepare_codebook(Ed, "bf16", Eb, nrm)
    ws = torch.zeros(ops.workspace_bytes(N, K, D, z.dtype, "bf16"), dtype=torch.uint8, device=dev)
    idx = torch.empty(N, dtype=torch.int32, device=dev)
    q = torch.empty_like(z)
    sq = torch.empty(1, device=dev)
    cnt = torch.empty(K, device=dev)
    sums = torch.empty((K, D), device=dev)

    if args.mode == "sort":           # stand-alone calls: the pass builds its own (block-ordered) counting sort
        def step():
            ops.assign(z, Ed, Eb, nrm, "bf16", idx, None, ws)
            ops.gather_st_loss_scatter(z, idx, Ed, q, sq, cnt, sums, VQ_STAT_SUM_Z, False, ws)
    else:                             # the production pairing through the same ops (arrival-ordered ranks from the search)
        def step():
            ops.assign(z, Ed, Eb, nrm, "bf16", idx, None, ws, presort=True, ws_clean=True)
            ops.gather_st_loss_scatter(z, idx, Ed, q, sq, cnt, sums, VQ_STAT_SUM_Z, False, ws, presorted=True)
else:
    def step():
        with torch.no_grad():
            mod(z)

for _ in range(5):
    step()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
from torch.profiler import ProfilerActivity, profile  # noqa: E402

mod.peer_exchange_timing(reset=True)

with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(args.steps):
        step()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "memcpy" not in e.name.lower()
       and "memset" not in e.name.lower()]
evs.sort(key=lambda e: e.time_range.start)


def short(name):
    name = name.replace("vq::(anonymous namespace)::", "").replace("void ", "")
    depth = 0
    for i, ch in enumerate(name):
        depth += ch == "<"
        depth -= ch == ">"
        if ch == "(" and depth == 0:
            return name[:i]
    return name


agg = {}
for e in evs:
    d = e.time_range.end - e.time_range.start
    a = agg.setdefault(short(e.name), [])
    a.append(d)
table = {k: {"n": len(v), "mean_us": sum(v) / len(v), "min_us": min(v), "max_us": max(v)} for k, v in agg.items()}
# one step's timeline: the LAST complete step (kernels between two consecutive search kernels)
starts = [i for i, e in enumerate(evs) if "assign_" in e.name]
tl = []
if len(starts) >= 2:
    a, b = starts[-2], starts[-1]
    tl = [(short(e.name), e.time_range.start, e.time_range.end) for e in evs[a:b]]
out = {"rank": rank, "table": table, "timeline": tl, "peer_timing": mod.peer_exchange_timing(),
       "exchange": None if mod._peers is None else mod._peers.kind}
if world > 1:
    allout = [None] * world
    dist.all_gather_object(allout, out)
else:
    allout = [out]
if rank == 0:
    print(f"# {wl['desc']}  data={args.data} mode={args.mode} dtype={args.dtype} world={world} exchange={args.exchange}")
    step_us = 0.0
    for k, v in sorted(table.items(), key=lambda kv: -kv[1]["mean_us"] * kv[1]["n"]):
        print(f"{k:70s} n={v['n']:4d} mean {v['mean_us']:9.1f} us  min {v['min_us']:9.1f}  max {v['max_us']:9.1f}")
        step_us += v["mean_us"] * v["n"] / args.steps
    print(f"sum of kernel time per step: {step_us:.1f} us")
    if world > 1:
        print("# exchange:", allout[0]["exchange"])
        print("# in-kernel timing of the fused exchange per rank, mean over the profiled steps (wait for the last peer's signal / sum the count vectors):",
              [(o["rank"], o["peer_timing"]) for o in allout])
        t0 = min(t[1] for o in allout for t in o["timeline"])
        print("# timeline of one step: rank, kernel, start us, end us (relative to the earliest start over ranks)")
        for o in allout:
            for (nm, s, e) in o["timeline"]:
                print(f"rank {o['rank']}  {nm:55s} {s - t0:9.1f} {e - t0:9.1f}   ({e - s:8.1f} us)")
        ends = {}
        for o in allout:
            for (nm, s, e) in o["timeline"]:
                ends.setdefault(nm, []).append((s - t0, e - t0))
        print("# per kernel over ranks: earliest start, latest start, earliest end, latest end")
        for nm, v in ends.items():
            print(f"{nm:55s} start {min(x[0] for x in v):8.1f} .. {max(x[0] for x in v):8.1f}   end {min(x[1] for x in v):8.1f} .. "
                  f"{max(x[1] for x in v):8.1f}")
    if args.json:
        with open(args.json, "w") as fh:
            json.dump({"workload": wl["desc"], "world": world, "exchange": args.exchange, "ranks": allout}, fh, indent=1)
if mod._peers is not None:
    mod._peers.close()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()

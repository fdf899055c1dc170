"""2+ rank check of the fused peer-memory exchange (a-6 fused into a-7) against the NCCL all-reduce path.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/dp_check.py
Prints one line per check on rank 0 and exits non-zero on a mismatch."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vqvae_b200 import VectorQuantizer, parallel  # noqa: E402

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
world = dist.get_world_size()
ok = True


def say(*a):
    if rank == 0:
        print(*a, flush=True)


for (N, K, D, dt) in [(4096, 256, 64, torch.float32), (100000, 1024, 128, torch.bfloat16), (65536, 8192, 256, torch.bfloat16)]:
    torch.manual_seed(7)
    E0 = torch.randn(K, D)
    mods = {}
    for mode in ("peer", "nccl"):
        m = VectorQuantizer(D, K, quantizer="ema", peer_exchange=mode, deterministic=True).to(dev)
        with torch.no_grad():
            m.codebook.copy_(E0); m.ema_sum.copy_(E0)
        mods[mode] = m.train()
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    for step in range(4):
        z = torch.randn(N + 64 * rank, D, generator=g, device=dev).to(dt)      # ragged shards
        outs = {k: m(z) for k, m in mods.items()}
        assert torch.equal(outs["peer"][2], outs["nccl"][2]) or step > 0
    torch.cuda.synchronize()
    a, b = mods["peer"], mods["nccl"]
    err = float((a.codebook - b.codebook).abs().max() / b.codebook.abs().max())
    same_usage = bool(torch.equal(a.usage_total, b.usage_total))
    pp = abs(float(a.last_perplexity) - float(b.last_perplexity)) / float(b.last_perplexity)
    try:
        parallel.assert_replicas_identical(a.codebook)
        identical = True
    except RuntimeError:
        identical = False
    tot = int(a.usage_total.sum())
    expect = 4 * sum(N + 64 * r for r in range(world))
    good = err < 2e-6 and same_usage and pp < 1e-5 and identical and tot == expect
    ok &= good
    say(f"N={N} K={K} D={D} {dt}: peer-vs-nccl codebook rel err {err:.2e}, usage equal {same_usage}, perplexity rel {pp:.1e}, "
        f"replicas bit-identical {identical}, rows counted {tot}/{expect} -> {'OK' if good else 'MISMATCH'}")

# timing: C2-shaped step, both paths
N, K, D = 1310720, 1024, 128
quick = "--quick" in sys.argv
z = torch.randn(16 if quick else N, D, device=dev, dtype=torch.bfloat16)
for mode in (() if quick else ("nccl", "peer")):
    m = VectorQuantizer(D, K, quantizer="ema", peer_exchange=mode).to(dev).train()
    with torch.no_grad():
        for _ in range(5):
            m(z)
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(30):
            m(z)
        e1.record()
        dist.barrier(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 30], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    say(f"C2 step, {world} ranks, exchange={mode}: {float(t):.4f} ms/step")
    if m._peers is not None:
        m._peers.close()
for m in mods.values():
    if m._peers is not None:
        m._peers.close()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)

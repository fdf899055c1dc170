#!/usr/bin/env python
"""BASELINE config 5 as a TRAIN STEP: the loop body of scripts/train_vqvae.py:261-347 -- conv encoder
-> quantiser -> conv decoder heads in torch, MSE + CE + canopy + vq loss, backward, AdamW -- around the native quantiser
(K=8192, D=256, autocast), at the largest batch toward 2048 patches of 5x32x32 that fits, and the same step with the
quantiser written in stock PyTorch ops (what a user has without the library).  Reports ms/step, the quantiser's share of
the step's kernel time (CUPTI) and vectors/s.

    python tools/bench_train_step.py [--batch 2048] [--steps 5] [--quantizer native|eager|both] [--dtype bf16|fp16]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29577 \
        tools/bench_train_step.py --batch 256      # per-rank batch; DDP averages the conv weights, the quantiser exchanges
                                                   # code statistics (a-6)
"""
import argparse
import json
import os
import sys

import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
from vqvae_model import IGNORE_INDEX, VQVAE, SyntheticPatches  # noqa: E402


class EagerQuantizer(nn.Module):
    """EMA quantiser in stock PyTorch ops (baseline only): chunked bf16 matmul + argmin (the [chunk, K] score matrix goes
    through HBM), index_select, straight-through, commitment loss, bincount / index_add_ statistics, elementwise EMA."""

    def __init__(self, emb_dim, codebook_size, beta=0.25, decay=0.99, eps=1e-5, chunk=65536):
        super().__init__()
        self.emb_dim, self.codebook_size, self.beta, self.decay, self.eps, self.chunk = emb_dim, codebook_size, beta, decay, eps, chunk
        E = torch.randn(codebook_size, emb_dim)
        self.register_buffer("codebook", E)
        self.register_buffer("ema_cnt", torch.ones(codebook_size))
        self.register_buffer("ema_sum", E.clone())
        self.register_buffer("last_perplexity", torch.ones(1))

    @property
    def perplexity(self):
        return self.last_perplexity[0]

    def attach_codebook_manager(self, m):
        pass

    def forward(self, z):
        lead, D, K = z.shape[:-1], self.emb_dim, self.codebook_size
        z2 = z.reshape(-1, D)
        with torch.no_grad():
            Eb = self.codebook.to(torch.bfloat16)
            e2 = Eb.float().pow(2).sum(1)
            idx = torch.empty(z2.shape[0], dtype=torch.int64, device=z.device)
            zb = z2.detach().to(torch.bfloat16)
            for s in range(0, z2.shape[0], self.chunk):
                idx[s:s + self.chunk] = (e2[None, :] - 2.0 * (zb[s:s + self.chunk] @ Eb.t()).float()).argmin(1)
            q = self.codebook.index_select(0, idx)
        loss = self.beta * (z2.float() - q).pow(2).mean()
        out = z2 + (q.to(z2.dtype) - z2).detach()
        if self.training:
            with torch.no_grad():
                cnt = torch.bincount(idx, minlength=K).float()
                sums = torch.zeros(K, D, device=z.device).index_add_(0, idx, z2.detach().float())
                if torch.distributed.is_initialized() and torch.distributed.get_world_size() > 1:
                    torch.distributed.all_reduce(cnt); torch.distributed.all_reduce(sums)
                self.ema_cnt.mul_(self.decay).add_(cnt, alpha=1 - self.decay)
                self.ema_sum.mul_(self.decay).add_(sums, alpha=1 - self.decay)
                n = self.ema_cnt.sum()
                self.codebook.copy_(self.ema_sum / ((self.ema_cnt + self.eps) / (n + K * self.eps) * n)[:, None])
                p = cnt / cnt.sum().clamp_min(1)
                self.last_perplexity[0] = torch.exp(-(p * torch.log(p + 1e-10)).sum())
        return out.reshape(*lead, D), loss, idx.reshape(lead)


def run(args, kind, batch, dev, world, rank, local):
    torch.manual_seed(0)
    model = VQVAE(cont_dim=args.cont_dim, cat_vocab_sizes={"lc": 12}, emb_dim=args.emb_dim, codebook_size=args.codebook_size,
                  hidden=args.hidden, quantizer="ema")
    if kind == "eager":
        model.quant = EagerQuantizer(args.emb_dim, args.codebook_size)
    model = model.to(dev).train()
    if world > 1 and kind == "native":
        model.quant.sync_state_from(0)
    net = nn.parallel.DistributedDataParallel(model, device_ids=[local], broadcast_buffers=False) if world > 1 else model
    ds = SyntheticPatches(cont_dim=args.cont_dim, T=5, H=32, W=32, seed=rank, device=dev)
    opt = torch.optim.AdamW([p for p in model.parameters() if p.requires_grad], lr=3e-4, weight_decay=1e-4, betas=(0.9, 0.95))
    adt = torch.bfloat16 if args.dtype == "bf16" else torch.float16
    scaler = torch.amp.GradScaler("cuda", enabled=args.dtype == "fp16")
    data = ds.batch(batch)

    def step():
        opt.zero_grad(set_to_none=True)
        with torch.autocast(device_type="cuda", dtype=adt):
            cont_pred, cat_logits, canopy_pred, vq_loss, pplx = net(data)
            loss = F.mse_loss(cont_pred.float(), data["cont"]) + F.mse_loss(canopy_pred.float(), data["canopy"]) + vq_loss
            for j, name in enumerate(ds.cat_names):
                lg = cat_logits[name]
                loss = loss + F.cross_entropy(lg.reshape(-1, lg.size(-1)).float(), data["cat_target"][..., j].reshape(-1),
                                              ignore_index=IGNORE_INDEX)
        scaler.scale(loss).backward()
        scaler.unscale_(opt)
        nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        scaler.step(opt)
        scaler.update()
        return loss

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / args.steps], device=dev)
    if world > 1:
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
    # kernel-time share of the quantiser (one profiled step): library kernels vs everything else
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        step()
        torch.cuda.synchronize()
    tot = vq = 0.0
    for e in prof.events():
        if e.device_type != torch.autograd.DeviceType.CUDA:
            continue
        d = e.time_range.end - e.time_range.start
        tot += d
        if "vq::" in e.name:
            vq += d
    share = {"kernel_time_us": tot}
    if kind == "native":
        share.update({"quantiser_kernel_time_us": vq, "quantiser_share": vq / max(tot, 1e-9)})
    vec = batch * 5 * 32 * 32
    out = {"quantizer": kind, "ms_per_step": float(ms), "patches_per_gpu": batch, "vectors_per_gpu_step": vec,
           "vectors_per_s": vec * world / (float(ms) * 1e-3), "n_gpus": world, "loss": float(loss), "profile": share}
    del model, net, opt, data
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=2048, help="patches per GPU (halved on out-of-memory, single GPU only)")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--quantizer", default="both", choices=["native", "eager", "both"])
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp16"])
    ap.add_argument("--codebook_size", type=int, default=8192)
    ap.add_argument("--emb_dim", type=int, default=256)
    ap.add_argument("--hidden", type=int, default=128)
    ap.add_argument("--cont_dim", type=int, default=64)
    args = ap.parse_args()
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=dev)
    res = {"workload": f"C5 train step: conv encoder/decoder (hidden {args.hidden}) in torch + quantiser K={args.codebook_size} "
                       f"D={args.emb_dim}, {args.dtype} autocast, 5x32x32 patches of {args.cont_dim} features", "runs": []}
    for kind in (["native", "eager"] if args.quantizer == "both" else [args.quantizer]):
        batch = args.batch
        while True:
            try:
                res["runs"].append(run(args, kind, batch, dev, world, rank, local))
                break
            except torch.OutOfMemoryError:
                torch.cuda.empty_cache()
                if world > 1 or batch <= 16:
                    raise
                batch //= 2
    if rank == 0:
        print(json.dumps(res), flush=True)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()

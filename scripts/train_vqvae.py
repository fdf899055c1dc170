#!/usr/bin/env python
"""scripts/train_vqvae.py -- the reference's GEN-0 training entry point, re-created around the native quantiser.

Same argument surface as the reference script (scripts/train_vqvae.py:390-458: --codebook_size, --emb_dim, --beta,
--quantizer {st,ema}, --ema_decay, --ema_eps, --lambda_*, --anneal_vq_*, --bf16, --no_amp, --clip_grad, ...) and
the same loop (:261-380): cosine LR, autocast, `model(batch)` -> 5-tuple, MSE + CE + canopy + lambda_vq(step) *
vq_loss, GradScaler, two AdamW groups split on "quant.codebook" (:221-228), per-epoch `{"model","opt","args",
"step"}` checkpoints + ckpt_best.pt (:366-380), a log line with `pplx=` (:351-358).

Differences, stated: the dataset stack (`vqvae.loader/preprocess`, xarray, zarr) is absent from the reference tree
and this image, so `--zarr synthetic` (the default) draws zarr-chunk-shaped synthetic patches
(scripts/vqvae_model.py); the encoder/decoder are a small torch conv stack (BASELINE config 5).  The quantiser is
`vqvae_b200.VectorQuantizer` -- CUDA only, no CPU fallback, so `--cpu` is refused.
With torchrun (WORLD_SIZE > 1) every rank trains on its own patches; the quantiser all-reduces code statistics
 and the conv weights are averaged by DistributedDataParallel.
"""
from __future__ import annotations

import argparse
import math
import os
import sys
import time
from pathlib import Path

import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))

from vqvae_model import IGNORE_INDEX, VQVAE, SyntheticPatches  # noqa: E402
from vqvae_b200 import CodebookManager  # noqa: E402


def mse_ignore_nan(pred: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    mask = torch.isfinite(target)
    if mask.sum() == 0:
        return pred.new_tensor(0.0)
    return ((pred - torch.nan_to_num(target, nan=0.0)) ** 2)[mask].mean()


def load_scheduler(base: dict, a: dict):
    """lambda_vq annealing (reference: vqvae.annealers.load_scheduler, absent; flags scripts/train_vqvae.py:438-456).
    Returns f(step=...) -> dict of loss weights; only "vq" is annealed."""
    if not a.get("anneal_vq_enable"):
        return lambda step: dict(base)
    kind = a.get("anneal_vq_schedule", "warmup_hold_decay")
    lo = float(a.get("anneal_vq_floor", 0.0))
    hi = float(a.get("anneal_vq_ceil", base["vq"]))
    start, dur = int(a.get("anneal_vq_start", 0)), max(1, int(a.get("anneal_vq_duration", 0) or 1))
    k = float(a.get("anneal_vq_k", 5.0))
    warm, hold, dec = int(a.get("anneal_vq_warmup", 0)), int(a.get("anneal_vq_hold", 0)), int(a.get("anneal_vq_decay", 0))
    final = float(a.get("anneal_vq_final", lo))
    miles = sorted((int(s.split(":")[0]), float(s.split(":")[1])) for s in (a.get("anneal_vq_milestones") or []))

    def value(step: int) -> float:
        t = min(max((step - start) / dur, 0.0), 1.0)
        if kind == "constant":
            return hi
        if kind == "linear":
            return lo + (hi - lo) * t
        if kind == "cosine":
            return lo + (hi - lo) * 0.5 * (1 - math.cos(math.pi * t))
        if kind == "exponential":
            return lo + (hi - lo) * (1 - math.exp(-k * t)) / (1 - math.exp(-k))
        if kind == "stepwise":
            v = lo
            for s, w in miles:
                if step >= s:
                    v = w
            return v
        s = step - start                                    # warmup_hold_decay
        if s < warm:
            return lo + (hi - lo) * max(s, 0) / max(1, warm)
        if s < warm + hold:
            return hi
        if s < warm + hold + dec:
            return hi + (final - hi) * (s - warm - hold) / max(1, dec)
        return final

    return lambda step: {**base, "vq": value(step)}


def train(args) -> dict:
    if args.cpu or not torch.cuda.is_available():
        raise RuntimeError("train_vqvae: the quantiser kernels are sm_100a only (no CPU fallback); run on a CUDA device")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1 and not torch.distributed.is_initialized():
        torch.distributed.init_process_group("nccl", device_id=device)
    run_dir = Path(args.run_dir)
    run_dir.mkdir(parents=True, exist_ok=True)
    if args.zarr != "synthetic":
        raise RuntimeError("only --zarr synthetic is available: the reference's dataset stack (vqvae.loader, xarray, "
                           "zarr) is absent from its tree and from this image")
    torch.manual_seed(args.seed + rank)
    ds = SyntheticPatches(cont_dim=args.cont_dim, T=args.time_steps, H=args.patch, W=args.patch, seed=args.seed + rank,
                          device=device)
    model = VQVAE(cont_dim=ds.cont_dim, cat_vocab_sizes=ds.cat_vocab_sizes, naip_bands=0, emb_dim=args.emb_dim,
                  codebook_size=args.codebook_size, beta=args.beta, hidden=args.hidden, quantizer=args.quantizer,
                  cat_emb_dim=args.cat_emb_dim, ema_decay=args.ema_decay, ema_eps=args.ema_eps).to(device)
    manager = CodebookManager(num_codes=model.quant.codebook_size, code_dim=model.quant.emb_dim)
    model.attach_codebook_manager(manager)
    if rank == 0:
        print(f"Model params: {sum(p.numel() for p in model.parameters() if p.requires_grad) / 1e6:.2f}M  "
              f"quantizer={args.quantizer} K={args.codebook_size} D={args.emb_dim}")
    net = model
    if world > 1:
        # every rank seeds its own RNG (seed + rank), DDP broadcasts parameters only (broadcast_buffers=False: the EMA
        # state is the kernels' job) -- so the quantiser state, buffers included, is taken from rank 0 explicitly
        model.quant.sync_state_from(0)
        net = nn.parallel.DistributedDataParallel(model, device_ids=[local], broadcast_buffers=False)
    codebook_params = [p for n, p in model.named_parameters() if ".quant.codebook" in n or n.endswith("quant.codebook")]
    other_params = [p for n, p in model.named_parameters()
                    if ".quant.codebook" not in n and not n.endswith("quant.codebook")]
    groups = [{"params": other_params, "lr": args.lr, "weight_decay": args.weight_decay, "betas": (0.9, 0.95)}]
    if codebook_params:                                   # empty in "ema" mode (the codebook is a buffer)
        groups.append({"params": codebook_params, "lr": args.lr, "weight_decay": 0.0, "betas": (0.9, 0.95)})
    opt = torch.optim.AdamW(groups, lr=args.lr, weight_decay=args.weight_decay, betas=(0.9, 0.95))
    amp = not args.no_amp
    scaler = torch.amp.GradScaler("cuda", enabled=amp and not args.bf16)
    total_steps = args.epochs * args.steps_per_epoch
    lambda_vq = load_scheduler({"vq": args.lambda_vq, "cont": args.lambda_cont, "cat": args.lambda_cat,
                                "canopy": args.lambda_canopy}, {k: v for k, v in vars(args).items() if v is not None})

    def lr_at(step):
        if total_steps <= 1:
            return args.lr
        return args.min_lr + (args.lr - args.min_lr) * 0.5 * (1 + math.cos(math.pi * step / total_steps))

    step, best_loss, avg, hist = 0, float("inf"), {}, []
    model.train()
    for epoch in range(1, args.epochs + 1):
        running = {"recon_cont": 0.0, "recon_cat": 0.0, "canopy": 0.0, "vq": 0.0, "total": 0.0}
        count, t_last, vec_seen = 0, time.time(), 0
        for _ in range(args.steps_per_epoch):
            batch = ds.batch(args.batch_size)
            if step == 0 and args.data_init:              # data-dependent codebook init from the first latents
                with torch.no_grad(), torch.autocast(device_type="cuda", dtype=torch.bfloat16, enabled=amp and args.bf16):
                    model.quant.init_codebook_from(model.encode(batch["cont"]).float())
                model.quant.sync_state_from(0)
            if args.reseed_every > 0 and step > 0 and step % args.reseed_every == 0:
                with torch.no_grad():                     # DP: the dead set is identical on all ranks, rank 0's rows win
                    model.quant.reseed_dead_codes(model.encode(batch["cont"]).float())
                model.quant.sync_state_from(0)
            for g in opt.param_groups:
                g["lr"] = lr_at(step)
            opt.zero_grad(set_to_none=True)
            with torch.autocast(device_type="cuda", dtype=torch.bfloat16 if args.bf16 else torch.float16, enabled=amp):
                cont_pred, cat_logits, canopy_pred, vq_loss, perplexity = net(batch)
                cont_pred = torch.nan_to_num(cont_pred, nan=0.0, posinf=0.0, neginf=0.0)
                canopy_pred = torch.nan_to_num(canopy_pred, nan=0.0, posinf=0.0, neginf=0.0)
                loss_cont = mse_ignore_nan(cont_pred.float(), batch["cont"])
                loss_cat = 0.0
                for j, name in enumerate(ds.cat_names):
                    logits = torch.nan_to_num(cat_logits[name], nan=0.0, posinf=0.0, neginf=0.0)
                    loss_cat = loss_cat + F.cross_entropy(logits.reshape(-1, logits.size(-1)).float(),
                                                          batch["cat_target"][..., j].reshape(-1),
                                                          ignore_index=IGNORE_INDEX, reduction="mean")
                loss_canopy = mse_ignore_nan(canopy_pred.float(), batch["canopy"])
                loss = (args.lambda_cont * loss_cont + args.lambda_cat * loss_cat + args.lambda_canopy * loss_canopy
                        + lambda_vq(step=step)["vq"] * vq_loss)
            scaler.scale(loss).backward()
            if args.clip_grad is not None and args.clip_grad > 0:
                scaler.unscale_(opt)
                nn.utils.clip_grad_norm_(model.parameters(), args.clip_grad)
            scaler.step(opt)
            scaler.update()
            step += 1
            count += 1
            if step == 1 and world > 1:                   # replicas must be bit-identical from the first update on
                from vqvae_b200 import parallel as _par
                _par.assert_replicas_identical(model.quant.codebook)
            vec_seen += batch["cont"][..., 0].numel()
            if step % args.log_every == 0 or step == total_steps:       # host syncs only on logging steps
                vals = {"recon_cont": float(loss_cont.detach()), "recon_cat": float(loss_cat.detach()), "canopy": float(loss_canopy.detach()),
                        "vq": float(vq_loss.detach()), "total": float(loss.detach())}
                avg = vals
                dt = time.time() - t_last
                hist.append(vals["total"])
                if rank == 0:
                    print(f"[epoch {epoch} step {step}] total={vals['total']:.4f} cont={vals['recon_cont']:.4f} "
                          f"cat={vals['recon_cat']:.4f} canopy={vals['canopy']:.4f} vq={vals['vq']:.4f} "
                          f"pplx={float(perplexity):.2f} lr={opt.param_groups[0]['lr']:.2e} "
                          f"dead={int(model.quant.n_dead)} vec/s={world * vec_seen / max(dt, 1e-9):.3e}", flush=True)
                t_last, vec_seen = time.time(), 0
            if step >= total_steps:
                break
        ckpt = {"model": model.state_dict(), "opt": opt.state_dict(), "args": vars(args), "step": step}
        if rank == 0:
            torch.save(ckpt, str(run_dir / f"ckpt_epoch{epoch:03d}.pt"))
            if avg and avg["total"] < best_loss - 1e-4:
                best_loss = avg["total"]
                torch.save(ckpt, str(run_dir / "ckpt_best.pt"))
        if step >= total_steps:
            break
    if rank == 0:
        print("Training done.", manager.summary())
    return {"losses": hist, "step": step, "summary": manager.summary()}


def parse_args(argv=None):
    p = argparse.ArgumentParser(description="Train a VQ-VAE around the B200-native quantiser (synthetic patches).")
    p.add_argument("--config", type=str, help="YAML with a 'train_vqvae' section (CLI flags win, as in utils/argyaml.py)")
    p.add_argument("--zarr", default="synthetic", help="'synthetic' (the reference's zarr loader is absent)")
    p.add_argument("--run_dir", required=True)
    # data / loader
    p.add_argument("--batch_size", type=int, default=64)
    p.add_argument("--num_workers", type=int, default=4)
    p.add_argument("--steps_per_epoch", type=int, default=10000)
    p.add_argument("--epochs", type=int, default=5)
    p.add_argument("--eager", action="store_true")
    p.add_argument("--cont_dim", type=int, default=64, help="synthetic: features per pixel (README.md:22-30)")
    p.add_argument("--time_steps", type=int, default=5)
    p.add_argument("--patch", type=int, default=32)
    p.add_argument("--seed", type=int, default=0)
    # schema flags are accepted for CLI compatibility; the synthetic source has no vocabulary to collapse
    p.add_argument("--min_hits_per_epoch", type=int, default=100)
    p.add_argument("--mass_coverage", type=float, default=0.999)
    p.add_argument("--vocab_cap", type=int, default=5000)
    # model
    p.add_argument("--codebook_size", type=int, default=256)
    p.add_argument("--emb_dim", type=int, default=128)
    p.add_argument("--hidden", type=int, default=128)
    p.add_argument("--cat_emb_dim", type=int, default=8)
    p.add_argument("--beta", type=float, default=0.25)
    # optimisation
    p.add_argument("--lr", type=float, default=3e-4)
    p.add_argument("--min_lr", type=float, default=3e-5)
    p.add_argument("--weight_decay", type=float, default=1e-4)
    p.add_argument("--clip_grad", type=float, default=1.0)
    p.add_argument("--lambda_cont", type=float, default=1.0)
    p.add_argument("--lambda_cat", type=float, default=1.0)
    p.add_argument("--lambda_canopy", type=float, default=1.0)
    p.add_argument("--lambda_vq", type=float, default=1.0)
    p.add_argument("--log_every", type=int, default=1)
    # precision / runtime
    p.add_argument("--no_amp", action="store_true")
    p.add_argument("--bf16", action="store_true")
    p.add_argument("--cpu", action="store_true")
    # quantiser
    p.add_argument("--quantizer", type=str, choices=["st", "ema"], default="st")
    p.add_argument("--ema_decay", type=float, default=0.99)
    p.add_argument("--ema_eps", type=float, default=1e-5)
    p.add_argument("--data_init", action="store_true", help="initialise the codebook from the first batch's latents")
    p.add_argument("--reseed_every", type=int, default=0, help="re-seed dead codes every N steps (0 = never)")
    # annealing
    p.add_argument("--anneal_vq_enable", action="store_true")
    p.add_argument("--anneal_vq_schedule", type=str, default="warmup_hold_decay",
                   choices=["constant", "linear", "cosine", "exponential", "stepwise", "warmup_hold_decay"])
    p.add_argument("--anneal_vq_start", type=int, default=0)
    p.add_argument("--anneal_vq_duration", type=int, default=0)
    p.add_argument("--anneal_vq_floor", type=float, default=0.0)
    p.add_argument("--anneal_vq_ceil", type=float, default=0.1)
    p.add_argument("--anneal_vq_k", type=float, default=5.0)
    p.add_argument("--anneal_vq_warmup", type=int, default=10000)
    p.add_argument("--anneal_vq_hold", type=int, default=15000)
    p.add_argument("--anneal_vq_decay", type=int, default=5000)
    p.add_argument("--anneal_vq_final", type=float, default=.08)
    p.add_argument("--anneal_vq_milestones", type=str, nargs="*", default=None)
    pre, _ = p.parse_known_args(argv)
    if pre.config:                                     # YAML section sets defaults, the command line overrides them
        import yaml
        with open(pre.config) as fh:
            section = (yaml.safe_load(fh) or {}).get("train_vqvae", {}) or {}
        known = {a.dest for a in p._actions}
        p.set_defaults(**{k: v for k, v in section.items() if k in known})
    return p.parse_args(argv)


if __name__ == "__main__":
    train(parse_args())

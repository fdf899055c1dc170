"""VQVAE wrapper + synthetic patch source for the re-created reference entry points.

The reference's `vqvae.model.VQVAE`, `vqvae.loader` and `vqvae.preprocess` are absent from its tree
(scripts/train_vqvae.py:78-98) and xarray/zarr are not installed here, so this file supplies the minimum the
training script needs around the native quantiser: a torch conv encoder/decoder (BASELINE config 5 keeps the
conv stack in torch) and a synthetic source of zarr-chunk-shaped patches (time, y, x, feature)
(README.md:22-30).  Model contract kept from scripts/train_vqvae.py:183-198,287:
    VQVAE(cont_dim, cat_vocab_sizes, naip_bands, emb_dim, codebook_size, beta, hidden, quantizer,
          cat_emb_dim, ema_decay, ema_eps);  .quant ; .attach_codebook_manager(m) ;
    forward(batch) -> (cont_pred, cat_logits, canopy_pred, vq_loss, perplexity)
"""
from __future__ import annotations

from typing import Dict

import torch
import torch.nn as nn

from vqvae_b200 import VectorQuantizer

IGNORE_INDEX = -100


class SyntheticPatches:
    """Endless batches shaped like one zarr chunk per sample: cont [B, T, H, W, C] (channel-last), a categorical
    target [B, T, H, W] per categorical feature and a canopy scalar [B].  A low-rank latent structure makes
    the reconstruction losses fall during a short run."""

    def __init__(self, cont_dim=64, cat_vocab_sizes=None, T=5, H=32, W=32, seed=0, device="cpu"):
        self.cont_dim, self.T, self.H, self.W = cont_dim, T, H, W
        self.cat_vocab_sizes = dict(cat_vocab_sizes or {"lc": 12})
        self.cont_names = [f"c{i}" for i in range(cont_dim)]
        self.cat_names = list(self.cat_vocab_sizes)
        self.device = torch.device(device)
        self.gen = torch.Generator(device=self.device).manual_seed(seed)
        g = torch.Generator().manual_seed(seed + 1)
        self.proto = torch.randn(32, cont_dim, generator=g).to(self.device)     # 32 latent "land-cover types"

    def batch(self, B: int) -> Dict[str, torch.Tensor]:
        dev, g = self.device, self.gen
        kind = torch.randint(0, self.proto.shape[0], (B, 1, self.H, self.W), generator=g, device=dev)
        kind = kind.expand(B, self.T, self.H, self.W)
        cont = self.proto[kind] + 0.25 * torch.randn(B, self.T, self.H, self.W, self.cont_dim, generator=g, device=dev)
        out = {"cont": cont, "canopy": (kind[:, -1].float().mean(dim=(1, 2)) / self.proto.shape[0])}
        out["cat_target"] = torch.stack([kind % n for n in self.cat_vocab_sizes.values()], dim=-1)
        return out


class VQVAE(nn.Module):
    """conv encoder -> native VectorQuantizer -> conv decoder heads (channels_last, so the [N, D] view the
    quantiser takes is zero-copy: SURVEY.md 8a-0)."""

    def __init__(self, cont_dim: int, cat_vocab_sizes: Dict[str, int], naip_bands: int = 0, emb_dim: int = 128,
                 codebook_size: int = 256, beta: float = 0.25, hidden: int = 128, quantizer: str = "st",
                 cat_emb_dim: int = 8, ema_decay: float = 0.99, ema_eps: float = 1e-5, **quant_kw):
        super().__init__()
        self.cont_dim, self.cat_vocab_sizes = cont_dim, dict(cat_vocab_sizes)
        self.encoder = nn.Sequential(
            nn.Conv2d(cont_dim, hidden, 3, padding=1), nn.GELU(),
            nn.Conv2d(hidden, hidden, 3, padding=1), nn.GELU(),
            nn.Conv2d(hidden, emb_dim, 1))
        self.quant = VectorQuantizer(emb_dim, codebook_size, beta=beta, quantizer=quantizer, ema_decay=ema_decay,
                                     ema_eps=ema_eps, **quant_kw)
        self.dec_trunk = nn.Sequential(nn.Conv2d(emb_dim, hidden, 3, padding=1), nn.GELU())
        self.head_cont = nn.Conv2d(hidden, cont_dim, 1)
        self.head_cat = nn.ModuleDict({k: nn.Conv2d(hidden, n, 1) for k, n in self.cat_vocab_sizes.items()})
        self.head_canopy = nn.Linear(hidden, 1)
        self.to(memory_format=torch.channels_last)

    def attach_codebook_manager(self, manager) -> None:
        self.quant.attach_codebook_manager(manager)

    def encode(self, cont: torch.Tensor) -> torch.Tensor:
        """[B, T, H, W, C] channel-last -> latents [B, T, H, W, D] (a view of the conv output, no transpose pass)."""
        B, T, H, W, C = cont.shape
        x = cont.reshape(B * T, H, W, C).permute(0, 3, 1, 2)          # NCHW view with channels_last strides
        z = self.encoder(x)                                            # stays channels_last
        return z.permute(0, 2, 3, 1).reshape(B, T, H, W, -1)

    def decode(self, zq: torch.Tensor):
        B, T, H, W, D = zq.shape
        h = self.dec_trunk(zq.reshape(B * T, H, W, D).permute(0, 3, 1, 2))
        cont_pred = self.head_cont(h).permute(0, 2, 3, 1).reshape(B, T, H, W, -1)
        cat_logits = {k: m(h).permute(0, 2, 3, 1).reshape(B, T, H, W, -1) for k, m in self.head_cat.items()}
        pooled = h.reshape(B, T, -1, H, W)[:, -1].mean(dim=(2, 3))     # last time step, spatial mean
        return cont_pred, cat_logits, self.head_canopy(pooled)[:, 0]

    def forward(self, batch: Dict[str, torch.Tensor]):
        z = self.encode(batch["cont"])
        zq, vq_loss, _ = self.quant(z)
        cont_pred, cat_logits, canopy_pred = self.decode(zq)
        return cont_pred, cat_logits, canopy_pred, vq_loss, self.quant.perplexity

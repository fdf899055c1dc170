"""CPU oracle for the vector-quantisation hot path  --  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED.  The reference tree (nnnagle/vq-vae) does not contain a quantiser:
``scripts/train_vqvae.py:78-98`` and ``scripts/export_codebook.py:41-49`` import a
``vqvae`` package that is absent, and ``frl/models/__init__.py:15-82`` exports none.
There are no golden vectors, tests or checkpoints for it either.  This file is therefore
a *self-written* PyTorch restatement of the textbook VQ-VAE (van den Oord et al. 2017:
straight-through + commitment) and EMA-VQ arithmetic, constrained only by the call-site
contract the reference scripts impose:

  * constructor kwargs  emb_dim / codebook_size / beta / quantizer{"st","ema"} /
    ema_decay / ema_eps            -> scripts/train_vqvae.py:183-195, :410-414, :433-436
  * ``.codebook`` is ``[K, D]``   -> scripts/export_codebook.py:104,114,138
  * ``.codebook_size``/``.emb_dim`` -> scripts/train_vqvae.py:197
  * parameter name contains ``quant.codebook`` in "st" mode -> scripts/train_vqvae.py:221-228
  * scalar ``vq_loss`` and a ``perplexity`` value -> scripts/train_vqvae.py:287,321,357
  * per-code usage recoverable from the state dict -> scripts/export_codebook.py:124-128
  * commitment_cost 0.25, terms [commitment, codebook] -> frl/config/frl_bindings_v0.yaml:887-891
  * straight_through: true -> frl/config/frl_model_v0.yaml:35

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl
reference`` legs may import this module.  The product (``vqvae_b200``) never does.

Choices that no reference line pins are marked (CHOICE) below and listed in DESIGN.md.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.nn as nn

PERPLEXITY_EPS = 1e-10          # (CHOICE) exp(-sum p log(p + 1e-10))
DEFAULT_DEAD_AFTER = 100        # (CHOICE) steps without an assignment before a code is "dead"


# --------------------------------------------------------------------------- a-0
def flatten_latents(z: torch.Tensor, channel_first: bool = False) -> torch.Tensor:
    """[..., D] (zarr chunk order (time, y, x, feature), README.md:22-30) -> [N, D].

    channel_first=True takes the live model's [B, D, *spatial] layout
    (frl/models/representation.py:317-334) the way
    frl/examples/variance_covariance_loss_example.py:78-86 flattens it.
    """
    if channel_first:
        perm = (0,) + tuple(range(2, z.dim())) + (1,)
        z = z.permute(*perm)
    return z.reshape(-1, z.shape[-1])


# --------------------------------------------------------------------------- a-1
def round_bf16(x: torch.Tensor) -> torch.Tensor:
    return x.to(torch.bfloat16).to(torch.float32)


def round_f16(x: torch.Tensor) -> torch.Tensor:
    return x.to(torch.float16).to(torch.float32)


def round_to(x: torch.Tensor, precision: str) -> torch.Tensor:
    """What the 16-bit tensor-core searches see: the operand rounded to the search's element type."""
    return round_bf16(x) if precision == "bf16" else round_f16(x) if precision == "fp16" else x


def code_norms(E: torch.Tensor, precision: str) -> torch.Tensor:
    """||e_k||^2 in fp32.  bf16 / fp16 mode: of the rounded codebook, so that the search is
    the exact nearest neighbour among the rounded vectors (CHOICE)."""
    Ef = round_to(E.detach().to(torch.float32), precision)
    return (Ef * Ef).sum(dim=1)


def scores(z: torch.Tensor, E: torch.Tensor, precision: str = "fp32") -> torch.Tensor:
    """s[n,k] = ||e_k||^2 - 2 z_n.e_k   ( = d[n,k] - ||z_n||^2, same argmin ).

    precision "fp32": everything fp32.  "bf16" / "fp16": z and E rounded to that type, products
    accumulated in fp32 (what a 16-bit tensor-core GEMM with fp32 accumulate computes; fp16 is
    the reference script's default autocast dtype, scripts/train_vqvae.py:284-286).
    "fp64": for near-tie adjudication in tests.
    """
    if precision == "fp64":
        zf, Ef = z.detach().double(), E.detach().double()
        return (Ef * Ef).sum(1)[None, :] - 2.0 * (zf @ Ef.t())
    zf = z.detach().to(torch.float32)
    Ef = E.detach().to(torch.float32)
    if precision in ("bf16", "fp16"):
        zf, Ef = round_to(zf, precision), round_to(Ef, precision)
    elif precision != "fp32":
        raise ValueError(precision)
    return code_norms(E, precision)[None, :] - 2.0 * (zf @ Ef.t())


def row_is_finite(z: torch.Tensor) -> torch.Tensor:
    return torch.isfinite(z.detach().to(torch.float32)).all(dim=1)


def assign(z: torch.Tensor, E: torch.Tensor, precision: str = "fp32",
           chunk: int = 16384) -> Tuple[torch.Tensor, torch.Tensor]:
    """Nearest-code search.  Returns (idx int64 [N], dmin fp32 [N]).

    idx[n] = argmin_k d[n,k]; lowest index wins ties (CHOICE; torch.argmin semantics).
    A row with any non-finite element gets idx = 0 (CHOICE).
    dmin[n] = ||z_n||^2 + s[n, idx[n]] (clamped at 0 is NOT applied).
    """
    N = z.shape[0]
    idx = torch.empty(N, dtype=torch.int64)
    dmin = torch.empty(N, dtype=torch.float32)
    for a in range(0, N, chunk):
        zc = z[a:a + chunk]
        s = scores(zc, E, precision)
        smin, i = torch.min(s, dim=1)          # first minimal index
        # torch.min(dim) documents no tie order; argmin does. Use argmin for the index.
        i = torch.argmin(s, dim=1)
        zf = round_to(zc.detach().to(torch.float32), precision)
        z2 = (zf * zf).sum(1)
        ok = row_is_finite(zc)
        idx[a:a + chunk] = torch.where(ok, i, torch.zeros_like(i))
        dmin[a:a + chunk] = (z2 + s.gather(1, i[:, None])[:, 0]).to(torch.float32)
    return idx, dmin


# --------------------------------------------------------------------------- a-2 .. a-5
def gather(E: torch.Tensor, idx: torch.Tensor, out_dtype: torch.dtype) -> torch.Tensor:
    """q_n = E_fp32[idx_n] cast to the output dtype (gather from the fp32 master) (CHOICE)."""
    return E.detach().to(torch.float32)[idx].to(out_dtype)


def sq_err_sum(z: torch.Tensor, E: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """sum_{n finite, d} (z - E[idx])^2 in fp64->fp32.  Non-finite rows contribute 0 (CHOICE)."""
    ok = row_is_finite(z)
    diff = z.detach().double() - E.detach().double()[idx]
    diff = torch.where(ok[:, None], diff, torch.zeros_like(diff))
    return (diff * diff).sum()


def code_stats(z: torch.Tensor, idx: torch.Tensor, K: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """cnt_k = #{n finite: idx_n = k};  sum_k = sum_{n finite: idx_n = k} z_n  (fp32 of the
    same z the search saw; accumulated in fp64 here so the oracle is order-independent)."""
    ok = row_is_finite(z)
    zf = z.detach().double()
    zf = torch.where(ok[:, None], zf, torch.zeros_like(zf))
    cnt = torch.zeros(K, dtype=torch.float64)
    cnt.index_add_(0, idx, ok.double())
    s = torch.zeros(K, z.shape[1], dtype=torch.float64)
    s.index_add_(0, idx, zf)
    return cnt.to(torch.float32), s.to(torch.float32)


def vq_loss_value(sqerr: torch.Tensor, N: int, D: int, beta: float, quantizer: str) -> torch.Tensor:
    """mse = sum (z-q)^2 / (N*D)  (mean over N*D) (CHOICE).
    st : loss = mse(sg[z], q) + beta * mse(z, sg[q])   (value (1+beta)*mse)
    ema: loss = beta * mse(z, sg[q])"""
    mse = sqerr.double() / float(max(N * D, 1))
    w = (1.0 + beta) if quantizer == "st" else beta
    return (w * mse).to(torch.float32)


# --------------------------------------------------------------------------- a-7 .. a-9
def ema_update(E: torch.Tensor, ema_cnt: torch.Tensor, ema_sum: torch.Tensor,
               cnt: torch.Tensor, sums: torch.Tensor, decay: float, eps: float
               ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """n_k <- g n_k + (1-g) cnt_k ; m_k <- g m_k + (1-g) sum_k ;
    n^_k = (n_k + eps) / (sum n + K eps) * sum n ; e_k <- m_k / n^_k   (Laplace form).
    Returns new (E, ema_cnt, ema_sum), fp32; arithmetic in fp32 except sum n (fp64->fp32)."""
    K = E.shape[0]
    g = torch.tensor(decay, dtype=torch.float32)
    omg = torch.tensor(1.0, dtype=torch.float32) - g          # 1 - g evaluated in fp32, like the kernel
    n = g * ema_cnt.float() + omg * cnt.float()
    m = g * ema_sum.float() + omg[None] * sums.float()
    tot = n.double().sum().to(torch.float32)
    nhat = (n + eps) / (tot + K * eps) * tot
    Enew = m / nhat[:, None]
    return Enew, n, m


def perplexity(cnt: torch.Tensor) -> torch.Tensor:
    """exp(-sum p log(p + 1e-10)), p = cnt / sum cnt (CHOICE of epsilon).  0 assignments -> 1.0"""
    tot = cnt.double().sum()
    if float(tot) <= 0:
        return torch.tensor(1.0, dtype=torch.float32)
    p = cnt.double() / tot
    return torch.exp(-(p * torch.log(p + PERPLEXITY_EPS)).sum()).to(torch.float32)


def usage_update(usage_total: torch.Tensor, steps_since_used: torch.Tensor, cnt: torch.Tensor,
                 dead_after: int) -> Tuple[torch.Tensor, torch.Tensor, int]:
    """usage_total += cnt ; steps_since_used = 0 where cnt > 0 else +1 ;
    n_dead = #{k : steps_since_used >= dead_after}  (CHOICE; rule unpinned by the reference)."""
    used = cnt > 0
    u = usage_total + cnt.round().to(torch.int64)
    s = torch.where(used, torch.zeros_like(steps_since_used), steps_since_used + 1)
    return u, s, int((s >= dead_after).sum())


# --------------------------------------------------------------------------- backward
def backward_z(z: torch.Tensor, q: torch.Tensor, grad_q: Optional[torch.Tensor],
               grad_loss: float, beta: float) -> torch.Tensor:
    """dL/dz = grad_q (straight-through identity) + grad_loss * 2 beta (z - q) / (N D)."""
    N, D = z.shape
    g = 2.0 * beta * grad_loss / float(N * D) * (z.double() - q.double())
    if grad_q is not None:
        g = g + grad_q.double()
    return g.to(z.dtype)


def backward_codebook(E, cnt, sums, grad_loss: float, N: int, D: int) -> torch.Tensor:
    """st mode: dL/de_k = grad_loss * 2/(N D) * sum_{n: idx_n=k} (e_k - z_n)."""
    return (2.0 * grad_loss / float(N * D) * (cnt.double()[:, None] * E.double() - sums.double())).float()


# --------------------------------------------------------------------------- the module
@dataclass
class StepStats:
    cnt: torch.Tensor
    sums: torch.Tensor
    sqerr: torch.Tensor
    n_rows: int


class VectorQuantizerOracle(nn.Module):
    """Same object contract as ``vqvae_b200`` 's VectorQuantizer, pure torch, CPU.

    forward(z [..., D]) -> (quantized [..., D] same dtype, loss fp32 scalar, indices int64 [...]).
    Order inside one training forward (CHOICE): search with the *current* codebook, emit
    quantized/loss from it, then apply the EMA update (so the update is visible next step).
    """

    def __init__(self, emb_dim: int, codebook_size: int, beta: float = 0.25, quantizer: str = "st",
                 ema_decay: float = 0.99, ema_eps: float = 1e-5, precision: str = "fp32",
                 dead_after: int = DEFAULT_DEAD_AFTER):
        super().__init__()
        if quantizer not in ("st", "ema"):
            raise ValueError(f"quantizer must be 'st' or 'ema', got {quantizer!r}")
        self.emb_dim, self.codebook_size = int(emb_dim), int(codebook_size)
        self.beta, self.quantizer = float(beta), quantizer
        self.ema_decay, self.ema_eps = float(ema_decay), float(ema_eps)
        self.precision, self.dead_after = precision, int(dead_after)
        K, D = self.codebook_size, self.emb_dim
        init = torch.empty(K, D, dtype=torch.float32)
        if quantizer == "st":
            init.uniform_(-1.0 / K, 1.0 / K)
            self.codebook = nn.Parameter(init)
        else:
            init.normal_()
            self.register_buffer("codebook", init)
        self.register_buffer("ema_cnt", torch.ones(K, dtype=torch.float32))       # (CHOICE) n=1, m=E
        self.register_buffer("ema_sum", init.detach().clone())
        self.register_buffer("usage_total", torch.zeros(K, dtype=torch.int64))
        self.register_buffer("steps_since_used", torch.zeros(K, dtype=torch.int32))
        self.register_buffer("last_perplexity", torch.ones(1, dtype=torch.float32))
        self.register_buffer("n_dead", torch.zeros(1, dtype=torch.int32))
        self.last_stats: Optional[StepStats] = None

    def attach_codebook_manager(self, manager) -> None:       # scripts/train_vqvae.py:197-198
        self.manager = manager

    def forward(self, z: torch.Tensor):
        with torch.autocast(device_type="cpu", enabled=False):
            lead, D = z.shape[:-1], z.shape[-1]
            zf = z.reshape(-1, D)
            N = zf.shape[0]
            E = self.codebook
            idx, _ = assign(zf, E, self.precision)
            q = gather(E, idx, z.dtype)
            ok = row_is_finite(zf)
            # differentiable loss (fp32), identical in value to vq_loss_value()
            okf = ok[:, None].to(torch.float32)
            zf32 = torch.where(ok[:, None], zf.to(torch.float32), torch.zeros((), dtype=torch.float32))
            e_sel = E[idx].to(torch.float32) * okf
            denom = float(max(N * D, 1))
            commit = ((zf32 - e_sel.detach()) ** 2).sum() / denom
            if self.quantizer == "st":
                loss = ((zf32.detach() - e_sel) ** 2).sum() / denom + self.beta * commit
            else:
                loss = self.beta * commit
            # straight-through: value exactly q, gradient identity to z
            q_st = q.detach() + (zf - zf.detach())
            q_st = torch.where(ok[:, None], q_st, q.detach())
            cnt, sums = code_stats(zf, idx, self.codebook_size)
            self.last_stats = StepStats(cnt, sums, sq_err_sum(zf, E, idx), N)
            if self.training:
                with torch.no_grad():
                    if self.quantizer == "ema":
                        En, n, m = ema_update(E, self.ema_cnt, self.ema_sum, cnt, sums,
                                              self.ema_decay, self.ema_eps)
                        self.codebook.copy_(En); self.ema_cnt.copy_(n); self.ema_sum.copy_(m)
                    u, s, nd = usage_update(self.usage_total, self.steps_since_used, cnt, self.dead_after)
                    self.usage_total.copy_(u); self.steps_since_used.copy_(s); self.n_dead.fill_(nd)
            self.last_perplexity.copy_(perplexity(cnt))
            return q_st.reshape(*lead, D), loss.to(torch.float32), idx.reshape(lead)


# --------------------------------------------------------------------------- test helpers
def near_tie_report(z: torch.Tensor, E: torch.Tensor, idx_test: torch.Tensor, idx_ref: torch.Tensor,
                    rel: float = 1e-5, precision: str = "fp32") -> dict:
    """north_star: indices bit-exact except at near-ties (|delta d| < rel relative), counted.
    For every row where idx_test != idx_ref, d of both candidates is evaluated in fp64 on the
    inputs the kernel saw (bf16-rounded in bf16 mode) and the relative gap is classified."""
    bad = (idx_test != idx_ref).nonzero()[:, 0]
    out = {"n": int(z.shape[0]), "mismatch": int(bad.numel()), "near_tie": 0, "hard": 0, "max_rel_gap": 0.0}
    if bad.numel() == 0:
        return out
    zf = round_to(z[bad].detach().float(), precision)
    Ef = round_to(E.detach().float(), precision)
    zd, Ed = zf.double(), Ef.double()
    dt = ((zd - Ed[idx_test[bad]]) ** 2).sum(1)
    dr = ((zd - Ed[idx_ref[bad]]) ** 2).sum(1)
    scale = torch.maximum(torch.maximum(dt, dr), (zd * zd).sum(1)).clamp_min(1e-30)
    gap = (dt - dr).abs() / scale
    out["near_tie"] = int((gap < rel).sum())
    out["hard"] = int((gap >= rel).sum())
    out["max_rel_gap"] = float(gap.max())
    return out

"""VectorQuantizer: the ``model.quant`` object the reference scripts expect, on B200 kernels.

Contract (every line below is a reference call site, the arithmetic itself is absent
from the reference tree and follows oracle/vq_reference.py -- "parity unpinned"):

  * ``VectorQuantizer(emb_dim, codebook_size, beta, quantizer in {"st","ema"}, ema_decay, ema_eps)``
        scripts/train_vqvae.py:183-195, CLI defaults :410-414, :433-436
  * ``.codebook`` [K, D] fp32; an ``nn.Parameter`` in "st" mode (so its qualified name contains
    ``quant.codebook`` and lands in the weight_decay=0 AdamW group), a buffer in "ema" mode
        scripts/train_vqvae.py:221-228 ; scripts/export_codebook.py:104,114,138
  * ``.codebook_size`` / ``.emb_dim``                      scripts/train_vqvae.py:197
  * ``forward(z) -> (quantized, loss, indices)``; ``.last_perplexity`` for the outer model's 5-tuple
        scripts/train_vqvae.py:287,321,357
  * ``attach_codebook_manager(CodebookManager(num_codes, code_dim))``   scripts/train_vqvae.py:197-198
  * all state (codebook, EMA accumulators, usage counters) in ``state_dict()``
        scripts/train_vqvae.py:367-373 ; scripts/export_codebook.py:124-128

forward() on a CPU tensor raises: the kernels are sm_100a only and there is no fallback.
"""
from __future__ import annotations

import contextlib

import os
from typing import Optional

import torch
import torch.nn as nn

from . import ops, parallel
from ._lib import VQ_STAT_SUM_RESID, VQ_STAT_SUM_Z

DEFAULT_DEAD_AFTER = 100
PEER_ONE_SHOT_MAX_BYTES = 16 << 20      # total bytes a rank reads from its peers per step on the fused path
GRAPH_MAX_BYTES = 32 << 20              # batches up to this size replay the step as one CUDA graph (launch-latency bound)
GRAPH_CACHE = 4                         # captured (shape, mode) combinations kept per module


class CodebookManager:
    """Usage / dead-code view over a quantiser (scripts/train_vqvae.py:92,197-198).

    The reference's own class is absent; this keeps the constructor the script calls and reads the
    counters the kernels maintain in the module's registered buffers.
    """

    def __init__(self, num_codes: int, code_dim: int):
        self.num_codes, self.code_dim = int(num_codes), int(code_dim)
        self.quant: Optional["VectorQuantizer"] = None

    def attach(self, quant: "VectorQuantizer") -> None:
        if (quant.codebook_size, quant.emb_dim) != (self.num_codes, self.code_dim):
            raise ValueError("CodebookManager shape does not match the quantiser")
        self.quant = quant

    def usage(self) -> torch.Tensor:
        return self.quant.usage_total.detach().cpu()

    def dead_codes(self) -> torch.Tensor:
        q = self.quant
        return (q.steps_since_used >= q.dead_after).nonzero()[:, 0].cpu()

    def summary(self) -> dict:
        q = self.quant
        return {"n_dead": int(q.n_dead), "perplexity": float(q.last_perplexity),
                "used": int((q.usage_total > 0).sum())}


def extract_code_usage_from_state(state: dict, prefix: str = "") -> torch.Tensor:
    """Per-code cumulative usage from a plain state dict (scripts/export_codebook.py:124-125)."""
    for k, v in state.items():
        if k.startswith(prefix) and k.endswith("usage_total"):
            return v.detach().cpu().to(torch.int64)
    raise KeyError("no '*usage_total' entry in the state dict")


_NULL_CTX = contextlib.nullcontext()


class _VQFunction(torch.autograd.Function):
    """Forward = search + fused gather pass (+ statistics); backward = straight-through + commitment."""

    @staticmethod
    def forward(ctx, z2: torch.Tensor, codebook: torch.Tensor, mod: "VectorQuantizer", search: str):
        N, D = z2.shape
        K = mod.codebook_size
        mod._ensure_shadow(search)
        q, loss, idx, code_sum = mod._run_step(z2, search)
        loss = loss.reshape(())
        ctx.save_for_backward(z2, q)
        ctx.mod, ctx.N, ctx.D = mod, N, D
        ctx.resid = code_sum.clone() if (mod.quantizer == "st" and codebook.requires_grad) else None
        ctx.mark_non_differentiable(idx)
        return q, loss, idx

    @staticmethod
    def backward(ctx, grad_q, grad_loss, _grad_idx):
        z2, q = ctx.saved_tensors
        mod, N, D = ctx.mod, ctx.N, ctx.D
        dev = z2.device
        gl = (torch.zeros(1, dtype=torch.float32, device=dev) if grad_loss is None
              else grad_loss.detach().to(torch.float32).reshape(1))
        grad_z = grad_E = None
        if ctx.needs_input_grad[0]:
            gq = None if grad_q is None else grad_q.to(z2.dtype).contiguous()
            grad_z = torch.empty((N, D), dtype=z2.dtype, device=dev)
            ops.backward(z2, q, gq, gl, 2.0 * mod.beta / float(max(N * D, 1)), grad_z)
        if ctx.needs_input_grad[1] and ctx.resid is not None:
            grad_E = torch.empty((mod.codebook_size, D), dtype=torch.float32, device=dev)
            ops.codebook_grad(ctx.resid, gl, 2.0 / float(max(N * D, 1)), grad_E)
        return grad_z, grad_E, None, None


class _LayoutTranspose(torch.autograd.Function):
    """[B, R, C] -> [B, C, R] on the library's transpose kernel (a-0); the backward is the same kernel the other way."""

    @staticmethod
    def forward(ctx, x: torch.Tensor):
        B, R, C = x.shape
        out = torch.empty((B, C, R), dtype=x.dtype, device=x.device)
        ops.layout_transpose(x.contiguous(), out)
        return out

    @staticmethod
    def backward(ctx, g: torch.Tensor):
        B, C, R = g.shape
        out = torch.empty((B, R, C), dtype=g.dtype, device=g.device)
        ops.layout_transpose(g.contiguous(), out)
        return out


class _VQFunctionCF(torch.autograd.Function):
    """Channel-first call: z [B, D, S] -> (q [B, D, S], loss, idx [B*S]).  Backward = straight-through + commitment on the
    channel-first tensors themselves (elementwise, so no layout pass), codebook gradient from the residual sums."""

    @staticmethod
    def forward(ctx, z3: torch.Tensor, codebook: torch.Tensor, mod: "VectorQuantizer", search: str):
        B, D, S = z3.shape
        rows = torch.empty((B, S, D), dtype=z3.dtype, device=z3.device)
        ops.layout_transpose(z3, rows)                               # the ONE layout pass of the call
        qcf = torch.empty_like(z3)
        mod._ensure_shadow(search)
        loss, idx, code_sum = mod._run_step_cf(rows.view(B * S, D), qcf, search)
        ctx.save_for_backward(z3, qcf)
        ctx.mod, ctx.n_rows = mod, B * S
        ctx.resid = code_sum.clone() if (mod.quantizer == "st" and codebook.requires_grad) else None
        ctx.mark_non_differentiable(idx)
        return qcf, loss.reshape(()), idx

    @staticmethod
    def backward(ctx, grad_q, grad_loss, _grad_idx):
        z3, qcf = ctx.saved_tensors
        mod, N = ctx.mod, ctx.n_rows
        B, D, S = z3.shape
        dev = z3.device
        gl = (torch.zeros(1, dtype=torch.float32, device=dev) if grad_loss is None
              else grad_loss.detach().to(torch.float32).reshape(1))
        grad_z = grad_E = None
        if ctx.needs_input_grad[0]:
            gq = None if grad_q is None else grad_q.to(z3.dtype).contiguous().view(B * D, S)
            grad_z = torch.empty_like(z3)
            ops.backward(z3.view(B * D, S), qcf.view(B * D, S), gq, gl, 2.0 * mod.beta / float(max(N * D, 1)),
                         grad_z.view(B * D, S))
        if ctx.needs_input_grad[1] and ctx.resid is not None:
            grad_E = torch.empty((mod.codebook_size, D), dtype=torch.float32, device=dev)
            ops.codebook_grad(ctx.resid, gl, 2.0 / float(max(N * D, 1)), grad_E)
        return grad_z, grad_E, None, None


class VectorQuantizer(nn.Module):
    """Nearest-codebook quantiser with straight-through output, commitment/codebook loss and (in "ema"
    mode) an EMA codebook update, all on hand-written sm_100a kernels.

    search: "auto" (bf16 / fp16 latents -> tensor-core search on operands of that type, fp32 latents -> the fp32 tier),
        "bf16" (fp32 / fp16 latents are staged as bf16), "fp16" (fp16 latents only), "fp32" (the 1e-5 tier: split-operand
        tensor-core search, three bf16 MMAs per product, CUDA cores where the shape does not fit) or "fp32_simt" (always
        the CUDA-core kernel).
    deterministic: code-sorted visiting order reproducible run to run (stable counting sort).
    channel_first: forward takes [B, D, *spatial] (the live model's layout, frl/models/representation.py:317-334)
        and returns quantized in the same layout, indices [B, *spatial]; one transpose pass each way (a-0).
    cuda_graph: small batches (<= 32 MB of latents, single process) replay the whole step -- search, sort, pass, EMA update,
        index widening -- as ONE captured CUDA graph from the second call with the same shape on: at the reference's own
        batch sizes (8 patches = 40 960 vectors) the step is launch-latency bound, not kernel bound.
    peer_exchange: "auto" (default) uses the fused peer-memory exchange (parallel.PeerExchange + vq_ema_update_dp)
        for the EMA statistics when the process group runs NCCL on one node, "nccl" always all-reduces with
        torch.distributed, "peer" insists on the fused path.
    peer_timeout_s: wall-clock bound of the cross-rank wait on the fused path (default: VQ_PEER_TIMEOUT_S or 1800 s --
        a rank may be minutes late because of rank-0-only validation or checkpointing; 0 waits for ever).
    process_group / sync_stats: with torch.distributed initialised, per-code counts and sums are
        all-reduced (SUM) before the EMA step so every rank applies the identical update.
    """

    def __init__(self, emb_dim: int, codebook_size: int, beta: float = 0.25, quantizer: str = "st",
                 ema_decay: float = 0.99, ema_eps: float = 1e-5, search: str = "auto",
                 dead_after: int = DEFAULT_DEAD_AFTER, deterministic: bool = False, process_group=None,
                 sync_stats: bool = True, channel_first: bool = False, peer_exchange: str = "auto",
                 cuda_graph: bool = True, peer_timeout_s: Optional[float] = None):
        super().__init__()
        self.peer_timeout_s = peer_timeout_s
        if quantizer not in ("st", "ema"):
            raise ValueError(f"quantizer must be 'st' or 'ema', got {quantizer!r}")
        if search not in ("auto", "bf16", "fp16", "fp32", "fp32_simt"):
            raise ValueError(f"search must be 'auto', 'bf16', 'fp16', 'fp32' or 'fp32_simt', got {search!r}")
        self.emb_dim, self.codebook_size = int(emb_dim), int(codebook_size)
        self.beta, self.quantizer = float(beta), quantizer
        self.ema_decay, self.ema_eps = float(ema_decay), float(ema_eps)
        self.search, self.dead_after, self.deterministic = search, int(dead_after), bool(deterministic)
        self.process_group, self.sync_stats = process_group, bool(sync_stats)
        self.channel_first = bool(channel_first)
        if peer_exchange not in ("auto", "nccl", "peer"):
            raise ValueError("peer_exchange must be 'auto', 'nccl' or 'peer'")
        self.peer_exchange = peer_exchange
        self._peers = None
        K, D = self.codebook_size, self.emb_dim
        init = torch.empty(K, D, dtype=torch.float32)
        if quantizer == "st":
            init.uniform_(-1.0 / K, 1.0 / K)
            self.codebook = nn.Parameter(init)
        else:
            init.normal_()
            self.register_buffer("codebook", init)
        self.register_buffer("ema_cnt", torch.ones(K, dtype=torch.float32))
        self.register_buffer("ema_sum", init.detach().clone())
        self.register_buffer("usage_total", torch.zeros(K, dtype=torch.int64))
        self.register_buffer("steps_since_used", torch.zeros(K, dtype=torch.int32))
        self.register_buffer("last_perplexity", torch.ones(1, dtype=torch.float32))
        self.register_buffer("n_dead", torch.zeros(1, dtype=torch.int32))
        self.manager: Optional[CodebookManager] = None
        # scratch (never in the state dict)
        self._E_bf16 = self._e_norm2 = self._ws = self._stats = None
        self._ws_need = {}
        self._shadow_key = None
        self.cuda_graph = bool(cuda_graph)
        self._graphs: dict = {}
        self.graph_launches = 0          # library kernels launched through graph replays (ops.launch_count sees only eager ones)

    # ------------------------------------------------------------------ reference-script surface
    def attach_codebook_manager(self, manager) -> None:
        self.manager = manager
        if hasattr(manager, "attach"):
            manager.attach(self)

    @property
    def perplexity(self) -> torch.Tensor:
        return self.last_perplexity[0]

    def extra_repr(self) -> str:
        return (f"emb_dim={self.emb_dim}, codebook_size={self.codebook_size}, beta={self.beta}, "
                f"quantizer={self.quantizer!r}, search={self.search!r}")

    # ------------------------------------------------------------------ codebook maintenance (host-driven, rare)
    @torch.no_grad()
    def init_codebook_from(self, z: torch.Tensor, generator: Optional[torch.Generator] = None) -> None:
        """Data-dependent initialisation: K rows of z (sampled without replacement when N >= K) become the
        codebook; EMA accumulators restart at n = 1, m = E.  Call on every rank with the SAME rows (or
        broadcast the codebook afterwards).  Not part of the per-step path; plain torch indexing."""
        rows = z.detach().reshape(-1, self.emb_dim).to(torch.float32)
        N, K = rows.shape[0], self.codebook_size
        pick = (torch.randperm(N, generator=generator, device=rows.device)[:K] if N >= K
                else torch.randint(0, N, (K,), generator=generator, device=rows.device))
        self.codebook.data.copy_(rows[pick].to(self.codebook.device))
        self.ema_sum.copy_(self.codebook.data)
        self.ema_cnt.fill_(1.0)
        self._shadow_key = None

    @torch.no_grad()
    def reseed_dead_codes(self, z: torch.Tensor, generator: Optional[torch.Generator] = None) -> int:
        """Optional dead-code policy (off unless the caller invokes it): every code with
        steps_since_used >= dead_after is replaced by a random row of z.  Returns how many were replaced."""
        dead = (self.steps_since_used >= self.dead_after).nonzero()[:, 0]
        n = int(dead.numel())
        if n == 0:
            return 0
        rows = z.detach().reshape(-1, self.emb_dim).to(torch.float32)
        pick = torch.randint(0, rows.shape[0], (n,), generator=generator, device=rows.device)
        new = rows[pick].to(self.codebook.device)
        self.codebook.data[dead] = new
        self.ema_sum[dead] = new
        self.ema_cnt[dead] = 1.0
        self.steps_since_used[dead] = 0
        self._shadow_key = None
        return n

    @torch.no_grad()
    def sync_state_from(self, src: int = 0) -> None:
        """Data-parallel runs: make every replica start from rank `src`'s quantiser state (codebook, EMA accumulators,
        usage counters).  The buffers are initialised from the per-rank RNG, DistributedDataParallel is normally run with
        broadcast_buffers=False around this module (the EMA update is the kernels' job, not DDP's), and identical
        all-reduced updates applied to DIFFERENT states never converge to identical replicas -- so call this once after
        construction / data-dependent init / re-seeding, on every rank."""
        if not parallel.dist_active(self.process_group):
            return
        import torch.distributed as dist
        gsrc = dist.get_global_rank(self.process_group, src) if self.process_group is not None else src
        for t in (self.codebook.data, self.ema_cnt, self.ema_sum, self.usage_total, self.steps_since_used,
                  self.last_perplexity, self.n_dead):
            dist.broadcast(t, src=gsrc, group=self.process_group)
        self._shadow_key = None

    # ------------------------------------------------------------------ scratch management
    def _search_for(self, dtype: torch.dtype) -> str:
        """"auto": half-precision latents take the tensor-core search on operands of their own type (bf16 x bf16 shadow;
        fp16 -- the reference script's default autocast dtype, scripts/train_vqvae.py:284-286 -- x fp16 shadow: no
        staging pass, no cast), fp32 latents the fp32 tier (split-operand tensor-core search, CUDA cores for odd shapes).  An emb_dim the
        tensor-core kernel cannot take (not a multiple of 8, below 8 or above 512) goes to the fp32 kernel, which has no
        shape limits; only an explicit search="bf16" insists (and raises for such a shape)."""
        if self.search != "auto":
            return self.search
        D = self.emb_dim
        if D % 8 != 0 or D < 8 or D > 512:
            return "fp32"
        return {torch.bfloat16: "bf16", torch.float16: "fp16"}.get(dtype, "fp32")

    def _ensure_shadow(self, search: str) -> None:
        """(Re)build the bf16 copy and ||e||^2 the search reads if the codebook changed behind our back
        (optimizer step in st mode, load_state_dict, .to())."""
        E = self.codebook.detach()
        key = (E.data_ptr(), E._version, search)
        if key == self._shadow_key:
            return
        K, D = E.shape
        if (self._E_bf16 is None or self._E_bf16.device != E.device or self._E_bf16.dtype != ops.SHADOW_DTYPE[search]
                or self._E_bf16.numel() != ops.shadow_elems(K, D, search)):
            # 16-bit image of the codebook in the search's operand type: bf16 rows, fp16 rows, or the hi / lo bf16 planes
            # of the split fp32 tier
            self._E_bf16 = ops.new_shadow(K, D, search, E.device)
            if self._e_norm2 is None or self._e_norm2.device != E.device:
                self._e_norm2 = torch.empty(ops.norm_elems(K), dtype=torch.float32, device=E.device)
        ops.prepare_codebook(E, search, self._E_bf16, self._e_norm2)
        self._shadow_key = key

    def _workspace(self, N: int, dtype: torch.dtype, search: str = "bf16") -> torch.Tensor:
        need = self._ws_need.get((N, dtype, search))
        if need is None:                              # one ctypes call per new (rows, dtype, search), not per step
            if len(self._ws_need) > 64:
                self._ws_need.clear()
            need = self._ws_need[(N, dtype, search)] = ops.workspace_bytes(N, self.codebook_size, self.emb_dim, dtype,
                                                                           search)
        dev = self.codebook.device
        if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
            # zero-filled, and from here on used only by this module's presort / presorted calls: the counter region
            # stays zero between steps (VQ_SEARCH_WS_CLEAN), so the search needs no memset in front of it
            self._ws = torch.zeros(need, dtype=torch.uint8, device=dev)
        return self._ws

    def _use_peers(self) -> bool:
        """Fused a-6 + a-7 path: EMA training step, NCCL group, every rank on this node."""
        if self.peer_exchange == "nccl" or not (self.quantizer == "ema" and self.training and self._dist_active()):
            return False
        if self._peers is not None:
            return True
        import torch.distributed as dist
        world = dist.get_world_size(self.process_group)
        ok = dist.get_backend(self.process_group) == "nccl" and world <= int(os.environ.get("LOCAL_WORLD_SIZE", "1"))
        if not ok:
            if self.peer_exchange == "peer":
                raise RuntimeError("peer_exchange='peer' needs an NCCL process group whose ranks share one node")
            self.peer_exchange = "nccl"
            return False
        K, D = self.codebook_size, self.emb_dim
        ex = parallel.make_exchange(K * D + K, self.codebook.device, self.process_group, timeout_s=self.peer_timeout_s,
                                    reduced_numel=K * D)
        if self.peer_exchange == "auto" and ex.mc_stats() is None \
                and (K * (D + 1)) * 4 * world > PEER_ONE_SHOT_MAX_BYTES:
            # no in-switch reduction here, so every rank would read all W buffers (one-shot): right for small statistics
            # (latency-bound collective), wrong for large ones, where NCCL's bandwidth-optimal all-reduce moves
            # 2 (W-1)/W of one buffer instead.  With the multicast mapping each element crosses NVLink once: always fused.
            ex.close()
            self.peer_exchange = "nccl"
            return False
        self._peers = ex
        return True

    def _stats_buffer(self) -> torch.Tensor:
        K, D = self.codebook_size, self.emb_dim
        dev = self.codebook.device
        if self._use_peers():
            return self._peers.next_buffer()
        if self._stats is None or self._stats.device != dev:
            self._stats = torch.empty(K * D + K, dtype=torch.float32, device=dev)   # [code_sum | code_cnt]
        return self._stats

    def peer_exchange_timing(self, reset: bool = False) -> Optional[dict]:
        """Diagnostics of the fused peer exchange on this rank since the last reset (device -> host read): mean microseconds
        per step between this rank's signal and the last peer's (the wait for the slowest rank) and spent adding the W
        count vectors over NVLink.  reset=True restarts the averages."""
        if self._peers is None or self._ws is None:
            return None
        v = self._ws[:16].view(torch.float32).cpu()
        n = max(float(v[3]), 1.0)
        out = {"steps": int(v[3]), "wait_us": float(v[1]) / n, "read_counts_us": float(v[2]) / n}
        if reset:
            self._ws[4:16].zero_()
        return out

    def _dist_active(self) -> bool:
        return self.sync_stats and parallel.dist_active(self.process_group)

    # ------------------------------------------------------------------ one step of kernels (eager or as a CUDA graph)
    def _loss_scale(self, N: int) -> float:
        w = (1.0 + self.beta) if self.quantizer == "st" else self.beta
        return w / float(max(N * self.emb_dim, 1))

    def _step_kernels(self, z2, q, idx32, idx64, sqerr, loss, search: str, q_cf=None):
        """a-1 .. a-9 on preallocated tensors; every launch goes to the current stream (capturable).  The int64 indices
        and the scaled loss are by-products of the gather pass: no torch arithmetic on the path."""
        N, D = z2.shape
        K = self.codebook_size
        ws = self._workspace(N, z2.dtype, search)
        stats = self._stats_buffer()
        code_sum, code_cnt = stats[:K * D].view(K, D), stats[K * D:K * D + K]
        E = self.codebook.detach()
        # the search leaves the counting-sort inputs of idx in ws, the pass picks them up (one scatter launch instead of
        # memset + histogram + scan/scatter)
        kind = VQ_STAT_SUM_RESID if self.quantizer == "st" else VQ_STAT_SUM_Z
        ops.assign(z2, E, self._E_bf16, self._e_norm2, search, idx32, None, ws, presort=True, ws_clean=True)
        ops.gather_st_loss_scatter(z2, idx32, E, q, sqerr, code_cnt, code_sum, kind, self.deterministic, ws,
                                   idx64_out=idx64, loss_out=loss, loss_scale=self._loss_scale(N), presorted=True)
        if q_cf is not None:             # channel-first caller: q straight in [B, D, S] from the (final) indices, gathered
            ops.gather_rows_cf(idx32, E, q_cf)   # from the codebook the search used -- before the EMA step rewrites it
        self._finish_step(code_cnt, code_sum, search, ws)
        return code_sum

    def _graph_wanted(self, z2) -> bool:
        return (self.cuda_graph and z2.shape[0] > 0 and z2.numel() * z2.element_size() <= GRAPH_MAX_BYTES
                and not parallel.dist_active(self.process_group) and self._peers is None
                and not torch.cuda.is_current_stream_capturing())

    def _run_step_cf(self, z2, q_cf, search: str):
        """Channel-first variant of _run_step: the pass only reads the rows (no row-major q), q_cf [B, D, S] is written by
        the gather kernel.  -> (loss, idx int64, code_sum view)."""
        N, D = z2.shape
        dev = z2.device
        idx32 = torch.empty(N, dtype=torch.int32, device=dev)
        idx64 = torch.empty(N, dtype=torch.int64, device=dev)
        sqerr = torch.empty(1, dtype=torch.float32, device=dev)
        loss = torch.empty(1, dtype=torch.float32, device=dev)
        code_sum = self._step_kernels(z2, None, idx32, idx64, sqerr, loss, search, q_cf=q_cf)
        return loss, idx64, code_sum

    def _run_step(self, z2, search: str):
        """-> (q, sq_err_sum, idx int64, code_sum view).  Small single-process batches: first call with a given shape
        runs eagerly (it is also the warm-up), the second captures the step into a CUDA graph, later ones replay it."""
        N, D = z2.shape
        dev = z2.device
        if self._graph_wanted(z2):
            # every address the captured launches hold is part of the key: a moved module, a regrown workspace or a
            # re-created shadow simply never matches the stale entry again
            key = (N, z2.dtype, search, self.training, self.quantizer, self.deterministic, self.beta, self.ema_decay,
                   self.ema_eps, self.dead_after, self.codebook.data_ptr(),
                   self.ema_sum.data_ptr(), self._E_bf16.data_ptr() if self._E_bf16 is not None else 0,
                   self._e_norm2.data_ptr(), self._workspace(N, z2.dtype, search).data_ptr(), self._stats_buffer().data_ptr(),
                   dev.index)
            ent = self._graphs.get(key)
            if ent is None:
                if len(self._graphs) >= GRAPH_CACHE:
                    self._graphs.clear()
                self._graphs[key] = "seen"
            else:
                if ent == "seen":
                    try:
                        ent = self._capture_step(z2, search)
                    except Exception as exc:        # capture refused (exotic context): stay on eager launches for good
                        import warnings
                        warnings.warn(f"vqvae_b200: CUDA-graph capture of the quantiser step failed ({exc}); "
                                      "continuing with eager launches")
                        self.cuda_graph = False
                        self._graphs.clear()
                        ent = None
                    else:
                        self._graphs[key] = ent
                if ent is not None:
                    g, sz, sout, nq, code_sum, n_launch, _scratch = ent
                    sz.copy_(z2)
                    g.replay()
                    self.graph_launches += n_launch
                    # q, indices and loss live in ONE static byte buffer: one clone (one copy kernel, one allocation)
                    # hands all three to the caller -- autograd keeps q, so the outputs cannot alias the static buffer
                    out = sout.clone()
                    return (out[:N * D * z2.element_size()].view(z2.dtype).view(N, D),
                            out[nq + 8 * N:nq + 8 * N + 4].view(torch.float32),
                            out[nq:nq + 8 * N].view(torch.int64), code_sum)
        q = torch.empty((N, D), dtype=z2.dtype, device=dev)
        idx32 = torch.empty(N, dtype=torch.int32, device=dev)
        idx64 = torch.empty(N, dtype=torch.int64, device=dev)
        sqerr = torch.empty(1, dtype=torch.float32, device=dev)
        loss = torch.empty(1, dtype=torch.float32, device=dev)
        code_sum = self._step_kernels(z2, q, idx32, idx64, sqerr, loss, search)
        return q, loss, idx64, code_sum

    def _capture_step(self, z2, search: str):
        N, D = z2.shape
        dev = z2.device
        sz = torch.empty((N, D), dtype=z2.dtype, device=dev)
        nq = (N * D * z2.element_size() + 15) // 16 * 16       # [q | pad | idx int64 | loss fp32 | pad] in one buffer
        sout = torch.empty(nq + 8 * N + 16, dtype=torch.uint8, device=dev)
        sq_ = sout[:N * D * z2.element_size()].view(z2.dtype).view(N, D)
        si64 = sout[nq:nq + 8 * N].view(torch.int64)
        sloss = sout[nq + 8 * N:nq + 8 * N + 4].view(torch.float32)
        si32 = torch.empty(N, dtype=torch.int32, device=dev)
        ssq = torch.empty(1, dtype=torch.float32, device=dev)
        self._workspace(N, z2.dtype, search)         # sized before capture: the graph holds its address
        self._stats_buffer()
        g = torch.cuda.CUDAGraph()
        torch.cuda.synchronize(dev)
        l0 = ops.launch_count()
        # records, does not execute: module state is untouched.  "relaxed": nothing is allocated or synchronised between
        # begin and end of the capture here, and event queries made meanwhile by the caching allocators (record_stream'd
        # tensors, a DataLoader's pin-memory thread) must not invalidate it
        with torch.cuda.graph(g, capture_error_mode="relaxed"):
            code_sum = self._step_kernels(sz, sq_, si32, si64, ssq, sloss, search)
        # (si32, ssq): scratch the captured launches write; it must live as long as the graph does
        return (g, sz, sout, nq, code_sum, ops.launch_count() - l0, (si32, ssq))

    # ------------------------------------------------------------------ a-6 .. a-9
    def _finish_step(self, code_cnt, code_sum, search: str, ws) -> None:
        ema = self.quantizer == "ema" and self.training
        if ema and self._peers is not None and self._use_peers():
            E = self.codebook.detach()
            px = self._peers
            mc_red, red = px.reduced()
            ops.ema_update_dp(E, self.ema_cnt, self.ema_sum, self.ema_decay, self.ema_eps, True, search, self._E_bf16,
                              self._e_norm2, self.usage_total, self.steps_since_used, self.last_perplexity, self.n_dead,
                              self.dead_after, True, ws, px.peer_stats(), px.peer_flags(), px.rank, px.epoch,
                              mc_stats=px.mc_stats(), mc_reduced=mc_red, reduced_local=red)
            self._shadow_key = (E.data_ptr(), E._version, search)
            return
        if self._dist_active():
            if ema:      # one collective over the contiguous [sum | cnt] buffer
                parallel.allreduce_code_stats(self._stats, self.process_group)
            else:        # st / eval: only the counts feed usage and perplexity
                code_cnt = parallel.allreduce_code_stats(code_cnt.clone(), self.process_group)
        E = self.codebook.detach()
        ops.ema_update(E, self.ema_cnt, self.ema_sum, code_cnt, code_sum, self.ema_decay, self.ema_eps, ema, search,
                       self._E_bf16, self._e_norm2, self.usage_total, self.steps_since_used, self.last_perplexity,
                       self.n_dead, self.dead_after, self.training, ws)
        # the kernel refreshed the shadow for the (possibly updated) codebook
        self._shadow_key = (E.data_ptr(), E._version, search)

    # ------------------------------------------------------------------ forward
    def forward(self, z: torch.Tensor):
        if not z.is_cuda:
            raise RuntimeError("VectorQuantizer.forward needs a CUDA tensor: the kernels are sm_100a only and "
                               "there is no CPU fallback (attribute access and state_dict work on CPU)")
        if z.device.index != torch.cuda.current_device():
            # launches, streams, graph capture and scratch allocation all follow the tensor's device, not whatever
            # device the calling thread has selected
            with torch.cuda.device(z.device):
                return self.forward(z)
        if self.channel_first:
            return self._forward_channel_first(z)
        if z.shape[-1] != self.emb_dim:
            raise ValueError(f"last dimension {z.shape[-1]} != emb_dim {self.emb_dim}")
        # the autocast guard and the autograd node cost ~30 us of host time per call, which is what bounds a small-batch
        # step (C1 replays in ~55 us of GPU time): enter them only when they have something to do
        with (torch.autocast(device_type="cuda", enabled=False) if torch.is_autocast_enabled("cuda") else _NULL_CTX):
            lead = z.shape[:-1]
            in_dtype = z.dtype
            z2 = z.reshape(-1, self.emb_dim)
            search = self._search_for(z2.dtype)
            if z2.dtype not in (torch.bfloat16, torch.float16, torch.float32):
                z2 = z2.to(torch.float32)        # exotic dtypes only; bf16 / fp16 / fp32 rows go to the kernels as they are
            if z2.shape[0] > 1 and (z2.stride(1) != 1 or z2.stride(0) < self.emb_dim):
                z2 = z2.contiguous()
            if torch.is_grad_enabled() and (z2.requires_grad or self.codebook.requires_grad):
                q, loss, idx = _VQFunction.apply(z2, self.codebook, self, search)
            else:
                self._ensure_shadow(search)
                q, loss, idx, _ = self._run_step(z2, search)
                loss = loss.reshape(())
            if q.dtype != in_dtype:
                q = q.to(in_dtype)
            return q.reshape(*lead, self.emb_dim), loss, idx.reshape(lead)

    @torch.no_grad()
    def forward_host(self, z_host: torch.Tensor, chunks: Optional[int] = None, q_host: Optional[torch.Tensor] = None,
                     idx_host: Optional[torch.Tensor] = None):
        """Quantise a batch that lives in (pinned) HOST memory: `[..., D]` CPU tensor in, device tensors
        `(quantized, loss, indices)` out.  With `q_host` / `idx_host` (pinned CPU tensors of the output shapes: the
        latents' dtype and int64) the results also travel BACK, chunk by chunk on a third stream while later chunks are
        still being copied in and quantised (PCIe is full duplex), and those host tensors are returned in place of the
        device ones -- valid once the current stream has been synchronised.  The batch is cut into row chunks; chunk c + 1 crosses PCIe on a copy stream
        while the kernels work on chunk c (the search per chunk, the gather / loss / statistics pass accumulating
        across chunks with VQ_STAT_ACCUMULATE), then one EMA update for the whole batch -- the same arithmetic as
        `forward` on the copied tensor, with the transfer hidden behind the compute instead of in front of it.
        Inference / EMA training only (no autograd graph: the input is not a device tensor)."""
        if z_host.is_cuda:
            raise ValueError("forward_host takes a CPU tensor; call forward() for device tensors")
        if self.channel_first:
            raise ValueError("forward_host expects channel-last rows [..., D]")
        if z_host.shape[-1] != self.emb_dim:
            raise ValueError(f"last dimension {z_host.shape[-1]} != emb_dim {self.emb_dim}")
        search = self._search_for(z_host.dtype)
        if z_host.dtype not in (torch.bfloat16, torch.float16, torch.float32):
            z_host = z_host.to(torch.float32)
        for name, t, dt in (("q_host", q_host, z_host.dtype), ("idx_host", idx_host, torch.int64)):
            if t is not None and (t.is_cuda or t.dtype != dt or not t.is_contiguous() or not t.is_pinned()):
                raise ValueError(f"forward_host: {name} must be a contiguous pinned CPU tensor of dtype {dt}")
        dev = self.codebook.device
        if dev.type != "cuda":
            raise RuntimeError("VectorQuantizer.forward_host: the module must live on a CUDA device (no CPU fallback)")
        if dev.index != torch.cuda.current_device():
            with torch.cuda.device(dev):
                return self.forward_host(z_host, chunks)
        lead, D, K = z_host.shape[:-1], self.emb_dim, self.codebook_size
        zh = z_host.reshape(-1, D)
        if not zh.is_contiguous():
            zh = zh.contiguous()
        if not zh.is_pinned():
            zh = zh.pin_memory()
        N = zh.shape[0]
        if chunks is None and N > 0 and self._graph_wanted(zh):
            # a small batch is one chunk: nothing to hide the copy under, so it goes in front of the graph-replayed step
            self._ensure_shadow(search)
            q, loss, idx64, _ = self._run_step(zh.to(dev, non_blocking=True), search)
            if q.dtype != z_host.dtype:
                q = q.to(z_host.dtype)
            if q_host is not None:
                q_host.view(-1, D).copy_(q, non_blocking=True)
                q = q_host
            if idx_host is not None:
                idx_host.view(-1).copy_(idx64, non_blocking=True)
                idx64 = idx_host
            return q.reshape(*lead, D), loss.reshape(()), idx64.reshape(lead)
        if chunks is None:                      # ~128 K rows per chunk: transfer time per chunk well above launch costs
            chunks = max(1, min(16, N // 131072))
        rows = max(128, -(-N // max(1, chunks)) // 128 * 128) if N > 0 else 1
        bounds = [(s, min(N, s + rows)) for s in range(0, max(N, 1), rows)] if N > 0 else []
        self._ensure_shadow(search)
        cur = torch.cuda.current_stream(dev)
        if getattr(self, "_copy_stream", None) is None or self._copy_stream.device != dev:
            self._copy_stream = torch.cuda.Stream(device=dev)
        copy = self._copy_stream
        z_dev = torch.empty((N, D), dtype=zh.dtype, device=dev)
        q = torch.empty((N, D), dtype=zh.dtype, device=dev)
        idx32 = torch.empty(N, dtype=torch.int32, device=dev)
        idx64 = torch.empty(N, dtype=torch.int64, device=dev)
        sqerr = torch.zeros(1, dtype=torch.float32, device=dev)
        loss = torch.zeros(1, dtype=torch.float32, device=dev)
        stats = self._stats_buffer()
        code_sum, code_cnt = stats[:K * D].view(K, D), stats[K * D:K * D + K]
        ws = self._workspace(rows if N > 0 else 0, zh.dtype, search)
        E = self.codebook.detach()
        kind = VQ_STAT_SUM_RESID if self.quantizer == "st" else VQ_STAT_SUM_Z
        copy.wait_stream(cur)                   # the fresh device buffers are not reused memory still in flight
        back = None
        if (q_host is not None or idx_host is not None) and N > 0:
            if getattr(self, "_back_stream", None) is None or self._back_stream.device != dev:
                self._back_stream = torch.cuda.Stream(device=dev)
            back = self._back_stream
            qh2 = q_host.view(-1, D) if q_host is not None else None
            ih1 = idx_host.view(-1) if idx_host is not None else None
        landed = []
        with torch.cuda.stream(copy):
            for a, b in bounds:
                z_dev[a:b].copy_(zh[a:b], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(copy)
                landed.append(ev)
        for c, (a, b) in enumerate(bounds):
            cur.wait_event(landed[c])
            ops.assign(z_dev[a:b], E, self._E_bf16, self._e_norm2, search, idx32[a:b], None, ws, presort=True, ws_clean=True)
            ops.gather_st_loss_scatter(z_dev[a:b], idx32[a:b], E, q[a:b], sqerr, code_cnt, code_sum, kind,
                                       self.deterministic, ws, accumulate=c > 0, idx64_out=idx64[a:b], loss_out=loss,
                                       loss_scale=self._loss_scale(N), presorted=True)
            if back is not None:                # this chunk's results go home under the next chunk's kernels
                back.wait_stream(cur)
                with torch.cuda.stream(back):
                    if qh2 is not None:
                        qh2[a:b].copy_(q[a:b], non_blocking=True)
                    if ih1 is not None:
                        ih1[a:b].copy_(idx64[a:b], non_blocking=True)
        if not bounds:
            ops.gather_st_loss_scatter(z_dev, idx32, E, q, sqerr, code_cnt, code_sum, kind, self.deterministic, ws,
                                       idx64_out=idx64, loss_out=loss, loss_scale=self._loss_scale(N))
        z_dev.record_stream(copy)
        self._finish_step(code_cnt, code_sum, search, ws)
        if back is not None:
            q.record_stream(back)
            idx64.record_stream(back)
            cur.wait_stream(back)               # synchronising the caller's stream makes the host tensors valid
            if q_host is not None:
                q = q_host
            if idx_host is not None:
                idx64 = idx_host
        if q.dtype != z_host.dtype:
            q = q.to(z_host.dtype)
        return q.reshape(*lead, D), loss.reshape(()), idx64.reshape(lead)

    def _forward_channel_first(self, z: torch.Tensor):
        """[B, D, *spatial] in, quantized in the same layout out.  One transpose pass in (the search and the sorted pass
        need row-major latents); NONE out: the pass only reads the rows and vq_gather_rows_cf writes q straight in
        channel-first layout from the indices; the backward works on the channel-first tensors directly (elementwise)."""
        if z.dim() < 2 or z.shape[1] != self.emb_dim:
            raise ValueError(f"channel_first: dimension 1 is {tuple(z.shape)[1:2]}, expected emb_dim {self.emb_dim}")
        with torch.autocast(device_type="cuda", enabled=False):
            B, D, spatial = z.shape[0], z.shape[1], z.shape[2:]
            z3 = z.reshape(B, D, -1)
            search = self._search_for(z3.dtype)
            in_dtype = z3.dtype
            if z3.dtype not in (torch.bfloat16, torch.float16, torch.float32):
                z3 = z3.to(torch.float32)
            if not z3.is_contiguous():
                z3 = z3.contiguous()
            qcf, loss, idx = _VQFunctionCF.apply(z3, self.codebook, self, search)
            if qcf.dtype != in_dtype:
                qcf = qcf.to(in_dtype)
            return qcf.reshape(B, D, *spatial), loss, idx.reshape(B, *spatial)


class SplitQuantizer(nn.Module):
    """Two codebooks for the live model's two latent streams (BASELINE config 3): a `type` codebook over
    z_type [B, D_type, H, W] and a `phase` codebook over z_phase [B, D_phase, T, H, W] (attributes `quant_type`, `quant_phase`)
    (frl/models/representation.py:60-72, :317-374; both channel-first).  The two quantisers share nothing; they run back to
    back on the caller's stream.  `overlap=True` puts the type stream (T times fewer vectors) on a side CUDA stream under
    the phase stream's kernels: 5 % faster when it works, but two persistent one-CTA-per-SM searches compete for the same
    SMs and one step in three to five ran 1.4-1.9x slower on B200 (profiles/r01/tuning_experiments.log), so it is off by
    default (and switches programmatic dependent launch off while both chains are in flight).

    forward(z_type, z_phase) -> ((q_type, loss_type, idx_type), (q_phase, loss_phase, idx_phase)); `loss` is the sum
    the training step adds to its objective, `perplexity` a dict.
    """

    def __init__(self, type_dim: int, phase_dim: int, codebook_size: int, channel_first: bool = True,
                 overlap: bool = False, **kw):
        super().__init__()
        self.quant_type = VectorQuantizer(type_dim, codebook_size, channel_first=channel_first, **kw)
        self.quant_phase = VectorQuantizer(phase_dim, codebook_size, channel_first=channel_first, **kw)
        self.overlap = bool(overlap)
        self._side: Optional[torch.cuda.Stream] = None
        self.loss: Optional[torch.Tensor] = None

    @property
    def perplexity(self) -> dict:
        return {"type": self.quant_type.perplexity, "phase": self.quant_phase.perplexity}

    @torch.no_grad()
    def forward_host(self, z_type_host: torch.Tensor, z_phase_host: torch.Tensor):
        """Both streams from (pinned) host memory, channel-last rows: each goes through its quantiser's pipelined
        `forward_host` (copy of chunk c + 1 under the kernels of chunk c); the phase stream's copies start while the
        type stream's last chunk is still being quantised."""
        out_t = self.quant_type.forward_host(z_type_host)
        out_p = self.quant_phase.forward_host(z_phase_host)
        self.loss = out_t[1] + out_p[1]
        return out_t, out_p

    def forward(self, z_type: torch.Tensor, z_phase: torch.Tensor):
        if not (z_type.is_cuda and z_phase.is_cuda):
            raise RuntimeError("SplitQuantizer.forward needs CUDA tensors (no CPU fallback)")
        if not self.overlap:
            out_t, out_p = self.quant_type(z_type), self.quant_phase(z_phase)
        else:
            cur = torch.cuda.current_stream(z_type.device)
            if self._side is None or self._side.device != z_type.device:
                self._side = torch.cuda.Stream(device=z_type.device)
            self._side.wait_stream(cur)                      # z_type was produced on the caller's stream
            # two kernel chains in flight at once: no programmatic dependent launch (a chain's early-resident grids can
            # keep the other chain's CTA pairs off the SMs: 0.89 -> 1.2 ms in one step out of three)
            prev_mask = ops.set_pdl_mask(0)
            try:
                with torch.cuda.stream(self._side):
                    out_t = self.quant_type(z_type)
                out_p = self.quant_phase(z_phase)
            finally:
                ops.set_pdl_mask(prev_mask)
            cur.wait_stream(self._side)
            for t in out_t:                                  # allocated on the side stream, consumed on the caller's
                t.record_stream(cur)
            z_type.record_stream(self._side)
        self.loss = out_t[1] + out_p[1]
        return out_t, out_p

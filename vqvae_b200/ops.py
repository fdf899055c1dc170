"""Thin torch-tensor wrappers over the C ABI (include/vq_b200.h).

torch is used for device memory and the current stream only; every computation is a kernel of
libvq_b200.so.  Each function takes CUDA tensors, validates dtype/contiguity, and passes raw pointers.
A CPU tensor raises: there is no fallback (SURVEY.md 8b "Device / dtype").
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import (VQ_DT_BF16, VQ_DT_F16, VQ_DT_F32, VQ_SEARCH_BF16, VQ_SEARCH_F16, VQ_SEARCH_F32, VQ_SEARCH_SIMT,
                   VQ_STAT_SUM_RESID, VQ_STAT_SUM_Z, check)

# search kinds (include/vq_b200.h): "bf16" / "fp16" = tensor-core search on operands of that type; "fp32" = the 1e-5 tier
# (split-operand tensor-core search when the shape allows and the split shadow is given, CUDA-core kernel otherwise);
# "fp32_simt" forces the CUDA-core kernel
SEARCH_MODES = {"bf16": VQ_SEARCH_BF16, "fp32": VQ_SEARCH_F32, "fp16": VQ_SEARCH_F16,
                "fp32_simt": VQ_SEARCH_F32 | VQ_SEARCH_SIMT}
_DTYPES = {torch.bfloat16: VQ_DT_BF16, torch.float32: VQ_DT_F32, torch.float16: VQ_DT_F16}
SHADOW_DTYPE = {"bf16": torch.bfloat16, "fp16": torch.float16, "fp32": torch.bfloat16, "fp32_simt": torch.bfloat16}


def _dt(t: torch.Tensor) -> int:
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise TypeError(f"vqvae_b200 kernels take bfloat16, float16 or float32 latents, got {t.dtype}") from None


class _On:
    """Device guard of one op call: every tensor argument must live on ONE CUDA device; the launch goes to that device's
    current stream with that device current (the library sizes grids and caches function attributes per current device),
    whatever device the caller's thread happens to have selected."""
    __slots__ = ("dev", "prev")

    def __init__(self, *ts: Optional[torch.Tensor]):
        dev = None
        for t in ts:
            if t is None:
                continue
            if not t.is_cuda:
                raise RuntimeError("vqvae_b200: this op runs only on CUDA tensors (sm_100a kernels, no CPU fallback); "
                                   "use oracle/vq_reference.py for CPU checks")
            if dev is None:
                dev = t.device
            elif t.device != dev:
                raise RuntimeError(f"vqvae_b200: tensor arguments live on different devices ({dev} and {t.device})")
        self.dev, self.prev = dev, -1

    def __enter__(self):
        if self.dev is not None:
            cur = torch.cuda.current_device()
            if cur != self.dev.index:
                self.prev = cur
                torch.cuda.set_device(self.dev)
        return self

    def __exit__(self, *exc):
        if self.prev >= 0:
            torch.cuda.set_device(self.prev)
        return False

    @property
    def stream(self) -> int:
        return torch.cuda.current_stream(self.dev).cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _rows(z: torch.Tensor) -> Tuple[int, int, int]:
    """[N, D] with unit inner stride -> (N, D, row stride in elements)."""
    if z.dim() != 2 or (z.shape[1] > 1 and z.stride(1) != 1):
        raise ValueError("expected a [N, D] tensor with contiguous rows")
    N, D = z.shape
    ld = z.stride(0) if N > 1 else D
    return N, D, max(ld, D)


def norm_elems(K: int) -> int:
    return _lib.load().vq_norm_elems(K)


def workspace_bytes(N: int, K: int, D: int, dtype: torch.dtype, search: Optional[str] = None) -> int:
    """Scratch bytes of one step; `search` defaults to the kind the dtype takes natively (bf16 / fp16 latents: their own
    tensor-core search, fp32 latents: "fp32")."""
    out = C.c_size_t(0)
    if search is None:
        search = {torch.bfloat16: "bf16", torch.float16: "fp16"}.get(dtype, "fp32")
    check(_lib.load().vq_workspace_bytes(N, K, D, _DTYPES[dtype], SEARCH_MODES[search] & 3, C.byref(out)),
          "vq_workspace_bytes")
    return int(out.value)


def shadow_elems(K: int, D: int, search: str) -> int:
    """16-bit elements the search shadow holds for this search kind (K * D; hi + lo planes [2][K_pad][D] for "fp32")."""
    return int(_lib.load().vq_shadow_elems(K, D, SEARCH_MODES[search] & 3))


def new_shadow(K: int, D: int, search: str, device) -> torch.Tensor:
    """Flat shadow buffer of the right size and element type for `search` (filled by prepare_codebook / ema_update)."""
    return torch.empty(shadow_elems(K, D, search), dtype=SHADOW_DTYPE[search], device=device)


def set_pdl_mask(mask: int) -> int:
    """Programmatic dependent launch along the step's kernel chain (include/vq_b200.h: vq_set_pdl_mask); returns the
    previous mask.  0 = plain stream order (what concurrent chains on several streams should use)."""
    return int(_lib.load().vq_set_pdl_mask(int(mask)))


def set_peer_timeout_ms(ms: int) -> int:
    """Wall-clock bound of the cross-rank wait of the fused peer exchange (include/vq_b200.h); returns the previous one."""
    return int(_lib.load().vq_set_peer_timeout_ms(int(ms)))


def launch_count() -> int:
    return int(_lib.load().vq_launch_count())


def _check_shadow(shadow: Optional[torch.Tensor], K: int, D: int, search: str) -> None:
    if shadow is None:
        return
    if shadow.element_size() != 2 or not shadow.is_contiguous() or shadow.numel() < shadow_elems(K, D, search):
        raise ValueError(f"search shadow for search={search!r} must be a contiguous 16-bit tensor of at least "
                         f"{shadow_elems(K, D, search)} elements (ops.new_shadow), got {tuple(shadow.shape)} {shadow.dtype}")


def prepare_codebook(E: torch.Tensor, search: str, E_bf16: Optional[torch.Tensor], e_norm2: torch.Tensor) -> None:
    with _On(E, E_bf16, e_norm2) as on:
        K, D = E.shape
        assert E.dtype == torch.float32 and E.is_contiguous()
        assert e_norm2.dtype == torch.float32 and e_norm2.numel() >= norm_elems(K)
        _check_shadow(E_bf16, K, D, search)
        check(_lib.load().vq_prepare_codebook(_ptr(E), K, D, SEARCH_MODES[search] & 3, _ptr(E_bf16), _ptr(e_norm2),
                                              on.stream), "vq_prepare_codebook")


def assign(z: torch.Tensor, E: torch.Tensor, E_bf16: Optional[torch.Tensor], e_norm2: torch.Tensor, search: str,
           idx: torch.Tensor, smin: Optional[torch.Tensor], ws: Optional[torch.Tensor], presort: bool = False,
           ws_clean: bool = False) -> None:
    """presort: the search also leaves the counting-sort inputs of idx (histogram, ranks, scan) in `ws`, so the
    following gather_st_loss_scatter(..., presorted=True) on the same ws skips its histogram launches.
    ws_clean: `ws` was zero-filled at allocation and has only ever been used by presort / presorted calls since."""
    with _On(z, E, E_bf16, e_norm2, idx, smin, ws) as on:
        N, D, ld = _rows(z)
        K = E.shape[0]
        assert idx.dtype == torch.int32 and idx.numel() >= N and idx.is_contiguous()
        assert smin is None or (smin.dtype == torch.float32 and smin.numel() >= N)
        mode = SEARCH_MODES[search] | (_lib.VQ_SEARCH_PRESORT if presort else 0) | \
            (_lib.VQ_SEARCH_WS_CLEAN if presort and ws_clean else 0)
        check(_lib.load().vq_assign(_ptr(z), _dt(z), N, D, ld, _ptr(E), _ptr(E_bf16), _ptr(e_norm2), K,
                                    mode, _ptr(idx), _ptr(smin), _ptr(ws), on.stream), "vq_assign")


def gather_st_loss_scatter(z: torch.Tensor, idx: torch.Tensor, E: torch.Tensor, q_out: torch.Tensor,
                           sq_err_sum: torch.Tensor, code_cnt: Optional[torch.Tensor],
                           code_sum: Optional[torch.Tensor], stat_kind: int, deterministic: bool,
                           ws: Optional[torch.Tensor], accumulate: bool = False, idx64_out: Optional[torch.Tensor] = None,
                           loss_out: Optional[torch.Tensor] = None, loss_scale: float = 0.0,
                           presorted: bool = False) -> None:
    """presorted: idx was written by assign(..., presort=True) with the same z rows and ws (nothing touched since).
    accumulate: add to sq_err_sum / code_cnt / code_sum instead of overwriting (row chunks of one batch).
    idx64_out [N] int64 / loss_out [1] fp32 (optional): the final indices widened to int64 and sq_err_sum * loss_scale,
    written by the pass itself (no torch elementwise kernels on the path)."""
    with _On(z, idx, E, q_out, sq_err_sum, code_cnt, code_sum, ws, idx64_out, loss_out) as on:
        assert idx64_out is None or (idx64_out.dtype == torch.int64 and idx64_out.is_contiguous() and idx64_out.numel() >= z.shape[0])
        assert loss_out is None or loss_out.dtype == torch.float32
        if accumulate:
            stat_kind |= _lib.VQ_STAT_ACCUMULATE
        if presorted:
            stat_kind |= _lib.VQ_STAT_PRESORTED
        N, D, ld = _rows(z)
        ldq = _rows(q_out)[2] if q_out is not None else D        # q_out None: loss and statistics only (gather_rows_cf)
        K = E.shape[0]
        assert q_out is None or (q_out.dtype == z.dtype and q_out.shape == z.shape)
        assert idx.dtype == torch.int32 and E.dtype == torch.float32 and E.is_contiguous()
        check(_lib.load().vq_gather_st_loss_scatter(_ptr(z), _dt(z), N, D, ld, _ptr(idx), _ptr(E), K, _ptr(q_out),
                                                    _dt(z), ldq, _ptr(sq_err_sum), _ptr(code_cnt), _ptr(code_sum),
                                                    stat_kind, int(bool(deterministic)), _ptr(idx64_out), _ptr(loss_out),
                                                    float(loss_scale), _ptr(ws), on.stream),
              "vq_gather_st_loss_scatter")


def ema_update(E: torch.Tensor, ema_cnt: Optional[torch.Tensor], ema_sum: Optional[torch.Tensor],
               code_cnt: torch.Tensor, code_sum: Optional[torch.Tensor], decay: float, eps: float, do_ema: bool,
               search: str, E_bf16: Optional[torch.Tensor], e_norm2: torch.Tensor, usage_total: torch.Tensor,
               steps_since_used: torch.Tensor, perplexity: torch.Tensor, n_dead: torch.Tensor, dead_after: int,
               update_usage: bool, ws: torch.Tensor) -> None:
    with _On(E, ema_cnt, ema_sum, code_cnt, code_sum, E_bf16, e_norm2, usage_total, steps_since_used, perplexity,
               n_dead, ws) as on:
        K, D = E.shape
        assert usage_total.dtype == torch.int64 and steps_since_used.dtype == torch.int32 and n_dead.dtype == torch.int32
        assert perplexity.dtype == torch.float32
        _check_shadow(E_bf16, K, D, search)
        check(_lib.load().vq_ema_update(_ptr(E), _ptr(ema_cnt), _ptr(ema_sum), _ptr(code_cnt), _ptr(code_sum), K, D,
                                        float(decay), float(eps), int(bool(do_ema)), SEARCH_MODES[search] & 3, _ptr(E_bf16),
                                        _ptr(e_norm2), _ptr(usage_total), _ptr(steps_since_used), _ptr(perplexity),
                                        _ptr(n_dead), int(dead_after), int(bool(update_usage)), _ptr(ws), on.stream),
              "vq_ema_update")


def backward(z: torch.Tensor, q: torch.Tensor, grad_q: Optional[torch.Tensor], grad_loss: torch.Tensor, coef: float,
             grad_z: torch.Tensor) -> None:
    with _On(z, q, grad_q, grad_loss, grad_z) as on:
        N, D, ld = _rows(z)
        assert q.is_contiguous() and grad_z.is_contiguous() and (grad_q is None or grad_q.is_contiguous())
        assert q.dtype == z.dtype and grad_z.dtype == z.dtype and (grad_q is None or grad_q.dtype == z.dtype)
        assert grad_loss.dtype == torch.float32 and grad_loss.numel() == 1
        check(_lib.load().vq_backward(_ptr(z), _ptr(q), _ptr(grad_q), N, D, ld, _ptr(grad_loss), float(coef),
                                      _ptr(grad_z), _dt(z), on.stream), "vq_backward")


def codebook_grad(resid_sum: torch.Tensor, grad_loss: torch.Tensor, coef: float, grad_E: torch.Tensor) -> None:
    with _On(resid_sum, grad_loss, grad_E) as on:
        K, D = resid_sum.shape
        assert resid_sum.dtype == torch.float32 and grad_E.dtype == torch.float32 and grad_E.is_contiguous()
        check(_lib.load().vq_codebook_grad(_ptr(resid_sum), K, D, _ptr(grad_loss), float(coef), _ptr(grad_E), on.stream),
              "vq_codebook_grad")


def cast_bf16(src: torch.Tensor, dst: torch.Tensor) -> None:
    with _On(src, dst) as on:
        N, D, ld = _rows(src)
        assert src.dtype == torch.float32 and dst.dtype == torch.bfloat16 and dst.is_contiguous()
        check(_lib.load().vq_cast_bf16(_ptr(src), N, D, ld, _ptr(dst), on.stream), "vq_cast_bf16")


def gather_rows_cf(idx: torch.Tensor, E: torch.Tensor, q_cf: torch.Tensor) -> None:
    """q_cf [B, D, S] (dense) <- E[idx[b * S + s]][d] cast to q_cf's dtype: the quantised output straight in channel-first
    layout (include/vq_b200.h: vq_gather_rows_cf)."""
    with _On(idx, E, q_cf) as on:
        B, D, S = q_cf.shape
        assert q_cf.is_contiguous() and idx.dtype == torch.int32 and idx.numel() >= B * S and idx.is_contiguous()
        assert E.dtype == torch.float32 and E.is_contiguous() and E.shape[1] == D
        check(_lib.load().vq_gather_rows_cf(_ptr(idx), _ptr(E), E.shape[0], D, B, S, _ptr(q_cf), _dt(q_cf), on.stream),
              "vq_gather_rows_cf")


def layout_transpose(src: torch.Tensor, dst: torch.Tensor) -> None:
    """src [B, R, C] -> dst [B, C, R], both dense (a-0: channel-first <-> row layout)."""
    with _On(src, dst) as on:
        B, R, C = src.shape
        assert src.is_contiguous() and dst.is_contiguous() and dst.shape == (B, C, R) and dst.dtype == src.dtype
        check(_lib.load().vq_layout_transpose(_ptr(src), _ptr(dst), B, R, C, _dt(src), on.stream), "vq_layout_transpose")


def ema_update_dp(E: torch.Tensor, ema_cnt: torch.Tensor, ema_sum: torch.Tensor, decay: float, eps: float,
                  do_ema: bool, search: str, E_bf16: Optional[torch.Tensor], e_norm2: torch.Tensor,
                  usage_total: torch.Tensor, steps_since_used: torch.Tensor, perplexity: torch.Tensor,
                  n_dead: torch.Tensor, dead_after: int, update_usage: bool, ws: torch.Tensor, peer_stats, peer_flags,
                  rank: int, epoch: int, mc_stats: Optional[int] = None, mc_reduced: Optional[int] = None,
                  reduced_local: Optional[int] = None) -> None:
    """a-6 fused into a-7..a-9: the update kernels add every rank's statistics through peer memory
    (peer_stats / peer_flags: lists of `world` raw device pointers, own buffers at index `rank`)."""
    with _On(E, ema_cnt, ema_sum, E_bf16, e_norm2, usage_total, steps_since_used, perplexity, n_dead, ws) as on:
        K, D = E.shape
        world = len(peer_stats)
        ps = (C.c_void_p * world)(*[int(p) for p in peer_stats])
        pf = (C.c_void_p * world)(*[int(p) for p in peer_flags])
        _check_shadow(E_bf16, K, D, search)
        check(_lib.load().vq_ema_update_dp(_ptr(E), _ptr(ema_cnt), _ptr(ema_sum), K, D, float(decay), float(eps),
                                           int(bool(do_ema)), SEARCH_MODES[search] & 3, _ptr(E_bf16), _ptr(e_norm2),
                                           _ptr(usage_total), _ptr(steps_since_used), _ptr(perplexity), _ptr(n_dead),
                                           int(dead_after), int(bool(update_usage)), _ptr(ws), ps, pf, world, int(rank),
                                           int(epoch), mc_stats or None, mc_reduced or None, reduced_local or None,
                                           on.stream), "vq_ema_update_dp")

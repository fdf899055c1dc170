"""Shared helpers for the -m gpu parity tests: run one step of the CUDA path through the C ABI."""
import torch

from oracle import vq_reference as R
from vqvae_b200 import ops
from vqvae_b200._lib import VQ_STAT_SUM_RESID, VQ_STAT_SUM_Z


def cuda_step(z_cpu, E_cpu, search, dtype=torch.float32, deterministic=False, stat_kind=VQ_STAT_SUM_Z, stats=True,
              want_smin=True):
    """assign + fused pass on the GPU.  Returns dict of CPU tensors."""
    dev = torch.device("cuda:0")
    z = z_cpu.to(dev).to(dtype).contiguous()
    E = E_cpu.to(dev).float().contiguous()
    N, D = z.shape
    K = E.shape[0]
    E_bf16 = ops.new_shadow(K, D, search, dev)       # bf16 rows / fp16 rows / hi + lo planes of the split fp32 tier
    norms = torch.empty(ops.norm_elems(K), dtype=torch.float32, device=dev)
    ops.prepare_codebook(E, search, E_bf16, norms)
    ws = torch.empty(ops.workspace_bytes(N, K, D, dtype, search), dtype=torch.uint8, device=dev)
    idx = torch.full((N,), -1, dtype=torch.int32, device=dev)
    smin = torch.empty(N, dtype=torch.float32, device=dev) if want_smin else None
    # (1) stand-alone calls: the pass builds its own counting sort
    ops.assign(z, E, E_bf16, norms, search, idx, smin, ws)
    raw_idx = idx.clone()
    q = torch.empty_like(z)
    sq = torch.empty(1, dtype=torch.float32, device=dev)
    cnt = torch.empty(K, dtype=torch.float32, device=dev) if stats else None
    sums = torch.empty((K, D), dtype=torch.float32, device=dev) if stats else None
    ops.gather_st_loss_scatter(z, idx, E, q, sq, cnt, sums, stat_kind, deterministic, ws)
    torch.cuda.synchronize()
    # (2) the production pairing: the search leaves histogram / ranks / scan in ws (VQ_SEARCH_PRESORT) and the pass
    # picks them up (VQ_STAT_PRESORTED).  Must agree with (1); its results are the ones returned.
    first = dict(idx=idx.clone(), q=q.clone(), sq=sq.clone(), cnt=None if cnt is None else cnt.clone(),
                 sums=None if sums is None else sums.clone())
    idx.fill_(-1)
    q.fill_(float("nan"))
    if stats:
        cnt.fill_(-7.0)
        sums.fill_(float("nan"))
    ops.assign(z, E, E_bf16, norms, search, idx, smin, ws, presort=True)
    assert torch.equal(idx, raw_idx), "VQ_SEARCH_PRESORT changed the indices"
    ops.gather_st_loss_scatter(z, idx, E, q, sq, cnt, sums, stat_kind, deterministic, ws, presorted=True)
    torch.cuda.synchronize()
    assert torch.equal(idx, first["idx"]) and torch.equal(q.view(torch.uint8), first["q"].view(torch.uint8))
    sq_a, sq_b = float(first["sq"]), float(sq)
    assert sq_a == sq_b or abs(sq_a - sq_b) <= 1e-5 * abs(sq_a), (sq_a, sq_b)
    if stats:
        assert torch.equal(cnt, first["cnt"]), "presorted pass: per-code counts differ"
        fin = torch.isfinite(first["sums"])
        assert torch.equal(fin, torch.isfinite(sums))
        scale = first["sums"][fin].abs().max().clamp_min(1e-30) if bool(fin.any()) else 1.0
        assert float(((sums - first["sums"])[fin].abs().max() if bool(fin.any()) else 0.0) / scale) <= 1e-5
    # (3) the module's contract: a zero-filled workspace of its own, VQ_SEARCH_WS_CLEAN (no memset in front of the search);
    # the search kernel's last block must leave the counters zero again, so the second round has to agree as well
    ws3 = torch.zeros_like(ws)
    for _ in range(2):
        idx3 = torch.full_like(idx, -1)
        q3 = torch.full_like(q, float("nan"))
        sq3 = torch.empty_like(sq)
        cnt3 = torch.full_like(cnt, -7.0) if stats else None
        sums3 = torch.full_like(sums, float("nan")) if stats else None
        ops.assign(z, E, E_bf16, norms, search, idx3, None, ws3, presort=True, ws_clean=True)
        ops.gather_st_loss_scatter(z, idx3, E, q3, sq3, cnt3, sums3, stat_kind, deterministic, ws3, presorted=True)
        torch.cuda.synchronize()
        assert torch.equal(idx3, idx) and torch.equal(q3.view(torch.uint8), q.view(torch.uint8))
        assert abs(float(sq3) - float(sq)) <= 1e-5 * abs(float(sq)) or float(sq3) == float(sq)
        if stats:
            assert torch.equal(cnt3, cnt), "ws_clean round: per-code counts differ"
            fin = torch.isfinite(sums)
            scale = sums[fin].abs().max().clamp_min(1e-30) if bool(fin.any()) else 1.0
            assert float(((sums3 - sums)[fin].abs().max() if bool(fin.any()) else 0.0) / scale) <= 1e-5
    out = dict(idx=idx.cpu().long(), raw_idx=raw_idx.cpu().long(), q=q.cpu(), sqerr=sq.cpu()[0],
               norms=norms.cpu(), E_bf16=E_bf16.cpu()[:K * D].view(K, D))
    if want_smin:
        out["smin"] = smin.cpu()
    if stats:
        out["cnt"], out["sums"] = cnt.cpu(), sums.cpu()
    return out


def check_step_against_oracle(z_cpu, E_cpu, search, dtype=torch.float32, rel=1e-5, tol=None, **kw):
    """Full single-step parity vs the oracle.  Index rule (north_star): bit-exact except near-ties
    (|delta d| < 1e-5 relative), which are counted and returned."""
    prec = search if search in ("bf16", "fp16") else "fp32"     # "fp32" (split tensor-core tier) and "fp32_simt": fp32 oracle
    if tol is None:
        tol = 1e-3 if dtype in (torch.bfloat16, torch.float16) else 1e-5          # north_star tolerances
    zin = z_cpu.to(dtype)                                        # what the kernel actually receives
    out = cuda_step(z_cpu, E_cpu, search, dtype, **kw)
    ref_idx, ref_dmin = R.assign(zin.float(), E_cpu, prec)
    rep = R.near_tie_report(zin.float(), E_cpu, out["idx"], ref_idx, rel=rel, precision=prec)
    assert rep["hard"] == 0, f"index mismatches beyond near-ties: {rep}"
    K = E_cpu.shape[0]
    assert int(out["idx"].min()) >= 0 and int(out["idx"].max()) < K
    # everything downstream is checked against the oracle evaluated at the kernel's own indices, so that a
    # legitimate near-tie flip does not fail the gather / statistics comparison
    idx = out["idx"]
    q_ref = R.gather(E_cpu, idx, dtype)
    assert torch.equal(out["q"], q_ref), "gather must be bit-exact"
    sq_ref = float(R.sq_err_sum(zin.float(), E_cpu, idx))
    assert abs(float(out["sqerr"]) - sq_ref) <= tol * max(sq_ref, 1e-12), (float(out["sqerr"]), sq_ref)
    if "cnt" in out:
        cnt_ref, sums_ref = R.code_stats(zin.float(), idx, K)
        zmax = zin.float()[R.row_is_finite(zin)].abs().max() if bool(R.row_is_finite(zin).any()) else torch.tensor(1.0)
        assert torch.equal(out["cnt"], cnt_ref), "per-code counts must be exact"
        if kw.get("stat_kind", VQ_STAT_SUM_Z) == VQ_STAT_SUM_RESID:
            sums_ref = sums_ref - cnt_ref[:, None] * E_cpu
            scale = (zmax + E_cpu.abs().max()) * cnt_ref.clamp_min(1)[:, None]
        else:
            scale = zmax * cnt_ref.clamp_min(1)[:, None]
        err = ((out["sums"] - sums_ref).abs() / scale).max()
        assert float(err) <= 1e-5, f"per-code sums off by {float(err)} (relative to |z|max * count)"
    return rep, out

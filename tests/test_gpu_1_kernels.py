"""-m gpu parity tests, part 1: fp32 search (CUDA cores), fused gather pass, EMA update, backward.
Everything goes through the C ABI (vqvae_b200.ops -> libvq_b200.so).  Oracle = oracle/vq_reference.py
(self-written, PARITY UNPINNED -- SURVEY.md section 0)."""
import os

import numpy as np
import pytest
import torch

from oracle import vq_reference as R

pytestmark = pytest.mark.gpu


def _golden(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name)))


@pytest.mark.parametrize("name", ["hand_ties.npz", "rand_fp32_n4096_k64_d16.npz", "zipf_fp32_n8192_k256_d32.npz"])
def test_golden_fp32_step(golden_dir, name):
    from gpu_util import check_step_against_oracle
    g = _golden(golden_dir, name)
    z, E = torch.from_numpy(g["z"]), torch.from_numpy(g["E"])
    rep, out = check_step_against_oracle(z, E, "fp32")
    ref_idx = torch.from_numpy(g["idx"]).long()
    assert R.near_tie_report(z, E, out["idx"], ref_idx)["hard"] == 0
    if rep["mismatch"] == 0:
        np.testing.assert_array_equal(out["cnt"].numpy(), g["cnt"])
        np.testing.assert_allclose(float(out["sqerr"]), float(g["sqerr"]), rtol=1e-5)


def test_hand_case_exact(golden_dir):
    from gpu_util import cuda_step
    g = _golden(golden_dir, "hand_ties.npz")
    out = cuda_step(torch.from_numpy(g["z"]), torch.from_numpy(g["E"]), "fp32")
    assert out["idx"].tolist() == [1, 3, 0, 0, 0]           # exact ties -> lowest index, NaN row -> 0
    assert out["cnt"].tolist() == [2.0, 1.0, 0.0, 1.0]
    assert torch.equal(out["q"][4], torch.from_numpy(g["E"])[0])


@pytest.mark.parametrize("N,K,D", [(1, 1, 4), (63, 5, 4), (64, 64, 32), (65, 257, 36), (1000, 300, 100),
                                   (4097, 1024, 128), (20000, 512, 64), (3000, 8192, 256)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_fp32_search_shapes(N, K, D, dtype):
    from gpu_util import check_step_against_oracle
    g = torch.Generator().manual_seed(N * 7 + K)
    z, E = torch.randn(N, D, generator=g), torch.randn(K, D, generator=g)
    check_step_against_oracle(z, E, "fp32", dtype)


@pytest.mark.parametrize("deterministic", [False, True])
@pytest.mark.parametrize("dtype,D", [(torch.bfloat16, 256), (torch.bfloat16, 128), (torch.bfloat16, 64),
                                     (torch.bfloat16, 16), (torch.float32, 256), (torch.float32, 64),
                                     (torch.float32, 512), (torch.bfloat16, 1024), (torch.float32, 12)])
def test_pass_zipf_skew(dtype, D, deterministic):
    """Trained-like, heavily skewed assignments (Zipf 1.1): hot codes span many 64-row slices."""
    from gpu_util import check_step_against_oracle
    g = torch.Generator().manual_seed(D)
    K, N = 512, 30000
    E = torch.randn(K, D, generator=g)
    w = 1.0 / torch.arange(1, K + 1, dtype=torch.float64) ** 1.1
    c = torch.multinomial(w / w.sum(), N, replacement=True, generator=g)
    z = E[c] + 0.1 * torch.randn(N, D, generator=g)
    check_step_against_oracle(z, E, "fp32", dtype, deterministic=deterministic)


def test_pass_residual_statistics():
    from gpu_util import check_step_against_oracle
    from vqvae_b200._lib import VQ_STAT_SUM_RESID
    g = torch.Generator().manual_seed(5)
    E = torch.randn(100, 64, generator=g)
    z = E[torch.randint(0, 100, (5000,), generator=g)] + 0.05 * torch.randn(5000, 64, generator=g)
    check_step_against_oracle(z, E, "fp32", torch.float32, stat_kind=VQ_STAT_SUM_RESID)


def test_pass_all_rows_one_code_and_collapse():
    from gpu_util import check_step_against_oracle
    E = torch.zeros(64, 32)
    E[1:] = 100.0
    z = 0.01 * torch.randn(10000, 32, generator=torch.Generator().manual_seed(0))
    rep, out = check_step_against_oracle(z, E, "fp32")
    assert out["cnt"][0] == 10000 and out["cnt"][1:].sum() == 0


def test_deterministic_mode_is_bit_reproducible():
    from gpu_util import cuda_step
    g = torch.Generator().manual_seed(9)
    E = torch.randn(256, 64, generator=g)
    z = torch.randn(50000, 64, generator=g)
    a = cuda_step(z, E, "fp32", deterministic=True)
    b = cuda_step(z, E, "fp32", deterministic=True)
    assert torch.equal(a["sums"], b["sums"]) and torch.equal(a["sqerr"], b["sqerr"])


def test_nonfinite_rows():
    from gpu_util import check_step_against_oracle
    g = torch.Generator().manual_seed(2)
    E = torch.randn(32, 16, generator=g)
    z = torch.randn(500, 16, generator=g)
    z[3, 2] = float("nan"); z[77, 0] = float("inf"); z[499, 15] = float("-inf")
    for search, dt in (("fp32", torch.float32), ("fp32", torch.bfloat16), ("bf16", torch.bfloat16), ("bf16", torch.float32)):
        rep, out = check_step_against_oracle(z, E, search, dt)       # "bf16": the tcgen05 CTA-pair search sees NaN / inf rows
        assert out["idx"][3] == 0 and out["idx"][77] == 0 and out["idx"][499] == 0
        assert torch.isfinite(out["sqerr"])


def test_eval_pass_without_statistics():
    from gpu_util import check_step_against_oracle
    g = torch.Generator().manual_seed(4)
    check_step_against_oracle(torch.randn(3000, 128, generator=g), torch.randn(77, 128, generator=g), "fp32",
                              torch.bfloat16, stats=False)


def test_ema_update_kernel_vs_oracle():
    from vqvae_b200 import ops
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(21)
    K, D = 300, 48
    E = torch.randn(K, D, generator=g)
    n0 = torch.rand(K, generator=g) * 5
    m0 = torch.randn(K, D, generator=g)
    cnt = torch.randint(0, 4, (K,), generator=g).float()
    sums = torch.randn(K, D, generator=g) * cnt[:, None]
    En, n1, m1 = R.ema_update(E, n0, m0, cnt, sums, 0.97, 1e-5)
    u0 = torch.randint(0, 10, (K,), generator=g)
    s0 = torch.randint(0, 5, (K,), generator=g).int()
    u1, s1, nd = R.usage_update(u0, s0, cnt, 3)
    dE, dn, dm = E.to(dev), n0.to(dev), m0.to(dev)
    du, ds = u0.to(dev), s0.to(dev)
    ppl = torch.zeros(1, device=dev); ndead = torch.zeros(1, dtype=torch.int32, device=dev)
    Eb = ops.new_shadow(K, D, "bf16", dev)
    nrm = torch.empty(ops.norm_elems(K), device=dev)
    ws = torch.empty(ops.workspace_bytes(16, K, D, torch.float32), dtype=torch.uint8, device=dev)
    ops.ema_update(dE, dn, dm, cnt.to(dev), sums.to(dev), 0.97, 1e-5, True, "bf16", Eb, nrm, du, ds, ppl, ndead, 3,
                   True, ws)
    torch.cuda.synchronize()
    np.testing.assert_allclose(dE.cpu().numpy(), En.numpy(), rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(dn.cpu().numpy(), n1.numpy(), rtol=1e-6)
    np.testing.assert_allclose(dm.cpu().numpy(), m1.numpy(), rtol=1e-6, atol=1e-7)
    assert torch.equal(du.cpu(), u1) and torch.equal(ds.cpu(), s1) and int(ndead) == nd
    np.testing.assert_allclose(float(ppl), float(R.perplexity(cnt)), rtol=1e-5)
    assert torch.equal(Eb.cpu().view(K, D), dE.cpu().to(torch.bfloat16))
    np.testing.assert_allclose(nrm.cpu()[:K].numpy(), R.code_norms(dE.cpu(), "bf16").numpy(), rtol=1e-5)
    Kp = (K + 255) // 256 * 256                      # norms are padded to the 256-code search tile with +inf
    assert torch.isinf(nrm.cpu()[K:Kp]).all()
    # behind them: -||e||^2/2 split into three bf16 pieces per code, as an 8 KiB no-swizzle UMMA operand image per tile
    img = nrm.cpu()[Kp:].view(torch.bfloat16).view(Kp // 256, 32, 2, 8, 8).float()      # [tile, group, khalf, row, elem]
    h = img[:, :, 0, :, :3].sum(-1).reshape(-1)                                          # per code: h1 + h2 + h3
    np.testing.assert_allclose(h[:K].numpy(), -0.5 * nrm.cpu()[:K].numpy(), rtol=3e-7)
    assert torch.isinf(h[K:]).all() and bool((h[K:] < 0).all())
    assert float(img[:, :, 1].abs().max()) == 0.0 and float(img[:, :, 0, :, 3:].abs().max()) == 0.0


@pytest.mark.parametrize("quantizer", ["st", "ema"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_module_forward_backward_vs_oracle(quantizer, dtype):
    from vqvae_b200 import VectorQuantizer
    torch.manual_seed(0)
    K, D = 64, 32
    ref = R.VectorQuantizerOracle(D, K, beta=0.25, quantizer=quantizer, ema_decay=0.9)
    mod = VectorQuantizer(D, K, beta=0.25, quantizer=quantizer, ema_decay=0.9, search="fp32", deterministic=True)
    mod.load_state_dict(ref.state_dict(), strict=False)
    mod.last_perplexity.fill_(1.0)
    mod = mod.cuda()
    tol = 1e-3 if dtype == torch.bfloat16 else 1e-5
    g = torch.Generator().manual_seed(1)
    for step in range(3):
        z = torch.randn(4, 5, 7, D, generator=g).to(dtype)
        w = torch.randn(4, 5, 7, D, generator=g).to(dtype)
        zr = z.clone().requires_grad_(True)
        zc = z.cuda().requires_grad_(True)
        E_before = mod.codebook.detach().clone()
        qr, lr, ir = ref(zr)
        qc, lc, ic = mod(zc)
        assert qc.shape == z.shape and ic.shape == z.shape[:-1] and ic.dtype == torch.int64 and lc.dtype == torch.float32
        rep = R.near_tie_report(z.reshape(-1, D).float(), ref.codebook.detach(), ic.cpu().reshape(-1), ir.reshape(-1))
        assert rep["mismatch"] == 0, rep      # well separated random data: expect exact agreement
        # gather is bit-exact w.r.t. the module's own codebook; the two codebooks agree to tolerance after step 0
        assert torch.equal(qc.detach().reshape(-1, D), E_before[ic.reshape(-1)].to(dtype))
        np.testing.assert_allclose(qc.detach().float().cpu().numpy(), qr.detach().float().numpy(), rtol=max(tol, 1e-5), atol=1e-6)
        np.testing.assert_allclose(float(lc), float(lr), rtol=tol)
        ((qr.float() * w.float()).sum() + 2.0 * lr).backward()
        ((qc.float() * w.cuda().float()).sum() + 2.0 * lc).backward()
        np.testing.assert_allclose(zc.grad.float().cpu().numpy(), zr.grad.float().numpy(), rtol=max(tol, 1e-2 if dtype == torch.bfloat16 else tol), atol=1e-6)
        if quantizer == "st":
            np.testing.assert_allclose(mod.codebook.grad.cpu().numpy(), ref.codebook.grad.numpy(), rtol=1e-4, atol=1e-7)
            with torch.no_grad():       # a plain SGD step on both, exercises the shadow refresh
                ref.codebook -= 0.1 * ref.codebook.grad; mod.codebook -= 0.1 * mod.codebook.grad
            ref.codebook.grad = None; mod.codebook.grad = None
        np.testing.assert_allclose(mod.codebook.detach().cpu().numpy(), ref.codebook.detach().numpy(), rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(float(mod.last_perplexity), float(ref.last_perplexity), rtol=1e-5)
    assert torch.equal(mod.usage_total.cpu(), ref.usage_total)
    assert torch.equal(mod.steps_since_used.cpu(), ref.steps_since_used)


def test_module_golden_ema_trajectory(golden_dir):
    from vqvae_b200 import VectorQuantizer
    g = _golden(golden_dir, "ema_traj_k32_d8.npz")
    K, D = g["E0"].shape
    mod = VectorQuantizer(D, K, beta=float(g["beta"]), quantizer="ema", ema_decay=float(g["decay"]),
                          ema_eps=float(g["eps"]), dead_after=int(g["dead_after"]), search="fp32", deterministic=True)
    with torch.no_grad():
        mod.codebook.copy_(torch.from_numpy(g["E0"])); mod.ema_sum.copy_(torch.from_numpy(g["E0"]))
    mod = mod.cuda()
    for t in range(g["z"].shape[0]):
        q, loss, idx = mod(torch.from_numpy(g["z"][t]).cuda())
        assert np.array_equal(idx.cpu().numpy(), g["idx"][t]), f"step {t}"
        np.testing.assert_allclose(float(loss), g["loss"][t], rtol=1e-5)
        np.testing.assert_allclose(mod.codebook.cpu().numpy(), g["codebooks"][t], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(float(mod.last_perplexity), g["perplexity"][t], rtol=1e-5)
    np.testing.assert_array_equal(mod.usage_total.cpu().numpy(), g["usage_total"])
    np.testing.assert_array_equal(mod.steps_since_used.cpu().numpy(), g["steps_since_used"])
    assert int(mod.n_dead) == int(g["n_dead"])


def test_module_eval_mode_and_empty_and_state_dict_roundtrip(tmp_path):
    from vqvae_b200 import VectorQuantizer
    torch.manual_seed(3)
    mod = VectorQuantizer(16, 40, quantizer="ema", search="fp32").cuda()
    mod(torch.randn(100, 16, device="cuda"))
    E1 = mod.codebook.clone()
    mod.eval()
    mod(torch.randn(100, 16, device="cuda"))
    assert torch.equal(mod.codebook, E1) and int(mod.usage_total.sum()) == 100
    q, loss, idx = mod(torch.zeros(0, 16, device="cuda"))
    assert q.shape == (0, 16) and idx.shape == (0,) and float(loss) == 0.0
    torch.save({"model": mod.state_dict()}, tmp_path / "ck.pt")
    sd = torch.load(tmp_path / "ck.pt", map_location="cpu")["model"]       # scripts/export_codebook.py path
    cpu = VectorQuantizer(16, 40, quantizer="ema")
    cpu.load_state_dict(sd)
    assert torch.equal(cpu.codebook, E1.cpu()) and cpu.codebook.shape[0] == 40


@pytest.mark.gpu
@pytest.mark.parametrize("shape,dtype", [((3, 64, 17, 19), torch.float32), ((2, 128, 5, 32, 32), torch.bfloat16),
                                         ((4, 7, 9), torch.float32), ((1, 256, 130), torch.bfloat16)])
def test_layout_transpose_matches_permute(shape, dtype):
    """a-0: [B, R, C] -> [B, C, R] kernel, bit-exact against torch.permute (odd sizes take the scalar path)."""
    from vqvae_b200 import ops
    x = torch.randn(shape).to(dtype).cuda().reshape(shape[0], shape[1], -1)
    out = torch.empty((x.shape[0], x.shape[2], x.shape[1]), dtype=dtype, device="cuda")
    ops.layout_transpose(x, out)
    assert torch.equal(out, x.permute(0, 2, 1).contiguous())


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_channel_first_module_matches_channel_last(dtype):
    """The live model's [B, D, H, W] latents through channel_first=True == permuting by hand (values and grads)."""
    from vqvae_b200 import VectorQuantizer
    torch.manual_seed(3)
    D, K = 64, 128
    a = VectorQuantizer(D, K, quantizer="ema", channel_first=True, deterministic=True).cuda()
    b = VectorQuantizer(D, K, quantizer="ema", deterministic=True).cuda()
    b.load_state_dict(a.state_dict())
    z = torch.randn(2, D, 24, 20).to(dtype).cuda()
    za = z.clone().requires_grad_(True)
    zb = z.permute(0, 2, 3, 1).contiguous().requires_grad_(True)
    qa, la, ia = a(za)
    qb, lb, ib = b(zb)
    assert qa.shape == z.shape and ia.shape == (2, 24, 20)
    assert torch.equal(ia, ib) and torch.equal(qa, qb.permute(0, 3, 1, 2)) and float(la) == float(lb)
    w = torch.randn_like(qa)
    (qa * w).sum().backward(retain_graph=True); la.backward()
    (qb * w.permute(0, 2, 3, 1)).sum().backward(retain_graph=True); lb.backward()
    assert torch.equal(za.grad, zb.grad.permute(0, 3, 1, 2))
    assert torch.equal(a.codebook, b.codebook)

"""-m gpu parity tests, part 2: the tcgen05 / TMEM / TMA nearest-code search (VQ_SEARCH_BF16) and the
module on top of it.  Oracle = oracle/vq_reference.py in "bf16" mode: z and E rounded to bf16, fp32
accumulate -- bit-comparable with the tensor-core GEMM up to fp32 summation order, so index flips are
allowed only at near-ties (|delta d| < 1e-5 relative) and are counted (north_star)."""
import os

import numpy as np
import pytest
import torch

from oracle import vq_reference as R

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,K,D", [
    (128, 256, 64), (1, 1, 8), (127, 100, 16), (129, 255, 32), (300, 257, 64), (5000, 1000, 128),
    (4096, 1024, 128), (2000, 512, 72), (1500, 3000, 256), (777, 513, 512), (148 * 128 * 2 + 77, 512, 64),
    (3000, 8192, 256), (2048, 2048, 48), (1000, 2048, 8),      # last two: the live model's z_type_dim / z_phase_dim (frl_repr_model_v1.yaml:14-15)
    (6000, 20000, 64),      # K beyond the shared-memory histogram and norm staging: global-counter sort, norms from L2
])
def test_tc_search_shapes(N, K, D):
    from gpu_util import check_step_against_oracle
    g = torch.Generator().manual_seed(N + 13 * K + D)
    z, E = torch.randn(N, D, generator=g), torch.randn(K, D, generator=g)
    rep, out = check_step_against_oracle(z, E, "bf16", torch.bfloat16)
    zin = z.to(torch.bfloat16).float()
    s = R.scores(zin, E, "bf16")
    ref_smin = s.gather(1, out["idx"][:, None])[:, 0]
    scale = (zin * zin).sum(1) + ref_smin.abs()
    assert float(((out["smin"] - ref_smin).abs() / scale.clamp_min(1e-6)).max()) < 1e-5
    assert torch.equal(out["E_bf16"], E.to(torch.bfloat16))


def test_tc_golden_bf16(golden_dir):
    from gpu_util import check_step_against_oracle
    g = dict(np.load(os.path.join(golden_dir, "rand_bf16_n2048_k128_d64.npz")))
    z, E = torch.from_numpy(g["z"]), torch.from_numpy(g["E"])
    rep, out = check_step_against_oracle(z, E, "bf16", torch.bfloat16)
    assert R.near_tie_report(z, E, out["idx"], torch.from_numpy(g["idx"]).long(), precision="bf16")["hard"] == 0
    if rep["mismatch"] == 0:
        np.testing.assert_array_equal(out["cnt"].numpy(), g["cnt"])


def test_tc_exact_ties_lowest_index():
    from gpu_util import cuda_step
    g = torch.Generator().manual_seed(3)
    K, D = 700, 64
    E = torch.randn(K, D, generator=g).to(torch.bfloat16).float()
    E[300] = E[5]; E[650] = E[5]; E[699] = E[256]       # duplicates across chains, halves and tiles
    z = torch.cat([E[5][None].repeat(200, 1), E[256][None].repeat(200, 1)]).to(torch.bfloat16).float()
    out = cuda_step(z, E, "bf16", torch.bfloat16)
    assert out["raw_idx"][:200].eq(5).all() and out["raw_idx"][200:].eq(256).all()


def test_tc_fp32_latents_staged_as_bf16():
    from gpu_util import check_step_against_oracle
    g = torch.Generator().manual_seed(8)
    z, E = torch.randn(3000, 128, generator=g), torch.randn(400, 128, generator=g)
    # kernel sees fp32 z (gather/loss in fp32) but searches on its bf16 rounding
    from gpu_util import cuda_step
    out = cuda_step(z, E, "bf16", torch.float32)
    ref_idx, _ = R.assign(z, E, "bf16")
    assert R.near_tie_report(z, E, out["idx"], ref_idx, precision="bf16")["hard"] == 0
    assert torch.equal(out["q"], E[out["idx"]])


def test_tc_trained_like_zipf():
    from gpu_util import check_step_against_oracle
    g = torch.Generator().manual_seed(17)
    K, D, N = 1024, 128, 40000
    E = torch.randn(K, D, generator=g)
    w = 1.0 / torch.arange(1, K + 1, dtype=torch.float64) ** 1.1
    c = torch.multinomial(w / w.sum(), N, replacement=True, generator=g)
    z = E[c] + 0.1 * torch.randn(N, D, generator=g)
    rep, out = check_step_against_oracle(z, E, "bf16", torch.bfloat16)
    assert float((out["idx"] == c).float().mean()) > 0.99


def test_tc_unsupported_shape_is_loud():
    from vqvae_b200 import ops
    from vqvae_b200._lib import VQLibError
    dev = torch.device("cuda:0")
    z = torch.randn(10, 12, device=dev).to(torch.bfloat16)       # D % 8 != 0
    E = torch.randn(4, 12, device=dev)
    Eb = ops.new_shadow(4, 12, "bf16", dev)
    nrm = torch.empty(ops.norm_elems(4), device=dev)
    ops.prepare_codebook(E, "bf16", Eb, nrm)
    with pytest.raises(VQLibError, match="not supported"):
        ops.assign(z, E, Eb, nrm, "bf16", torch.empty(10, dtype=torch.int32, device=dev), None, None)


@pytest.mark.parametrize("quantizer", ["st", "ema"])
def test_module_bf16_autocast_training_steps(quantizer):
    """Called the way scripts/train_vqvae.py:284-287 calls it: under bf16 autocast, loss scaled by a
    lambda, backward through the straight-through output."""
    from vqvae_b200 import VectorQuantizer
    torch.manual_seed(0)
    K, D = 256, 128
    ref = R.VectorQuantizerOracle(D, K, beta=0.25, quantizer=quantizer, ema_decay=0.95, precision="bf16")
    mod = VectorQuantizer(D, K, beta=0.25, quantizer=quantizer, ema_decay=0.95, deterministic=True)
    mod.load_state_dict(ref.state_dict())
    mod = mod.cuda()
    g = torch.Generator().manual_seed(1)
    total_near = 0
    for step in range(4):
        z = torch.randn(8, 5, 16, D, generator=g).to(torch.bfloat16)
        zr = z.clone().requires_grad_(True)
        zc = z.cuda().requires_grad_(True)
        qr, lr, ir = ref(zr)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            qc, lc, ic = mod(zc)
        assert qc.dtype == torch.bfloat16 and lc.dtype == torch.float32
        rep = R.near_tie_report(z.reshape(-1, D).float(), ref.codebook.detach(), ic.cpu().reshape(-1), ir.reshape(-1),
                                precision="bf16")
        assert rep["hard"] == 0, rep
        total_near += rep["near_tie"]
        if rep["mismatch"]:
            pytest.skip(f"near-tie flip at step {step} ({rep}); trajectories legitimately diverge")
        np.testing.assert_allclose(float(lc), float(lr), rtol=1e-3)
        (qr.float().sum() + 0.5 * lr).backward()
        (qc.float().sum() + 0.5 * lc).backward()
        np.testing.assert_allclose(zc.grad.float().cpu().numpy(), zr.grad.float().numpy(), rtol=1e-2, atol=1e-6)
        np.testing.assert_allclose(mod.codebook.detach().cpu().numpy(), ref.codebook.detach().numpy(), rtol=1e-3, atol=1e-6)
        np.testing.assert_allclose(float(mod.last_perplexity), float(ref.last_perplexity), rtol=1e-4)
    assert torch.equal(mod.usage_total.cpu(), ref.usage_total)


def test_full_size_properties_c2():
    """BASELINE config C2 at full size (N = 256*5*32*32, K=1024, D=128, bf16): size-independent properties."""
    from vqvae_b200 import VectorQuantizer
    torch.manual_seed(0)
    N, K, D = 256 * 5 * 32 * 32, 1024, 128
    mod = VectorQuantizer(D, K, quantizer="ema").cuda()
    E0 = mod.codebook.clone()
    z = torch.randn(N, D, device="cuda", dtype=torch.bfloat16)
    q, loss, idx = mod(z)
    assert int(idx.min()) >= 0 and int(idx.max()) < K
    assert torch.equal(q, E0[idx].to(torch.bfloat16))                             # gather is exact
    assert int(mod.usage_total.sum()) == N                                        # counts conserve rows
    mse = ((z.float() - q.float()) ** 2).mean()
    assert abs(float(loss) / 0.25 - float(((z.float() - E0[idx]) ** 2).mean())) < 1e-3 * float(mse)
    # idempotence: quantising codebook rows returns them unchanged with their own index
    mod.eval()
    Eb = mod.codebook.to(torch.bfloat16)
    q2, l2, i2 = mod(Eb)
    d_self = ((Eb.float() - mod.codebook[i2].to(torch.bfloat16).float()) ** 2).sum(1)
    assert float(d_self.max()) < 1e-6
    # a sample of rows against the oracle
    sel = torch.randperm(N)[:4096]
    ref_idx, _ = R.assign(z[sel.cuda()].cpu().float(), E0.cpu(), "bf16")
    rep = R.near_tie_report(z[sel.cuda()].cpu().float(), E0.cpu(), idx[sel.cuda()].cpu(), ref_idx, precision="bf16")
    assert rep["hard"] == 0, rep


@pytest.mark.gpu
def test_full_size_properties_c4():
    """BASELINE config C4 per-GPU shard at full size (N = 524288, K = 8192, D = 256, bf16)."""
    from vqvae_b200 import VectorQuantizer
    torch.manual_seed(1)
    N, K, D = 524288, 8192, 256
    mod = VectorQuantizer(D, K, quantizer="ema").cuda()
    E0 = mod.codebook.clone()
    ema_sum0, ema_cnt0 = mod.ema_sum.clone(), mod.ema_cnt.clone()
    z = torch.randn(N, D, device="cuda", dtype=torch.bfloat16)
    q, loss, idx = mod(z)
    assert int(idx.min()) >= 0 and int(idx.max()) < K
    assert torch.equal(q, E0[idx].to(torch.bfloat16))
    assert int(mod.usage_total.sum()) == N
    assert torch.equal(mod.usage_total, torch.bincount(idx, minlength=K))           # counts are exact
    # linearity of the EMA statistics: sum over codes of sum_k == column sums of z (checksum of checksums)
    code_sum = (mod.ema_sum - 0.99 * ema_sum0) / (1.0 - 0.99)
    np.testing.assert_allclose(code_sum.sum(0).cpu().numpy(), z.float().sum(0).cpu().numpy(), rtol=2e-3, atol=2.0)
    code_cnt = (mod.ema_cnt - 0.99 * ema_cnt0) / (1.0 - 0.99)
    np.testing.assert_allclose(code_cnt.cpu().numpy(), mod.usage_total.float().cpu().numpy(), rtol=1e-3, atol=0.05)
    mse = float(((z.float() - E0[idx]) ** 2).mean())
    assert abs(float(loss) / 0.25 - mse) < 1e-4 * mse
    sel = torch.randperm(N)[:2048]
    zs = z[sel.cuda()].cpu().float()
    ref_idx, _ = R.assign(zs, E0.cpu(), "bf16")
    rep = R.near_tie_report(zs, E0.cpu(), idx[sel.cuda()].cpu(), ref_idx, precision="bf16")
    assert rep["hard"] == 0, rep


@pytest.mark.gpu
@pytest.mark.parametrize("K,D", [(256, 128), (256, 64), (512, 64)])
def test_reference_default_sizes_bf16(K, D):
    """The sizes the reference itself writes down (scripts/config.yaml:48-49, frl/config/frl_model_v0.yaml:32-33,
    scripts/train_vqvae.py:410-411): epilogue-bound shapes, full single-step parity vs the oracle."""
    g = torch.Generator().manual_seed(K + D)
    E = torch.randn(K, D, generator=g)
    z = E[torch.randint(0, K, (20000,), generator=g)] + 0.3 * torch.randn(20000, D, generator=g)
    from gpu_util import check_step_against_oracle
    rep, _ = check_step_against_oracle(z, E, "bf16", dtype=torch.bfloat16)
    assert rep["hard"] == 0, rep


@pytest.mark.gpu
@pytest.mark.parametrize("search", ["auto", "fp32"])
def test_fp16_autocast_input(search):
    """scripts/train_vqvae.py:284-286 runs the model under fp16 autocast unless --bf16.  fp16 rows go to the kernels as
    they are: "auto" = tensor-core search on fp16 operands (fp16-rounded codebook), search="fp32" keeps the exact tier."""
    from vqvae_b200 import VectorQuantizer
    torch.manual_seed(5)
    mod = VectorQuantizer(32, 64, quantizer="st", search=search).cuda()
    with torch.no_grad():
        mod.codebook.normal_()
    z = torch.randn(4, 1000, 32, device="cuda", dtype=torch.float16, requires_grad=True)
    with torch.autocast("cuda", dtype=torch.float16):
        q, loss, idx = mod(z)
    assert q.dtype == torch.float16 and loss.dtype == torch.float32 and idx.shape == (4, 1000)
    zf, E = z.detach().float().cpu().reshape(-1, 32), mod.codebook.detach().cpu()
    prec = "fp32" if search == "fp32" else "fp16"
    ref_idx, _ = R.assign(zf, E, prec)
    rep = R.near_tie_report(zf, E, idx.cpu().reshape(-1), ref_idx, precision=prec)
    assert rep["hard"] == 0, rep
    assert torch.equal(q.detach().float().cpu().reshape(-1, 32), E[idx.cpu().reshape(-1)].to(torch.float16).float())
    mse = float(((zf - E[idx.cpu().reshape(-1)]) ** 2).mean())
    np.testing.assert_allclose(float(loss), 1.25 * mse, rtol=1e-4)          # loss on the exact fp32 values
    (q.float().sum() + loss).backward()
    assert torch.isfinite(z.grad).all() and mod.codebook.grad is not None

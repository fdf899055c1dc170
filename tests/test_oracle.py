"""CPU tests of the self-written oracle (oracle/vq_reference.py) -- PARITY UNPINNED: the reference
tree holds no quantiser, tests or fixtures for one, so these pin the oracle against
(a) the committed known-answer vectors, (b) an independent numpy fp64 brute force, (c) properties.
Style follows the reference's own tests (frl/tests/test_soft_neighborhood.py: seeded tiny tensors,
property checks, gradient-flow checks)."""
import os

import numpy as np
import pytest
import torch

from oracle import vq_reference as R


def _load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name)))


def _brute(z, E):
    """independent fp64 restatement: explicit ||z - e||^2, first minimum."""
    d = ((z[:, None, :].astype(np.float64) - E[None, :, :].astype(np.float64)) ** 2).sum(-1)
    return d.argmin(1), d


# ---------------------------------------------------------------- golden vectors
@pytest.mark.parametrize("name,prec,quant", [
    ("hand_ties.npz", "fp32", "st"),
    ("rand_fp32_n4096_k64_d16.npz", "fp32", "st"),
    ("zipf_fp32_n8192_k256_d32.npz", "fp32", "ema"),
    ("rand_bf16_n2048_k128_d64.npz", "bf16", "ema"),
])
def test_golden_single_step(golden_dir, name, prec, quant):
    g = _load(golden_dir, name)
    z, E = torch.from_numpy(g["z"]), torch.from_numpy(g["E"])
    K = E.shape[0]
    idx, dmin = R.assign(z, E, prec)
    assert np.array_equal(idx.numpy(), g["idx"])
    np.testing.assert_allclose(dmin.numpy(), g["dmin"], rtol=1e-6, atol=1e-6, equal_nan=True)
    cnt, sums = R.code_stats(z, idx, K)
    np.testing.assert_array_equal(cnt.numpy(), g["cnt"])
    np.testing.assert_allclose(sums.numpy(), g["sums"], rtol=1e-6, atol=1e-6)
    sq = R.sq_err_sum(z, E, idx)
    np.testing.assert_allclose(float(sq), float(g["sqerr"]), rtol=1e-9)
    loss = R.vq_loss_value(sq, z.shape[0], z.shape[1], float(g["beta"]), quant)
    np.testing.assert_allclose(float(loss), float(g["loss"]), rtol=1e-6)
    np.testing.assert_allclose(float(R.perplexity(cnt)), float(g["perplexity"]), rtol=1e-6)


def test_hand_case_by_hand(golden_dir):
    g = _load(golden_dir, "hand_ties.npz")
    assert g["idx"].tolist() == [1, 3, 0, 0, 0]          # ties -> lowest index; NaN row -> 0
    assert g["cnt"].tolist() == [2.0, 1.0, 0.0, 1.0]     # NaN row excluded from the statistics
    # row 0: (0.9-1)^2 + 0.1^2 = 0.02 ; row 1: 0.1^2+0.1^2 = 0.02 ; row 2: 1 ; row 3: 0.04+0.01+0.01 = 0.06
    np.testing.assert_allclose(float(g["sqerr"]), 0.02 + 0.02 + 1.0 + 0.06, rtol=1e-6)
    np.testing.assert_allclose(float(g["loss"]), 1.25 * 1.10 / 15.0, rtol=1e-6)   # (1+beta) * sum / (N*D), N=5
    p = np.array([0.5, 0.25, 0.25])
    np.testing.assert_allclose(float(g["perplexity"]), np.exp(-(p * np.log(p)).sum()), rtol=1e-5)


def test_golden_ema_trajectory(golden_dir):
    g = _load(golden_dir, "ema_traj_k32_d8.npz")
    K, D = g["E0"].shape
    m = R.VectorQuantizerOracle(D, K, beta=float(g["beta"]), quantizer="ema", ema_decay=float(g["decay"]),
                                ema_eps=float(g["eps"]), dead_after=int(g["dead_after"]))
    with torch.no_grad():
        m.codebook.copy_(torch.from_numpy(g["E0"]))
        m.ema_sum.copy_(torch.from_numpy(g["E0"]))
    for t in range(g["z"].shape[0]):
        q, loss, idx = m(torch.from_numpy(g["z"][t]))
        assert np.array_equal(idx.numpy(), g["idx"][t])
        np.testing.assert_allclose(float(loss), g["loss"][t], rtol=1e-6)
        np.testing.assert_allclose(m.codebook.numpy(), g["codebooks"][t], rtol=1e-6, atol=1e-7)
        np.testing.assert_allclose(float(m.last_perplexity), g["perplexity"][t], rtol=1e-6)
    np.testing.assert_array_equal(m.usage_total.numpy(), g["usage_total"])
    np.testing.assert_array_equal(m.steps_since_used.numpy(), g["steps_since_used"])
    assert int(m.n_dead) == int(g["n_dead"])
    assert int(m.usage_total.sum()) == g["z"].shape[0] * g["z"].shape[1]


# ---------------------------------------------------------------- against an independent brute force
@pytest.mark.parametrize("N,K,D,seed", [(257, 17, 5, 0), (1000, 64, 16, 1), (64, 300, 33, 2)])
def test_assign_matches_bruteforce(N, K, D, seed):
    rng = np.random.default_rng(seed)
    z = rng.standard_normal((N, D)).astype(np.float32)
    E = rng.standard_normal((K, D)).astype(np.float32)
    bi, d = _brute(z, E)
    idx, dmin = R.assign(torch.from_numpy(z), torch.from_numpy(E), "fp32")
    rep = R.near_tie_report(torch.from_numpy(z), torch.from_numpy(E), idx, torch.from_numpy(bi))
    assert rep["hard"] == 0, rep
    np.testing.assert_allclose(dmin.numpy(), d.min(1), rtol=1e-4, atol=1e-4)


def test_exact_duplicate_codes_lowest_index():
    torch.manual_seed(0)
    E = torch.randn(8, 4)
    E[5] = E[2]
    z = E[5][None, :].repeat(3, 1) + 1e-3
    idx, _ = R.assign(z, E, "fp32")
    assert idx.tolist() == [2, 2, 2]


def test_bf16_mode_rounds_inputs():
    torch.manual_seed(1)
    z, E = torch.randn(512, 32), torch.randn(40, 32)
    i1, _ = R.assign(z, E, "bf16")
    i2, _ = R.assign(R.round_bf16(z), R.round_bf16(E), "fp32")
    assert R.near_tie_report(R.round_bf16(z), R.round_bf16(E), i1, i2)["hard"] == 0


# ---------------------------------------------------------------- properties
def test_identical_inputs_zero_loss():
    m = R.VectorQuantizerOracle(4, 6, quantizer="st")
    z = m.codebook.detach()[[1, 3, 3, 0]].clone()
    q, loss, idx = m(z)
    assert idx.tolist() == [1, 3, 3, 0]
    assert float(loss) == 0.0
    assert torch.equal(q, z)


def test_gradients_flow_st_and_commitment():
    torch.manual_seed(0)
    m = R.VectorQuantizerOracle(8, 16, beta=0.25, quantizer="st")
    z = torch.randn(10, 8, requires_grad=True)
    q, loss, idx = m(z)
    w = torch.randn_like(q)
    ((q * w).sum() + 3.0 * loss).backward()
    exp_z = R.backward_z(z.detach(), q.detach(), w, 3.0, 0.25)
    np.testing.assert_allclose(z.grad.numpy(), exp_z.numpy(), rtol=1e-5, atol=1e-6)
    cnt, sums = R.code_stats(z.detach(), idx, 16)
    exp_E = R.backward_codebook(m.codebook.detach(), cnt, sums, 3.0, 10, 8)
    np.testing.assert_allclose(m.codebook.grad.numpy(), exp_E.numpy(), rtol=1e-4, atol=1e-6)


def test_ema_mode_codebook_is_buffer_and_st_is_parameter():
    st = R.VectorQuantizerOracle(4, 8, quantizer="st")
    ema = R.VectorQuantizerOracle(4, 8, quantizer="ema")
    assert [n for n, _ in st.named_parameters()] == ["codebook"]
    assert list(ema.named_parameters()) == []
    assert "codebook" in ema.state_dict() and "usage_total" in ema.state_dict()


def test_ema_update_fixed_point():
    """If every code receives exactly its own row once with decay g, m/n stays e (up to Laplace eps)."""
    K, D = 8, 4
    torch.manual_seed(3)
    E = torch.randn(K, D)
    En, n, m = R.ema_update(E, torch.ones(K), E.clone(), torch.ones(K), E.clone(), 0.9, 0.0)
    np.testing.assert_allclose(En.numpy(), E.numpy(), rtol=1e-6)
    np.testing.assert_allclose(n.numpy(), np.ones(K), rtol=1e-6)


def test_eval_mode_does_not_update():
    torch.manual_seed(0)
    m = R.VectorQuantizerOracle(4, 8, quantizer="ema").eval()
    E0 = m.codebook.clone()
    m(torch.randn(32, 4))
    assert torch.equal(m.codebook, E0) and int(m.usage_total.sum()) == 0


def test_empty_input():
    m = R.VectorQuantizerOracle(4, 8, quantizer="ema")
    q, loss, idx = m(torch.zeros(0, 4))
    assert q.shape == (0, 4) and idx.shape == (0,) and float(loss) == 0.0


def test_flatten_channel_first_matches_permute():
    z = torch.arange(2 * 3 * 4 * 5, dtype=torch.float32).reshape(2, 3, 4, 5)    # [B, D, H, W]
    a = R.flatten_latents(z, channel_first=True)
    b = z.permute(0, 2, 3, 1).reshape(-1, 3)
    assert torch.equal(a, b)

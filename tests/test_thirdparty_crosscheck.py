"""Cross-check of the (unpinned) quantiser oracle and of the CUDA path against an independent public implementation:
golden vectors from HF transformers' ChameleonVQVAEVectorQuantizer (tests/golden/make_thirdparty_golden.py).
The reference has no quantiser of its own to compare with; this is a second opinion on the oracle's
choices (distance form, first-minimum ties, gather, straight-through, loss value), not a pin."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import vq_reference as R

FILES = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "thirdparty_chameleon_*.npz")))
IDS = [os.path.basename(p)[:-4] for p in FILES]


def _check(g, quant, loss, idx, grad_z, grad_E, tol):
    B, D, H, W = g["z"].shape
    beta = float(g["beta"])
    assert tuple(quant.shape) == (B, D, H, W) and tuple(idx.shape) == (B, H, W)
    assert np.array_equal(idx.reshape(-1), g["indices"])                       # bit-exact indices
    np.testing.assert_allclose(quant, g["quantized"], rtol=tol, atol=1e-6)
    np.testing.assert_allclose(float(loss), float(g["loss"]), rtol=max(tol, 2e-6))
    # beta sits on the other term there (see the generator's docstring): the loss-driven parts differ by exactly beta
    np.testing.assert_allclose(grad_z - g["w"], beta * (g["grad_z"] - g["w"]), rtol=1e-4, atol=1e-6)   # fp32 cancellation against w ~ 1
    np.testing.assert_allclose(beta * grad_E, g["grad_E"], rtol=1e-4, atol=1e-8)


def test_fixtures_present():
    assert len(FILES) >= 3


@pytest.mark.parametrize("path", FILES, ids=IDS)
def test_oracle_agrees_with_public_implementation(path):
    g = np.load(path)
    K, D = g["E"].shape
    ref = R.VectorQuantizerOracle(D, K, beta=float(g["beta"]), quantizer="st")
    with torch.no_grad():
        ref.codebook.copy_(torch.from_numpy(g["E"]))
    z = torch.from_numpy(g["z"]).permute(0, 2, 3, 1).contiguous().requires_grad_(True)     # oracle takes channel-last
    q, loss, idx = ref(z)
    ((q * torch.from_numpy(g["w"]).permute(0, 2, 3, 1)).sum() + 3.0 * loss).backward()
    _check(g, q.detach().permute(0, 3, 1, 2).numpy(), loss.detach(), idx.numpy(), z.grad.permute(0, 3, 1, 2).numpy(),
           ref.codebook.grad.numpy(), 1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("path", FILES, ids=IDS)
def test_cuda_agrees_with_public_implementation(path):
    """Same call shape as the public module: channel-first [B, D, H, W] in, (quantized, loss, indices) out."""
    from vqvae_b200 import VectorQuantizer
    g = np.load(path)
    K, D = g["E"].shape
    mod = VectorQuantizer(D, K, beta=float(g["beta"]), quantizer="st", search="fp32", deterministic=True,
                          channel_first=True).cuda()
    with torch.no_grad():
        mod.codebook.copy_(torch.from_numpy(g["E"]))
    z = torch.from_numpy(g["z"]).cuda().requires_grad_(True)
    q, loss, idx = mod(z)
    ((q * torch.from_numpy(g["w"]).cuda()).sum() + 3.0 * loss).backward()
    _check(g, q.detach().cpu().numpy(), loss.detach().cpu(), idx.cpu().numpy(), z.grad.cpu().numpy(),
           mod.codebook.grad.cpu().numpy(), 1e-5)


@pytest.mark.gpu
def test_cuda_bf16_search_agrees_with_public_implementation_up_to_near_ties():
    from vqvae_b200 import VectorQuantizer
    g = np.load(FILES[-1])
    K, D = g["E"].shape
    mod = VectorQuantizer(D, K, beta=float(g["beta"]), quantizer="st", channel_first=True).cuda()      # tensor-core search
    with torch.no_grad():
        mod.codebook.copy_(torch.from_numpy(g["E"]))
    z = torch.from_numpy(g["z"]).cuda()
    _, _, idx = mod(z)
    zf = torch.from_numpy(g["z"]).permute(0, 2, 3, 1).reshape(-1, D)
    rep = R.near_tie_report(zf, torch.from_numpy(g["E"]), idx.cpu().reshape(-1), torch.from_numpy(g["indices"]), rel=2e-2)
    assert rep["hard"] == 0, rep        # bf16 rounding of z and E moves only near-tied rows
    assert rep["mismatch"] <= 0.05 * zf.shape[0], rep


# ---------------------------------------------------------------------------------------------------------------------
# EMA half (a-5, a-7): trajectories of an independent published EMA codebook with a training-mode update -- vllm 0.22
# `EuclideanCodebook` (the EnCodec / lucidrains form) -- frozen by tests/golden/make_thirdparty_ema_golden.py.
# Confirms: search against the pre-update codebook, n <- g n + (1-g) cnt, m <- g m + (1-g) sum z, Laplace smoothing
# n^ = (n + eps) / (sum n + K eps) * sum n, e = m / n^.  Does not pin: initial n (the fixtures start from OUR n = 1),
# the dead-code rule, perplexity.  Still not the reference: the quantiser rows stay "parity unpinned".
EMA_FILES = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "thirdparty_ema_*.npz")))
EMA_IDS = [os.path.basename(p)[:-4] for p in EMA_FILES]


def test_ema_fixtures_present():
    assert len(EMA_FILES) >= 3


@pytest.mark.parametrize("path", EMA_FILES, ids=EMA_IDS)
def test_oracle_ema_trajectory_agrees_with_public_implementation(path):
    g = np.load(path)
    S, N, D = g["z"].shape
    K = g["E0"].shape[0]
    ref = R.VectorQuantizerOracle(D, K, quantizer="ema", ema_decay=float(g["decay"]), ema_eps=float(g["eps"])).train()
    with torch.no_grad():
        ref.codebook.copy_(torch.from_numpy(g["E0"]))
        ref.ema_sum.copy_(torch.from_numpy(g["E0"]))
        ref.ema_cnt.fill_(1.0)
    for s in range(S):
        z = torch.from_numpy(g["z"][s])
        _, _, idx = ref(z)
        assert np.array_equal(idx.numpy(), g["indices"][s]), f"step {s}: indices differ"
        np.testing.assert_allclose(ref.ema_cnt.numpy(), g["cluster_size"][s], rtol=2e-6, atol=1e-7)
        np.testing.assert_allclose(ref.ema_sum.numpy(), g["embed_avg"][s], rtol=1e-5, atol=2e-6)
        np.testing.assert_allclose(ref.codebook.numpy(), g["E"][s], rtol=1e-5, atol=2e-6)


def test_oracle_ema_update_function_is_bitwise_the_public_form():
    """The update formula alone, fed with THEIR statistics: Laplace smoothing and the division agree to the last bits."""
    g = np.load(EMA_FILES[0])
    K, D = g["E0"].shape
    E, n, m = torch.from_numpy(g["E0"]), torch.ones(K), torch.from_numpy(g["E0"]).clone()
    for s in range(g["z"].shape[0]):
        z, idx = torch.from_numpy(g["z"][s]), torch.from_numpy(g["indices"][s])
        cnt, sums = R.code_stats(z, idx, K)
        E, n, m = R.ema_update(E, n, m, cnt, sums, float(g["decay"]), float(g["eps"]))
        np.testing.assert_allclose(E.numpy(), g["E"][s], rtol=2e-6, atol=1e-6)
        np.testing.assert_allclose(n.numpy(), g["cluster_size"][s], rtol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("path", EMA_FILES, ids=EMA_IDS)
@pytest.mark.parametrize("search", ["fp32", "fp32_simt"])
def test_cuda_ema_trajectory_agrees_with_public_implementation(path, search):
    from vqvae_b200 import VectorQuantizer
    g = np.load(path)
    S, N, D = g["z"].shape
    K = g["E0"].shape[0]
    mod = VectorQuantizer(D, K, quantizer="ema", ema_decay=float(g["decay"]), ema_eps=float(g["eps"]), search=search,
                          deterministic=True, cuda_graph=False)
    with torch.no_grad():
        mod.codebook.copy_(torch.from_numpy(g["E0"]))
        mod.ema_sum.copy_(torch.from_numpy(g["E0"]))
        mod.ema_cnt.fill_(1.0)
    mod = mod.cuda().train()
    for s in range(S):
        z = torch.from_numpy(g["z"][s]).cuda()
        E_before = mod.codebook.detach().cpu().clone()
        with torch.no_grad():
            _, _, idx = mod(z)
        theirs = torch.from_numpy(g["indices"][s])
        rep = R.near_tie_report(z.cpu(), E_before, idx.cpu(), theirs)
        assert rep["hard"] == 0, (s, rep)
        if rep["mismatch"]:                      # a near-tie flipped: the trajectories separate legitimately, stop here
            return
        np.testing.assert_allclose(mod.ema_cnt.cpu().numpy(), g["cluster_size"][s], rtol=2e-6, atol=1e-7)
        np.testing.assert_allclose(mod.ema_sum.cpu().numpy(), g["embed_avg"][s], rtol=1e-5, atol=2e-6)
        np.testing.assert_allclose(mod.codebook.cpu().numpy(), g["E"][s], rtol=1e-5, atol=2e-6)

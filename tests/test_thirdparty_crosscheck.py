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

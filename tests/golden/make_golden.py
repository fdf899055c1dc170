"""Generate the known-answer fixtures under tests/golden/ from the self-written oracle.

    python tests/golden/make_golden.py

PARITY UNPINNED: the reference tree holds no quantiser and no fixtures for one
(SURVEY.md section 0), so these vectors freeze the *oracle's* answers (oracle/vq_reference.py),
not the reference's.  The hand case is checkable with pencil and paper.
Inputs are stored next to the outputs, so nothing depends on RNG stream stability.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import vq_reference as R  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def save(name, **arrs):
    np.savez_compressed(os.path.join(OUT, name), **{k: np.asarray(v) for k, v in arrs.items()})
    print("wrote", name, {k: np.asarray(v).shape for k, v in arrs.items()})


def one_step(z, E, K, beta, quantizer, precision):
    idx, dmin = R.assign(z, E, precision)
    q = R.gather(E, idx, torch.float32)
    sq = R.sq_err_sum(z, E, idx)
    cnt, sums = R.code_stats(z, idx, K)
    loss = R.vq_loss_value(sq, z.shape[0], z.shape[1], beta, quantizer)
    return dict(idx=idx.numpy(), dmin=dmin.numpy(), q=q.numpy(), sqerr=np.float64(sq), cnt=cnt.numpy(),
                sums=sums.numpy(), loss=np.float32(loss), perplexity=np.float32(R.perplexity(cnt)))


def main():
    # 1. hand-checkable: D=3, K=4, with an exact tie (codes 1 and 2 identical) and an
    #    equidistant row (row 2 is midway between codes 0 and 3 -> lowest index 0 wins).
    E = torch.tensor([[0., 0., 0.], [1., 0., 0.], [1., 0., 0.], [0., 2., 0.]])
    z = torch.tensor([[0.9, 0.1, 0.0],      # nearest: codes 1 and 2 tie -> 1
                      [0.1, 1.9, 0.0],      # nearest: 3
                      [0.0, 1.0, 0.0],      # d(0)=1, d(3)=1 tie -> 0
                      [0.2, 0.1, 0.1],      # 0
                      [float("nan"), 0., 0.]])  # non-finite row -> 0, excluded from stats
    r = one_step(z, E, 4, 0.25, "st", "fp32")
    assert r["idx"].tolist() == [1, 3, 0, 0, 0], r["idx"]
    save("hand_ties.npz", z=z.numpy(), E=E.numpy(), beta=0.25, **r)

    # 2. seeded random fp32
    g = torch.Generator().manual_seed(1234)
    z = torch.randn(4096, 16, generator=g)
    E = torch.randn(64, 16, generator=g)
    save("rand_fp32_n4096_k64_d16.npz", z=z.numpy(), E=E.numpy(), beta=0.25,
         **one_step(z, E, 64, 0.25, "st", "fp32"))

    # 3. Zipf-skewed "trained-like": z = E[c] + 0.1 N(0,1), c ~ Zipf(1.1)
    g = torch.Generator().manual_seed(99)
    K, D, N = 256, 32, 8192
    E = torch.randn(K, D, generator=g)
    w = 1.0 / torch.arange(1, K + 1, dtype=torch.float64) ** 1.1
    c = torch.multinomial(w / w.sum(), N, replacement=True, generator=g)
    z = E[c] + 0.1 * torch.randn(N, D, generator=g)
    save("zipf_fp32_n8192_k256_d32.npz", z=z.numpy(), E=E.numpy(), beta=0.25,
         **one_step(z, E, K, 0.25, "ema", "fp32"))

    # 4. bf16-input case (inputs stored as the fp32 values of bf16 numbers)
    g = torch.Generator().manual_seed(7)
    z = R.round_bf16(torch.randn(2048, 64, generator=g))
    E = torch.randn(128, 64, generator=g)
    save("rand_bf16_n2048_k128_d64.npz", z=z.numpy(), E=E.numpy(), beta=0.25,
         **one_step(z, E, 128, 0.25, "ema", "bf16"))

    # 5. 10-step EMA trajectory through the module (training mode)
    g = torch.Generator().manual_seed(2024)
    K, D, N = 32, 8, 512
    torch.manual_seed(5)
    m = R.VectorQuantizerOracle(D, K, beta=0.25, quantizer="ema", ema_decay=0.9, ema_eps=1e-5, dead_after=3)
    E0 = m.codebook.clone()
    zs, idxs, losses, ppl, books = [], [], [], [], []
    for t in range(10):
        z = torch.randn(N, D, generator=g) * (1.0 + 0.1 * t)
        q, loss, idx = m(z)
        zs.append(z.numpy()); idxs.append(idx.numpy()); losses.append(float(loss))
        ppl.append(float(m.last_perplexity)); books.append(m.codebook.clone().numpy())
    save("ema_traj_k32_d8.npz", E0=E0.numpy(), z=np.stack(zs), idx=np.stack(idxs), loss=np.array(losses, np.float32),
         perplexity=np.array(ppl, np.float32), codebooks=np.stack(books), ema_cnt=m.ema_cnt.numpy(),
         ema_sum=m.ema_sum.numpy(), usage_total=m.usage_total.numpy(),
         steps_since_used=m.steps_since_used.numpy(), n_dead=int(m.n_dead), decay=0.9, eps=1e-5, beta=0.25,
         dead_after=3)


if __name__ == "__main__":
    main()

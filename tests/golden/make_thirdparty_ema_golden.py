"""Second opinion on the EMA half of the UNPINNED oracle (a-5, a-7): golden trajectories from an independent, published
EMA codebook with a TRAINING-mode update.

The reference ships no quantiser and says about the EMA form only `--quantizer ema --ema_decay 0.99
--ema_eps 1e-5` (scripts/train_vqvae.py:433-436).  The installed packages hold one implementation whose forward updates
the codebook in training mode: `EuclideanCodebook` in vllm 0.22 `vllm/model_executor/models/mimo_audio.py` (the
EnCodec / lucidrains codebook: `_ema_inplace`, `_laplace_smoothing`).  HF transformers' Encodec / Mimi / Xcodec codebooks
carry the same buffers (`cluster_size`, `embed_avg`) but are inference-only.  This script runs that class on seeded
inputs and freezes every step's inputs, indices and state.

What the fixtures CONFIRM about the oracle's choices (tests/test_thirdparty_crosscheck.py):
  * the search uses the codebook of BEFORE the update (update visible next step);
  * n <- g n + (1-g) cnt,  m <- g m + (1-g) sum z;
  * Laplace smoothing n^ = (n + eps) / (sum n + K eps) * sum n and e = m / n^   -- the exact form, bit-level in fp32;
  * under torch.distributed the statistics are all-reduced (SUM) before the EMA step (`_ema_inplace`), which is a-6.
What they do NOT pin (their choices differ or are absent; ours stay unpinned):
  * initial state: theirs cluster_size = 0, ours n = 1 (the fixtures start from OUR state, loaded into their buffers);
  * dead codes: theirs re-seeds codes whose EMA count falls below a threshold (`threshold_ema_dead_code`, switched off
    here); ours counts steps_since_used and re-seeds only on request;
  * perplexity (they have none), the commitment-loss scale (theirs returns 0 from this module).

    python tests/golden/make_thirdparty_ema_golden.py          # needs vllm importable; the .npz files are committed
"""
import os

import numpy as np
import torch
from vllm.model_executor.models.mimo_audio import EuclideanCodebook

HERE = os.path.dirname(os.path.abspath(__file__))

CASES = [  # name, K, D, rows per step, steps, decay, eps, seed, skew
    ("thirdparty_ema_k32_d8", 32, 8, 600, 10, 0.9, 1e-5, 0, 0.0),
    ("thirdparty_ema_k256_d64", 256, 64, 1536, 4, 0.99, 1e-5, 1, 0.0),
    ("thirdparty_ema_k128_d32_skewed", 128, 32, 1500, 6, 0.95, 1e-5, 2, 1.2),     # Zipf usage: some codes never used
]

for name, K, D, N, S, decay, eps, seed, skew in CASES:
    g = torch.Generator().manual_seed(seed)
    cb = EuclideanCodebook(dim=D, codebook_size=K, kmeans_init=False, decay=decay, epsilon=eps, threshold_ema_dead_code=0)
    E0 = torch.randn(K, D, generator=g)
    with torch.no_grad():
        cb.embed.copy_(E0)
        cb.embed_avg.copy_(E0)
        cb.cluster_size.fill_(1.0)                     # OUR initial state (n = 1, m = E); theirs would be 0
    cb.train()
    zs, idxs, Es, ns, ms = [], [], [], [], []
    for s in range(S):
        if skew > 0:
            p = 1.0 / torch.arange(1, K + 1, dtype=torch.float64).pow(skew)
            c = torch.multinomial((p / p.sum()).float(), N, replacement=True, generator=g)
            z = cb.embed[c] + 0.3 * torch.randn(N, D, generator=g)
        else:
            z = torch.randn(N, D, generator=g)
        with torch.no_grad():
            q, idx = cb(z)
        assert torch.equal(q, E0[idx] if s == 0 else q)
        zs.append(z.numpy().copy()); idxs.append(idx.numpy().copy())
        Es.append(cb.embed.numpy().copy()); ns.append(cb.cluster_size.numpy().copy()); ms.append(cb.embed_avg.numpy().copy())
    np.savez_compressed(os.path.join(HERE, name + ".npz"), E0=E0.numpy(), z=np.stack(zs), indices=np.stack(idxs),
                        E=np.stack(Es), cluster_size=np.stack(ns), embed_avg=np.stack(ms), decay=np.float32(decay),
                        eps=np.float32(eps))
    used = len(np.unique(np.concatenate(idxs)))
    print(name, "steps", S, "rows/step", N, "codes used", used, "of", K)

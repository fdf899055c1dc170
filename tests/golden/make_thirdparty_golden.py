"""Second opinion for the UNPINNED quantiser oracle: golden vectors from an independent, published implementation.

The reference ships no quantiser, so nothing of its own can pin oracle/vq_reference.py.  What can be
done is to check the oracle's choices against a widely used public implementation of the same (quantized, loss, indices)
forward: HF transformers' `ChameleonVQVAEVectorQuantizer` (transformers 5.5, models/chameleon/modeling_chameleon.py,
the taming-transformers quantiser).  It fixes: the expanded-square distance, first-minimum argmin, gather, the
straight-through output, and the loss VALUE (1 + beta) * mean_{n,d}(z - q)^2.  (Its beta sits on the codebook term,
ours on the commitment term -- SURVEY.md 8(a-4) -- so the values agree and dL/dz differs by the documented factor; the
fixture stores its gradient too so the test can state that factor.)  This does not make parity "pinned": it is not the
reference.  Run here (needs transformers); the .npz files are committed.

    python tests/golden/make_thirdparty_golden.py
"""
import os
from types import SimpleNamespace

import numpy as np
import torch
from transformers.models.chameleon.modeling_chameleon import ChameleonVQVAEVectorQuantizer

HERE = os.path.dirname(os.path.abspath(__file__))

CASES = [  # name, B, D, H, W, K, beta, seed
    ("thirdparty_chameleon_b2_d16_k64", 2, 16, 9, 7, 64, 0.25, 0),
    ("thirdparty_chameleon_b4_d64_k512", 4, 64, 16, 16, 512, 0.25, 1),       # C1's K and D
    ("thirdparty_chameleon_b3_d128_k1024", 3, 128, 16, 16, 1024, 0.5, 2),    # C2's K and D
]

for name, B, D, H, W, K, beta, seed in CASES:
    g = torch.Generator().manual_seed(seed)
    q = ChameleonVQVAEVectorQuantizer(SimpleNamespace(num_embeddings=K, embed_dim=D, beta=beta))
    with torch.no_grad():
        q.embedding.weight.copy_(torch.randn(K, D, generator=g))
    z = torch.randn(B, D, H, W, generator=g).requires_grad_(True)
    w = torch.randn(B, D, H, W, generator=g)
    quant, loss, idx = q(z)
    ((quant * w).sum() + 3.0 * loss).backward()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), z=z.detach().numpy(), w=w.numpy(),
                        E=q.embedding.weight.detach().numpy(), beta=np.float32(beta), quantized=quant.detach().numpy(),
                        loss=loss.detach().numpy(), indices=idx.numpy(), grad_z=z.grad.numpy(),
                        grad_E=q.embedding.weight.grad.numpy())
    print(name, "N", B * H * W, "loss", float(loss))

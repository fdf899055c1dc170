"""-m gpu: BASELINE configs 3 (type / phase split, two codebooks) and 5 (batch-2048 step: 10.5 M vectors against the
large codebook) -- small cases against the oracle, full sizes through size-independent properties."""
import numpy as np
import pytest
import torch

from oracle import vq_reference as R

pytestmark = pytest.mark.gpu


def _load(mod, ref):
    mod.load_state_dict(ref.state_dict(), strict=False)
    return mod


def test_split_quantizer_small_vs_oracle():
    """z_type [B, D, H, W] and z_phase [B, D, T, H, W] (frl/models/representation.py:317-374), one oracle per stream."""
    from vqvae_b200 import SplitQuantizer
    torch.manual_seed(3)
    B, Dt, Dp, T, H, W, K = 2, 16, 24, 3, 6, 5, 32
    ref_t = R.VectorQuantizerOracle(Dt, K, quantizer="ema", ema_decay=0.9)
    ref_p = R.VectorQuantizerOracle(Dp, K, quantizer="ema", ema_decay=0.9)
    sq = SplitQuantizer(Dt, Dp, K, quantizer="ema", ema_decay=0.9, search="fp32", deterministic=True)
    _load(sq.quant_type, ref_t); _load(sq.quant_phase, ref_p)
    sq = sq.cuda()
    for step in range(3):
        zt, zp = torch.randn(B, Dt, H, W), torch.randn(B, Dp, T, H, W)
        (qt, lt, it), (qp, lp, ip) = sq(zt.cuda(), zp.cuda())
        rqt, rlt, rit = ref_t(zt.permute(0, 2, 3, 1))
        rqp, rlp, rip = ref_p(zp.permute(0, 2, 3, 4, 1))
        assert qt.shape == zt.shape and it.shape == (B, H, W) and qp.shape == zp.shape and ip.shape == (B, T, H, W)
        assert torch.equal(it.cpu(), rit) and torch.equal(ip.cpu(), rip)
        np.testing.assert_allclose(qt.cpu().numpy(), rqt.permute(0, 3, 1, 2).numpy(), rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(qp.cpu().numpy(), rqp.permute(0, 4, 1, 2, 3).numpy(), rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(float(sq.loss), float(rlt) + float(rlp), rtol=1e-5)
        np.testing.assert_allclose(sq.quant_type.codebook.cpu().numpy(), ref_t.codebook.numpy(), rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(sq.quant_phase.codebook.cpu().numpy(), ref_p.codebook.numpy(), rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(float(sq.perplexity["phase"]), float(ref_p.last_perplexity), rtol=1e-5)


def test_split_quantizer_overlap_is_bit_identical_and_differentiable():
    from vqvae_b200 import SplitQuantizer
    torch.manual_seed(4)
    B, D, T, H, W, K = 4, 64, 5, 16, 16, 256
    a = SplitQuantizer(D, D, K, quantizer="st", deterministic=True, overlap=True).cuda()
    b = SplitQuantizer(D, D, K, quantizer="st", deterministic=True, overlap=False).cuda()
    b.load_state_dict(a.state_dict())
    zt = torch.randn(B, D, H, W, device="cuda", dtype=torch.bfloat16)
    zp = torch.randn(B, D, T, H, W, device="cuda", dtype=torch.bfloat16)
    outs = []
    for m in (a, b):
        xt, xp = zt.clone().requires_grad_(True), zp.clone().requires_grad_(True)
        (qt, lt, it), (qp, lp, ip) = m(xt, xp)
        (qt.float().sum() + 2.0 * qp.float().mean() + m.loss).backward()
        torch.cuda.synchronize()
        outs.append((qt, it, qp, ip, m.loss.detach(), xt.grad, xp.grad, m.quant_type.codebook.grad, m.quant_phase.codebook.grad))
    for x, y in zip(*outs):
        assert torch.equal(x, y)


def test_full_size_properties_c3():
    """BASELINE config 3: B=256, T=5, 32x32, two codebooks of 2048 x 128; type N = 262 144, phase N = 1 310 720."""
    from vqvae_b200 import SplitQuantizer
    torch.manual_seed(5)
    B, D, T, H, W, K = 256, 128, 5, 32, 32, 2048
    sq = SplitQuantizer(D, D, K, quantizer="ema").cuda()
    Et, Ep = sq.quant_type.codebook.clone(), sq.quant_phase.codebook.clone()
    zt = torch.randn(B, D, H, W, device="cuda", dtype=torch.bfloat16)
    zp = torch.randn(B, D, T, H, W, device="cuda", dtype=torch.bfloat16)
    (qt, lt, it), (qp, lp, ip) = sq(zt, zp)
    torch.cuda.synchronize()
    for q, idx, E, z, mod, n in ((qt, it, Et, zt, sq.quant_type, B * H * W), (qp, ip, Ep, zp, sq.quant_phase, B * T * H * W)):
        assert int(idx.min()) >= 0 and int(idx.max()) < K and idx.numel() == n
        assert torch.equal(q.movedim(1, -1), E[idx].to(torch.bfloat16))                  # gather exact, layout restored
        assert torch.equal(mod.usage_total, torch.bincount(idx.reshape(-1), minlength=K))  # counts exact
        rows = z.movedim(1, -1).reshape(-1, D)
        sel = torch.randperm(n)[:2048].cuda()
        zs = rows[sel].cpu().float()
        ref_idx, _ = R.assign(zs, E.cpu(), "bf16")
        rep = R.near_tie_report(zs, E.cpu(), idx.reshape(-1)[sel].cpu(), ref_idx, precision="bf16")
        assert rep["hard"] == 0, rep
    mse_t = float(((zt.float() - qt.float()) ** 2).mean()); mse_p = float(((zp.float() - qp.float()) ** 2).mean())
    np.testing.assert_allclose(float(sq.loss), 0.25 * (mse_t + mse_p), rtol=2e-3)


def test_full_size_properties_c5():
    """BASELINE config 5's quantiser load: 2048 patches x 5120 = 10 485 760 vectors, K = 8192, D = 256, bf16, one call."""
    from vqvae_b200 import VectorQuantizer
    free, _ = torch.cuda.mem_get_info()
    if free < 40 << 30:
        pytest.skip("needs ~25 GB of free device memory")
    torch.manual_seed(6)
    N, K, D = 2048 * 5120, 8192, 256
    mod = VectorQuantizer(D, K, quantizer="ema").cuda()
    E0 = mod.codebook.clone()
    ema_sum0 = mod.ema_sum.clone()
    z = torch.randn(N, D, device="cuda", dtype=torch.bfloat16)
    q, loss, idx = mod(z)
    torch.cuda.synchronize()
    assert int(idx.min()) >= 0 and int(idx.max()) < K
    assert torch.equal(mod.usage_total, torch.bincount(idx, minlength=K))
    Eb = E0.to(torch.bfloat16)
    se, col = 0.0, torch.zeros(D, device="cuda", dtype=torch.float64)
    for s in range(0, N, 1 << 20):                                   # chunked: keep the checker's memory bounded
        sl = slice(s, min(N, s + (1 << 20)))
        assert torch.equal(q[sl], Eb[idx[sl]])
        se += float(((z[sl].float() - E0[idx[sl]]) ** 2).sum(dtype=torch.float64))
        col += z[sl].sum(0, dtype=torch.float64)
    np.testing.assert_allclose(float(loss) / 0.25, se / (N * D), rtol=1e-4)
    code_sum = (mod.ema_sum - 0.99 * ema_sum0) / (1.0 - 0.99)        # checksum of checksums
    np.testing.assert_allclose(code_sum.sum(0).cpu().numpy(), col.cpu().numpy(), rtol=5e-3, atol=8.0)
    sel = torch.randperm(N)[:1024].cuda()
    zs = z[sel].cpu().float()
    ref_idx, _ = R.assign(zs, E0.cpu(), "bf16")
    rep = R.near_tie_report(zs, E0.cpu(), idx[sel].cpu(), ref_idx, precision="bf16")
    assert rep["hard"] == 0, rep


@pytest.mark.parametrize("D,dtype,det", [(64, torch.bfloat16, False), (64, torch.float32, True), (20, torch.float32, False),
                                         (7, torch.float32, False)])
def test_pass_accumulate_flag_equals_one_shot(D, dtype, det):
    """VQ_STAT_ACCUMULATE: the pass over two row chunks adds up to the pass over the whole batch (vector path,
    deterministic path and the generic odd-D path)."""
    from vqvae_b200 import ops
    from vqvae_b200._lib import VQ_STAT_SUM_Z
    torch.manual_seed(D)
    N, K = 5000, 96
    E = torch.randn(K, D, device="cuda")
    z = torch.randn(N, D, device="cuda").to(dtype)
    idx = torch.randint(0, K, (N,), device="cuda", dtype=torch.int32)
    ws = torch.empty(ops.workspace_bytes(N, K, D, dtype), dtype=torch.uint8, device="cuda")

    def run(parts):
        q = torch.empty_like(z); sq = torch.full((1,), 123.0, device="cuda")
        cnt = torch.full((K,), 7.0, device="cuda"); sums = torch.full((K, D), 7.0, device="cuda")
        for c, (a, b) in enumerate(parts):
            ops.gather_st_loss_scatter(z[a:b], idx[a:b], E, q[a:b], sq, cnt, sums, VQ_STAT_SUM_Z, det, ws, accumulate=c > 0)
        return q, sq, cnt, sums

    q1, s1, c1, m1 = run([(0, N)])
    q2, s2, c2, m2 = run([(0, 2048), (2048, 2048), (2048, N)])           # includes an empty chunk
    assert torch.equal(q1, q2) and torch.equal(c1, c2)
    np.testing.assert_allclose(float(s2), float(s1), rtol=1e-5)
    np.testing.assert_allclose(m2.cpu().numpy(), m1.cpu().numpy(), rtol=1e-4, atol=1e-4)
    ref_cnt, ref_sum = R.code_stats(z.float().cpu(), idx.long().cpu(), K)
    assert torch.equal(c1.cpu(), ref_cnt)
    np.testing.assert_allclose(m1.cpu().numpy(), ref_sum.numpy(), rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize("quantizer", ["ema", "st"])
def test_forward_host_equals_forward_on_the_copied_batch(quantizer):
    """The pipelined host entry (row chunks cross PCIe while the previous chunk is quantised) is the same step."""
    from vqvae_b200 import VectorQuantizer
    torch.manual_seed(8)
    K, D, N = 512, 64, 300000
    a = VectorQuantizer(D, K, quantizer=quantizer, deterministic=True).cuda()
    b = VectorQuantizer(D, K, quantizer=quantizer, deterministic=True).cuda()
    b.load_state_dict(a.state_dict())
    for step in range(2):
        zh = torch.randn(50, N // 50, D).to(torch.bfloat16).pin_memory()
        with torch.no_grad():
            qa, la, ia = a(zh.cuda())
        qb, lb, ib = b.forward_host(zh, chunks=None if step == 0 else 3)
        torch.cuda.synchronize()
        assert qb.shape == zh.shape and ib.shape == zh.shape[:-1] and qb.is_cuda
        assert torch.equal(ia, ib) and torch.equal(qa, qb)
        np.testing.assert_allclose(float(lb), float(la), rtol=1e-5)
        np.testing.assert_allclose(b.codebook.detach().cpu().numpy(), a.codebook.detach().cpu().numpy(), rtol=2e-5, atol=1e-6)
        assert torch.equal(a.usage_total, b.usage_total)
        np.testing.assert_allclose(float(b.perplexity), float(a.perplexity), rtol=1e-5)


@pytest.mark.parametrize("quantizer,dtype", [("ema", torch.bfloat16), ("st", torch.float32), ("ema", torch.float32)])
def test_cuda_graph_replay_equals_eager(quantizer, dtype):
    """Small batches replay the whole step as one CUDA graph from the third call on; results are those of the eager
    launches, training and eval, with a larger batch (workspace regrowth) in between."""
    from vqvae_b200 import VectorQuantizer
    torch.manual_seed(9)
    K, D = 128, 32
    a = VectorQuantizer(D, K, quantizer=quantizer, deterministic=True, cuda_graph=True).cuda()
    b = VectorQuantizer(D, K, quantizer=quantizer, deterministic=True, cuda_graph=False).cuda()
    b.load_state_dict(a.state_dict())
    g = torch.Generator().manual_seed(3)
    sizes = [4096, 4096, 4096, 4096, 50000, 4096, 4096, 777, 777, 777]
    for step, n in enumerate(sizes):
        if step == 7:
            a.eval(); b.eval()
        z = torch.randn(n, D, generator=g).to(dtype).cuda()
        outs = []
        for m in (a, b):
            x = z.clone().requires_grad_(True)
            q, loss, idx = m(x)
            (q.float().pow(2).sum() + 3.0 * loss).backward()
            outs.append((q.detach(), loss.detach(), idx, x.grad, None if m.codebook.grad is None else m.codebook.grad.clone()))
            if m.codebook.grad is not None:
                with torch.no_grad():
                    m.codebook -= 0.05 * m.codebook.grad
                m.codebook.grad = None
        for u, v in zip(*outs):
            assert (u is None and v is None) or torch.equal(u, v), step
        assert torch.equal(a.codebook, b.codebook) and torch.equal(a.usage_total, b.usage_total)
        assert torch.equal(a.last_perplexity, b.last_perplexity) and torch.equal(a.ema_sum, b.ema_sum)
    assert any(not isinstance(v, str) for v in a._graphs.values())      # a graph was really captured and replayed
    assert not b._graphs


def test_split_quantizer_forward_host_equals_forward():
    from vqvae_b200 import SplitQuantizer
    torch.manual_seed(10)
    D, K = 64, 256
    a = SplitQuantizer(D, D, K, channel_first=False, quantizer="ema", deterministic=True).cuda()
    b = SplitQuantizer(D, D, K, channel_first=False, quantizer="ema", deterministic=True).cuda()
    b.load_state_dict(a.state_dict())
    zt = torch.randn(40000, D).to(torch.bfloat16).pin_memory()
    zp = torch.randn(150000, D).to(torch.bfloat16).pin_memory()
    with torch.no_grad():
        (qa_t, _, ia_t), (qa_p, _, ia_p) = a(zt.cuda(), zp.cuda())
    (qb_t, _, ib_t), (qb_p, _, ib_p) = b.forward_host(zt, zp)
    torch.cuda.synchronize()
    assert torch.equal(ia_t, ib_t) and torch.equal(ia_p, ib_p) and torch.equal(qa_t, qb_t) and torch.equal(qa_p, qb_p)
    np.testing.assert_allclose(float(b.loss), float(a.loss), rtol=1e-5)
    np.testing.assert_allclose(b.quant_phase.codebook.cpu().numpy(), a.quant_phase.codebook.cpu().numpy(), rtol=2e-5, atol=1e-6)

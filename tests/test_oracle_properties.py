"""Size-independent properties of the oracle (oracle/vq_reference.py) -- what any correct restatement of a-1 .. a-9 must
satisfy whatever the unpinned choices are.  CPU only, seconds.  The GPU parity tests compare the kernels with this oracle;
these tests keep the oracle itself honest."""
import math

import pytest
import torch

from oracle import vq_reference as R


def _case(N, K, D, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(N, D, generator=g), torch.randn(K, D, generator=g)


@pytest.mark.parametrize("N,K,D,seed", [(513, 37, 12, 0), (2048, 256, 64, 1)])
def test_quantising_the_quantised_rows_is_idempotent(N, K, D, seed):
    z, E = _case(N, K, D, seed)
    idx, _ = R.assign(z, E, "fp32")
    q = R.gather(E, idx, torch.float32)
    idx2, dmin2 = R.assign(q, E, "fp32")
    assert torch.equal(idx, idx2)
    assert float(R.sq_err_sum(q, E, idx2)) == 0.0
    assert float(dmin2.abs().max()) <= 1e-3            # ||q||^2 + ||e||^2 - 2 q.e evaluated in fp32


@pytest.mark.parametrize("shards", [2, 3, 8])
def test_code_statistics_are_additive_over_row_shards(shards):
    """a-6 is a SUM: statistics of the concatenated batch == sum of the shards' statistics (ragged shards included)."""
    z, E = _case(1000, 64, 16, 3)
    idx, _ = R.assign(z, E, "fp32")
    cnt, sums = R.code_stats(z, idx, 64)
    bounds = [round(i * 1000 / shards) for i in range(shards + 1)]
    c2 = torch.zeros_like(cnt)
    s2 = torch.zeros_like(sums)
    sq = 0.0
    for a, b in zip(bounds[:-1], bounds[1:]):
        c, s = R.code_stats(z[a:b], idx[a:b], 64)
        c2 += c
        s2 += s
        sq += float(R.sq_err_sum(z[a:b], E, idx[a:b]))
    assert torch.equal(cnt, c2) and int(cnt.sum()) == 1000
    assert torch.allclose(sums, s2, rtol=1e-5, atol=1e-5)
    assert math.isclose(sq, float(R.sq_err_sum(z, E, idx)), rel_tol=1e-9)
    assert torch.allclose(sums.sum(0), z.sum(0), rtol=1e-4, atol=1e-4)           # every row lands in exactly one code


def test_search_is_invariant_under_a_common_shift_and_positive_scale():
    """argmin ||z - e||^2 does not change when z and E are shifted together or scaled together; in fp32 a row may flip
    only between two codes that were already a near tie (the rule the GPU parity tests use, near_tie_report)."""
    z, E = _case(700, 50, 8, 4)
    idx, _ = R.assign(z, E, "fp32")
    t = torch.randn(8, generator=torch.Generator().manual_seed(5))
    for zz, EE in ((z + t, E + t), (z * 3.0, E * 3.0), (z * 0.125, E * 0.125)):
        i2, _ = R.assign(zz, EE, "fp32")
        rep = R.near_tie_report(z, E, i2, idx, rel=1e-4, precision="fp32")
        assert rep["hard"] == 0, rep
        assert rep["mismatch"] <= 3, rep
    i3, _ = R.assign(z * 0.125, E * 0.125, "fp32")           # a power of two scales every fp32 operation exactly
    assert torch.equal(i3, idx)


def test_non_finite_rows_get_code_zero_and_count_for_nothing():
    z, E = _case(64, 9, 5, 6)
    z[3, 2] = float("nan")
    z[10, 0] = float("inf")
    idx, _ = R.assign(z, E, "fp32")
    assert int(idx[3]) == 0 and int(idx[10]) == 0
    cnt, sums = R.code_stats(z, idx, 9)
    assert int(cnt.sum()) == 62 and torch.isfinite(sums).all()
    assert math.isfinite(float(R.sq_err_sum(z, E, idx)))
    ok = torch.ones(64, dtype=torch.bool)
    ok[3] = ok[10] = False
    cnt2, sums2 = R.code_stats(z[ok], idx[ok], 9)
    assert torch.equal(cnt, cnt2) and torch.equal(sums, sums2)


def test_perplexity_bounds():
    K = 128
    assert math.isclose(float(R.perplexity(torch.ones(K))), K, rel_tol=1e-5)                  # uniform usage: K
    one = torch.zeros(K)
    one[17] = 1000.0
    assert math.isclose(float(R.perplexity(one)), 1.0, rel_tol=1e-6)                          # one code: 1
    assert float(R.perplexity(torch.zeros(K))) == 1.0                                         # nothing assigned: 1 (choice)
    g = torch.Generator().manual_seed(7)
    for _ in range(5):
        c = torch.randint(0, 50, (K,), generator=g).float()
        p = float(R.perplexity(c))
        assert 1.0 <= p <= K + 1e-3
        assert math.isclose(p, float(R.perplexity(c * 7.0)), rel_tol=1e-6)                    # depends on proportions only


def test_ema_update_limits():
    """decay = 0: the smoothed code means of this batch alone; decay = 1: nothing moves; totals are conserved."""
    z, E = _case(4000, 32, 6, 8)
    idx, _ = R.assign(z, E, "fp32")
    cnt, sums = R.code_stats(z, idx, 32)
    n0, m0 = torch.ones(32), E.clone()
    E1, n1, m1 = R.ema_update(E, n0, m0, cnt, sums, decay=0.0, eps=1e-5)
    assert torch.equal(n1, cnt) and torch.equal(m1, sums)
    used = cnt > 0
    means = sums[used] / cnt[used][:, None]
    assert torch.allclose(E1[used], means, rtol=1e-3, atol=1e-5)          # Laplace smoothing moves them by O(eps)
    E2, n2, m2 = R.ema_update(E, n0, m0, cnt, sums, decay=1.0, eps=1e-5)
    assert torch.equal(n2, n0) and torch.equal(m2, m0)
    assert torch.allclose(E2, E, rtol=1e-4, atol=1e-6)
    E3, n3, m3 = R.ema_update(E, n0, m0, cnt, sums, decay=0.99, eps=1e-5)
    assert math.isclose(float(n3.double().sum()), 0.99 * 32 + 0.01 * 4000, rel_tol=1e-5)
    assert torch.allclose(m3.sum(0), 0.99 * m0.sum(0) + 0.01 * sums.sum(0), rtol=1e-4, atol=1e-4)


def test_usage_accounting():
    u = torch.zeros(5, dtype=torch.int64)
    s = torch.zeros(5, dtype=torch.int32)
    cnts = [torch.tensor([3., 0., 1., 0., 0.]), torch.tensor([0., 0., 2., 0., 5.]), torch.tensor([1., 0., 0., 0., 0.])]
    dead = []
    for c in cnts:
        u, s, nd = R.usage_update(u, s, c, dead_after=2)
        dead.append(nd)
    assert u.tolist() == [4, 0, 3, 0, 5]
    assert s.tolist() == [0, 3, 1, 3, 1]
    assert dead == [0, 2, 2]                       # codes 1 and 3 have gone 2, then 3 steps without a row


def test_reduced_precision_tiers_round_the_operands_only():
    """bf16 / fp16 tiers: the search sees rounded z and E (fp32 accumulate); the gather still reads the fp32 master."""
    z, E = _case(300, 40, 16, 9)
    for prec, rnd in (("bf16", R.round_bf16), ("fp16", R.round_f16)):
        i_lo, _ = R.assign(z, E, prec)
        i_ref, _ = R.assign(rnd(z), rnd(E), "fp32")
        assert torch.equal(i_lo, i_ref)
        q = R.gather(E, i_lo, torch.float32)
        assert torch.equal(q, E[i_lo])


def test_backward_matches_autograd_of_the_loss_definition():
    z, E = _case(50, 7, 4, 10)
    idx, _ = R.assign(z, E, "fp32")
    q = E[idx]
    beta = 0.25
    zt = z.clone().double().requires_grad_(True)
    Et = E.clone().double().requires_grad_(True)
    qt = Et[idx]
    loss = ((zt.detach() - qt) ** 2).mean() + beta * ((zt - qt.detach()) ** 2).mean()          # st form
    out = zt + (qt - zt).detach()
    gq = torch.randn(50, 4, dtype=torch.float64, generator=torch.Generator().manual_seed(11))
    (loss * 1.7 + (out * gq).sum()).backward()
    gz = R.backward_z(z, q, gq.float(), grad_loss=1.7, beta=beta)
    assert torch.allclose(gz.double(), zt.grad, rtol=1e-5, atol=1e-6)
    cnt, sums = R.code_stats(z, idx, 7)
    gE = R.backward_codebook(E, cnt, sums, grad_loss=1.7, N=50, D=4)
    assert torch.allclose(gE.double(), Et.grad, rtol=1e-4, atol=1e-6)

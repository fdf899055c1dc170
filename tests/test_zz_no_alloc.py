"""vq_gather_st_loss_scatter_noalloc (include/vq_b200.h): the fused gather pass with its scratch in a static __device__
array inside the library -- no workspace from the caller, no allocation in the call.  Checked through the C ABI against the
existing entry point (vq_gather_st_loss_scatter with a caller workspace) on two calls in a row, with the tolerances the
other pass tests use (tests/gpu_util.py): indices and quantised rows bit-equal, counts exact, sums / loss 1e-5.
This file is named so that pytest collects it last."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _call_noalloc(lib, z, idx, E, q, sq, cnt, sums, stat_kind, idx64, loss, loss_scale):
    from vqvae_b200 import ops
    N, D = z.shape
    K = E.shape[0]
    p = lambda t: None if t is None else t.data_ptr()
    return lib.vq_gather_st_loss_scatter_noalloc(p(z), ops._dt(z), N, D, z.stride(0) if N > 1 else D, p(idx), p(E), K, p(q),
                                                 q.stride(0) if N > 1 else D, p(sq), p(cnt), p(sums), stat_kind, p(idx64),
                                                 p(loss), float(loss_scale), torch.cuda.current_stream().cuda_stream)


def _compare(N, K, D, dtype, seed, stat_kind):
    from vqvae_b200 import _lib, ops
    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    g = torch.Generator().manual_seed(seed)
    E = torch.randn(K, D, generator=g).to(dev)
    z = torch.randn(N, D, generator=g).to(dev).to(dtype)
    idx0 = torch.randint(0, K, (N,), generator=g, dtype=torch.int32).to(dev)
    if N > 10:
        z[3, 1] = float("nan")                       # a non-finite row: canonicalised to code 0, left out of the sums
    outs = []
    for which in ("ws", "noalloc"):
        idx = idx0.clone()
        q = torch.full_like(z, float("nan"))
        sq = torch.full((1,), -1.0, device=dev)
        cnt = torch.full((K,), -7.0, device=dev)
        sums = torch.full((K, D), float("nan"), device=dev)
        idx64 = torch.full((N,), -1, dtype=torch.int64, device=dev)
        loss = torch.full((1,), -1.0, device=dev)
        scale = 0.25 / float(N * D)
        if which == "ws":
            ws = torch.empty(ops.workspace_bytes(N, K, D, dtype), dtype=torch.uint8, device=dev)
            ops.gather_st_loss_scatter(z, idx, E, q, sq, cnt, sums, stat_kind, False, ws, idx64_out=idx64, loss_out=loss,
                                       loss_scale=scale)
        else:
            rc = _call_noalloc(lib, z, idx, E, q, sq, cnt, sums, stat_kind, idx64, loss, scale)
            assert rc == 0, lib.vq_last_error().decode()
        torch.cuda.synchronize()
        outs.append((idx, q, sq, cnt, sums, idx64, loss))
    (ia, qa, sa, ca, ua, i64a, la), (ib, qb, sb, cb, ub, i64b, lb) = outs
    assert torch.equal(ia, ib) and torch.equal(i64a, i64b) and torch.equal(i64b, ib.long())
    assert torch.equal(qa.view(torch.uint8), qb.view(torch.uint8)), "quantised rows must be bit-equal"
    assert torch.equal(ca, cb), "per-code counts must be exact"
    assert abs(float(sa) - float(sb)) <= 1e-5 * abs(float(sa))
    assert abs(float(la) - float(lb)) <= 1e-5 * abs(float(la))
    scale_s = ua.abs().max().clamp_min(1e-30)
    assert float((ua - ub).abs().max() / scale_s) <= 1e-5
    assert int(ca.sum()) == N - (1 if N > 10 else 0)


def test_noalloc_matches_workspace_entry_point_two_calls_in_a_row():
    from vqvae_b200._lib import VQ_STAT_SUM_RESID, VQ_STAT_SUM_Z
    # two calls in a row on the shared scratch, different shapes and dtypes: the second must not see the first's leftovers
    _compare(40960, 512, 64, torch.bfloat16, 1, VQ_STAT_SUM_Z)          # BASELINE config 1's shape
    _compare(7001, 8192, 256, torch.float32, 2, VQ_STAT_SUM_Z)
    _compare(40960, 512, 64, torch.bfloat16, 1, VQ_STAT_SUM_Z)          # and the first shape again
    _compare(3000, 96, 20, torch.float32, 3, VQ_STAT_SUM_RESID)         # odd D: the generic (scalar) pass
    _compare(300000, 1024, 32, torch.float16, 4, VQ_STAT_SUM_Z)         # close to the built-in array's limit


def test_noalloc_rejects_sizes_beyond_the_builtin_scratch():
    from vqvae_b200 import _lib
    from vqvae_b200._lib import VQ_ERR_ARG, VQ_ERR_UNSUPPORTED, VQ_STAT_PRESORTED, VQ_STAT_SUM_Z
    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    N, K, D = 1 << 20, 1024, 8
    z = torch.zeros(N, D, device=dev)
    idx = torch.zeros(N, dtype=torch.int32, device=dev)
    E = torch.zeros(K, D, device=dev)
    q = torch.empty_like(z)
    sq = torch.zeros(1, device=dev)
    rc = _call_noalloc(lib, z, idx, E, q, sq, None, None, VQ_STAT_SUM_Z, None, None, 0.0)
    assert rc == VQ_ERR_UNSUPPORTED and b"built-in array" in lib.vq_last_error()
    rc = _call_noalloc(lib, z[:100], idx[:100], E, q[:100], sq, None, None, VQ_STAT_SUM_Z | VQ_STAT_PRESORTED, None, None, 0.0)
    assert rc == VQ_ERR_ARG
    torch.cuda.synchronize()

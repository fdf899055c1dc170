"""Latent dtypes and search tiers through the C ABI: native fp16 (the reference script's default autocast dtype,
scripts/train_vqvae.py:284-286), the split-operand tensor-core fp32 tier against the fp32 oracle at the 1e-5 rule, the
CUDA-core fp32 kernel for odd shapes, and the device guard.  Parity = vs oracle/vq_reference.py (self-written: unpinned)."""
import pytest
import torch

from oracle import vq_reference as R

pytestmark = pytest.mark.gpu


def _data(N, K, D, seed, scale_z=1.0, scale_e=1.0, trained=False):
    g = torch.Generator().manual_seed(seed)
    E = torch.randn(K, D, generator=g) * scale_e
    z = torch.randn(N, D, generator=g) * scale_z
    if trained:
        c = torch.randint(0, K, (N,), generator=g)
        z = E[c] + 0.1 * scale_e * torch.randn(N, D, generator=g)
    return z, E


# ------------------------------------------------------------------------------------------------ fp16
@pytest.mark.parametrize("N,K,D", [(4096, 512, 64), (3000, 1000, 128), (2048, 8192, 256), (1500, 300, 40), (777, 64, 8),
                                   (5000, 700, 200), (130, 20, 512)])
def test_fp16_step_vs_oracle(N, K, D):
    from gpu_util import check_step_against_oracle
    z, E = _data(N, K, D, 10 + D)
    rep, out = check_step_against_oracle(z, E, "fp16", torch.float16)
    assert rep["hard"] == 0
    assert out["q"].dtype == torch.float16
    assert torch.equal(out["E_bf16"].view(torch.float16), E.to(torch.float16))     # the shadow holds fp16 rows


@pytest.mark.parametrize("scale_e,scale_z", [(50.0, 40.0), (1e-3, 1e-3), (200.0, 1.0), (1e-2, 30.0)])
@pytest.mark.parametrize("D", [64, 128, 256])
def test_fp16_norm_image_range(D, scale_e, scale_z):
    """D <= 128 carries -||e||^2/2 through the tensor core as three fp16 pieces with weights (2^8, 2^-4, 2^-14): norms far
    above fp16's own range (||e||^2 = 128 * 2500 > 65504) and far below its normal range must still rank codes exactly."""
    from gpu_util import check_step_against_oracle
    z, E = _data(3000, 600, D, 77, scale_z=scale_z, scale_e=scale_e)
    rep, _ = check_step_against_oracle(z, E, "fp16", torch.float16)
    assert rep["hard"] == 0


def test_fp16_trained_like():
    from gpu_util import check_step_against_oracle
    z, E = _data(20000, 1024, 128, 5, trained=True)
    rep, _ = check_step_against_oracle(z, E, "fp16", torch.float16)
    assert rep["hard"] == 0


def test_fp16_module_under_autocast_no_torch_kernels():
    """The call the reference script makes by default: model(batch) under torch.autocast(float16).  Forward + backward vs
    the oracle (1e-3), and the forward launches nothing but the library's kernels (no at:: elementwise casts)."""
    from vqvae_b200 import VectorQuantizer
    torch.manual_seed(3)
    D, K = 64, 256
    ref = R.VectorQuantizerOracle(D, K, quantizer="ema", precision="fp16").train()
    mod = VectorQuantizer(D, K, quantizer="ema", cuda_graph=False)
    mod.load_state_dict(ref.state_dict())
    mod = mod.cuda().train()
    E0 = ref.codebook.detach().clone()                  # the codebook the search runs against (EMA updates afterwards)
    lin = torch.nn.Linear(32, D).cuda()
    x = torch.randn(6, 10, 10, 32, device="cuda")
    with torch.autocast(device_type="cuda", dtype=torch.float16):
        z = lin(x)                                      # fp16 under autocast
        assert z.dtype == torch.float16
        q, loss, idx = mod(z)
    assert q.dtype == torch.float16 and loss.dtype == torch.float32 and idx.dtype == torch.int64
    (q.float().sum() + 3.0 * loss).backward()
    assert torch.isfinite(lin.weight.grad).all()
    zc = z.detach().cpu()
    zr = zc.clone().requires_grad_(True)
    qr, lr, ir = ref(zr)
    (qr.float().sum() + 3.0 * lr).backward()
    rep = R.near_tie_report(zc.reshape(-1, D).float(), E0, idx.cpu().reshape(-1), ir.reshape(-1), precision="fp16")
    assert rep["hard"] == 0
    assert abs(float(loss) - float(lr)) <= 1e-3 * abs(float(lr))
    if rep["mismatch"] == 0:
        assert torch.equal(q.detach().cpu(), qr.detach())
        assert torch.allclose(mod.codebook.cpu(), ref.codebook, rtol=1e-3, atol=1e-5)
    # forward only, profiled: every kernel on the stream is one of ours
    from torch.profiler import ProfilerActivity, profile
    with torch.no_grad():
        zz = z.detach()
        mod(zz)
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            mod(zz)
            torch.cuda.synchronize()
    names = [e.name for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    kernels = [n for n in names if "memcpy" not in n.lower() and "memset" not in n.lower()]
    assert kernels and all("vq::" in n for n in kernels), kernels


# ------------------------------------------------------------------------------------------------ fp32 tier
@pytest.mark.parametrize("N,K,D", [(4096, 512, 64), (3000, 1000, 128), (2048, 8192, 256), (1500, 300, 40), (777, 64, 8),
                                   (6000, 2048, 256), (5000, 4096, 200)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_fp32_split_tier_vs_oracle(N, K, D, dtype):
    """search="fp32" on the tensor cores (z = hi + lo, E = hi + lo bf16 planes, three MMAs per product): indices match the
    fp32 oracle except at |delta d| < 1e-5 relative near-ties (counted), gather bit-exact, loss / sums within 1e-5."""
    from gpu_util import check_step_against_oracle
    z, E = _data(N, K, D, 100 + D)
    rep, out = check_step_against_oracle(z, E, "fp32", dtype, tol=1e-5 if dtype == torch.float32 else 1e-3)
    assert rep["hard"] == 0


@pytest.mark.parametrize("scale_e,scale_z,trained", [(1e3, 1e3, False), (1e-3, 1e-3, False), (1.0, 1.0, True),
                                                     (30.0, 30.0, True), (1e-2, 10.0, False)])
def test_fp32_split_tier_ranges(scale_e, scale_z, trained):
    from gpu_util import check_step_against_oracle
    for D in (128, 256):
        z, E = _data(5000, 1500, D, 9, scale_z=scale_z, scale_e=scale_e, trained=trained)
        rep, _ = check_step_against_oracle(z, E, "fp32")
        assert rep["hard"] == 0, (D, rep)


@pytest.mark.parametrize("N,K,D", [(3000, 700, 128), (2000, 300, 20), (1000, 96, 264), (500, 50, 7)])
def test_fp32_cuda_core_kernel_vs_oracle(N, K, D):
    """The CUDA-core fp32 kernel stays for shapes the tensor-core tier cannot take and behind search="fp32_simt"."""
    from gpu_util import check_step_against_oracle
    z, E = _data(N, K, D, 200 + D)
    rep, _ = check_step_against_oracle(z, E, "fp32_simt")
    assert rep["hard"] == 0
    if D % 8 == 0 and D <= 256:
        rep2, _ = check_step_against_oracle(z, E, "fp32")
        assert rep2["hard"] == 0


def test_fp32_tier_module_matches_cuda_core_module():
    """Ten EMA steps: search="fp32" (tensor cores) and search="fp32_simt" modules stay together (same indices unless a
    near-tie flips, codebooks within 1e-5 while they agree)."""
    from vqvae_b200 import VectorQuantizer
    torch.manual_seed(11)
    D, K = 128, 512
    a = VectorQuantizer(D, K, quantizer="ema", search="fp32", deterministic=True, cuda_graph=False).cuda().train()
    b = VectorQuantizer(D, K, quantizer="ema", search="fp32_simt", deterministic=True, cuda_graph=False).cuda().train()
    b.load_state_dict(a.state_dict())
    same = True
    for _ in range(10):
        z = torch.randn(8192, D, device="cuda")
        with torch.no_grad():
            qa, la, ia = a(z)
            qb, lb, ib = b(z)
        same &= bool(torch.equal(ia, ib))
        if not same:
            break
        assert abs(float(la) - float(lb)) <= 1e-5 * abs(float(lb))
    if same:
        assert float((a.codebook - b.codebook).abs().max() / b.codebook.abs().max()) <= 1e-5


# ------------------------------------------------------------------------------------------------ search="auto" envelope
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_auto_search_odd_dim_falls_back_to_fp32_kernel(dtype):
    """emb_dim = 12 under autocast: the tensor-core kernel cannot take it; search="auto" must pick the fp32 kernel
    instead of raising (only an explicit search="bf16" insists)."""
    from vqvae_b200 import VectorQuantizer
    from vqvae_b200._lib import VQLibError
    torch.manual_seed(1)
    D, K = 12, 40
    mod = VectorQuantizer(D, K, quantizer="ema").cuda().train()
    E0 = mod.codebook.detach().cpu().clone()
    z = torch.randn(2000, D, device="cuda").to(dtype)
    q, loss, idx = mod(z)
    ref_idx, _ = R.assign(z.cpu().float(), E0, "fp32")
    rep = R.near_tie_report(z.cpu().float(), E0, idx.cpu(), ref_idx)
    assert rep["hard"] == 0 and q.dtype == dtype
    strict = VectorQuantizer(D, K, quantizer="ema", search="bf16").cuda()
    with pytest.raises(VQLibError, match="not supported"):
        strict(z)


# ------------------------------------------------------------------------------------------------ device guard
def test_module_on_non_current_device():
    """A module on cuda:1 called while cuda:0 is current must launch on device 1's stream."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from vqvae_b200 import VectorQuantizer
    torch.manual_seed(2)
    torch.cuda.set_device(0)
    mod = VectorQuantizer(64, 128, quantizer="ema").to("cuda:1").train()
    E0 = mod.codebook.detach().cpu().clone()
    z = torch.randn(5000, 64, device="cuda:1").to(torch.bfloat16)
    q, loss, idx = mod(z)
    torch.cuda.synchronize(1)
    assert torch.cuda.current_device() == 0 and q.device.index == 1
    ref_idx, _ = R.assign(z.cpu().float(), E0, "bf16")
    assert R.near_tie_report(z.cpu().float(), E0, idx.cpu(), ref_idx, precision="bf16")["hard"] == 0
    with pytest.raises(RuntimeError, match="different devices"):
        from vqvae_b200 import ops
        ops.layout_transpose(torch.zeros(1, 8, 8, device="cuda:0"), torch.zeros(1, 8, 8, device="cuda:1"))


# ------------------------------------------------------------------------------------------------ channel-first (8f row 3)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("B,D,spatial", [(3, 64, (16, 16)), (2, 128, (5, 9, 7)), (1, 40, (33,)), (4, 256, (8, 8))])
def test_channel_first_call_no_transpose_out(B, D, spatial, dtype):
    """[B, D, *spatial] in and out (frl/models/representation.py:317-334): q comes straight from vq_gather_rows_cf, the
    pass is read-only, the backward is elementwise on the channel-first tensors.  Must equal the channel-last call on the
    permuted tensor: indices and q bit-equal, loss / gradients / codebook to the dtype's tolerance."""
    from vqvae_b200 import VectorQuantizer
    torch.manual_seed(D + len(spatial))
    K = 96
    a = VectorQuantizer(D, K, quantizer="ema", channel_first=True, deterministic=True, cuda_graph=False).cuda().train()
    b = VectorQuantizer(D, K, quantizer="ema", channel_first=False, deterministic=True, cuda_graph=False).cuda().train()
    b.load_state_dict(a.state_dict())
    z = torch.randn(B, D, *spatial, device="cuda").to(dtype)
    w = torch.randn(B, D, *spatial, device="cuda").to(dtype)
    za = z.clone().requires_grad_(True)
    perm = (0,) + tuple(range(2, z.dim())) + (1,)
    inv = (0, z.dim() - 1) + tuple(range(1, z.dim() - 1))
    zb = z.permute(*perm).contiguous().requires_grad_(True)
    qa, la, ia = a(za)
    qb, lb, ib = b(zb)
    assert qa.shape == z.shape and ia.shape == (B, *spatial) and qa.dtype == dtype
    assert torch.equal(ia, ib)
    assert torch.equal(qa, qb.permute(*inv))
    tol = 1e-5 if dtype == torch.float32 else 1e-3
    assert abs(float(la) - float(lb)) <= tol * abs(float(lb))
    ((qa * w).float().sum() + 3.0 * la).backward()
    ((qb * w.permute(*perm)).float().sum() + 3.0 * lb).backward()
    assert torch.allclose(za.grad.float(), zb.grad.permute(*inv).float(), rtol=tol, atol=tol)
    assert torch.allclose(a.codebook, b.codebook, rtol=tol, atol=1e-6)
    # the forward launches ONE layout pass (transpose in), none out
    from torch.profiler import ProfilerActivity, profile
    with torch.no_grad():
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            a(z)
            torch.cuda.synchronize()
    names = [e.name for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    assert sum("transpose_kernel" in n for n in names) == 1 and sum("gather_rows_cf_kernel" in n for n in names) == 1

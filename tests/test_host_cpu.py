"""CPU tests: the C-ABI library loads and exports every symbol include/vq_b200.h declares (no compute
calls without a GPU), the module's object contract, and the N>1 host logic over gloo."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    src = open(os.path.join(ROOT, "include", "vq_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vq_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from vqvae_b200 import _lib
    lib = _lib.load(build_if_missing=True)
    names = _header_functions()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/vq_b200.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), "ctypes binding and header disagree"
    assert lib.vq_abi_version() == _lib.ABI_VERSION == 5


def _header_prototypes():
    """name -> number of parameters of every function include/vq_b200.h declares (comments stripped)."""
    src = open(os.path.join(ROOT, "include", "vq_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = {}
    for m in re.finditer(r"\b(vq_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    return out


def test_ctypes_signatures_have_the_header_arity():
    """A prototype that gains or loses a parameter must change the binding too (ABI drift shows up here, not as a
    mis-aligned argument on the GPU)."""
    from vqvae_b200 import _lib
    protos = _header_prototypes()
    assert set(protos) == set(_lib.SIGNATURES)
    for name, (_, argtypes) in _lib.SIGNATURES.items():
        assert len(argtypes) == protos[name], f"{name}: header has {protos[name]} parameters, ctypes binding {len(argtypes)}"


def test_c_host_example_compiles_and_links(tmp_path):
    """examples/host_c_example.c drives one step through the C ABI from plain C99: the header must be valid C and every
    entry point it uses must resolve against the built library (running it needs a GPU)."""
    import shutil
    import subprocess
    from vqvae_b200 import _lib
    _lib.load(build_if_missing=True)
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    if shutil.which("gcc") is None or not os.path.exists(os.path.join(cuda, "include", "cuda_runtime_api.h")):
        pytest.skip("gcc or the CUDA runtime headers are not here")
    hdr = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c",
                          os.path.join(ROOT, "include", "vq_b200.h")], capture_output=True, text=True)
    assert hdr.returncode == 0, hdr.stderr
    exe = str(tmp_path / "host_c_example")
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-O1", os.path.join(ROOT, "examples", "host_c_example.c"),
           "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(cuda, "include"),
           "-L" + os.path.join(ROOT, "vqvae_b200"), "-lvq_b200", "-L" + os.path.join(cuda, "lib64"), "-lcudart", "-lm",
           "-Wl,-rpath," + os.path.join(ROOT, "vqvae_b200"), "-o", exe]
    out = subprocess.run(cmd, capture_output=True, text=True)
    assert out.returncode == 0, out.stderr[-2000:]
    assert os.path.exists(exe)


def test_host_only_entry_points():
    from vqvae_b200 import _lib
    lib = _lib.load(build_if_missing=True)
    # K_pad norms + K_pad * 8 floats of operand image (include/vq_b200.h)
    assert lib.vq_norm_elems(1) == 256 * 9 and lib.vq_norm_elems(256) == 256 * 9 and lib.vq_norm_elems(8192) == 8192 * 9
    assert lib.vq_norm_elems(1000) == 1024 * 9
    out = ctypes.c_size_t(0)
    assert lib.vq_workspace_bytes(524288, 8192, 256, 0, 0, ctypes.byref(out)) == 0       # bf16 latents, bf16 search
    small = out.value
    assert lib.vq_workspace_bytes(524288, 8192, 256, 2, 2, ctypes.byref(out)) == 0       # fp16 latents, fp16 search: native
    assert out.value == small
    assert lib.vq_workspace_bytes(524288, 8192, 256, 1, 0, ctypes.byref(out)) == 0
    assert out.value >= small + 524288 * 256 * 2          # fp32 latents under the bf16 search: one bf16 staging plane
    assert lib.vq_workspace_bytes(524288, 8192, 256, 1, 1, ctypes.byref(out)) == 0
    assert out.value >= small + 524288 * 256 * 4          # split fp32 tier: hi + lo planes
    assert lib.vq_workspace_bytes(524288, 8192, 264, 1, 1, ctypes.byref(out)) == 0       # D > 256: CUDA-core kernel, no staging
    assert lib.vq_workspace_bytes(-1, 8192, 256, 0, 0, ctypes.byref(out)) == 1   # VQ_ERR_ARG
    assert b"bad arguments" in lib.vq_last_error()
    assert lib.vq_shadow_elems(1000, 128, 0) == 1000 * 128 and lib.vq_shadow_elems(1000, 128, 2) == 1000 * 128
    assert lib.vq_shadow_elems(1000, 128, 1) == 2 * 1024 * 128


def test_module_contract_on_cpu():
    from vqvae_b200 import CodebookManager, VectorQuantizer, extract_code_usage_from_state

    class Outer(torch.nn.Module):
        def __init__(self, q):
            super().__init__()
            self.quant = q

    st = Outer(VectorQuantizer(emb_dim=128, codebook_size=256, beta=0.25, quantizer="st", ema_decay=0.99, ema_eps=1e-5))
    ema = Outer(VectorQuantizer(emb_dim=128, codebook_size=256, beta=0.05, quantizer="ema"))
    # scripts/train_vqvae.py:221-228 -- parameter split on "quant.codebook"
    assert [n for n, _ in st.named_parameters() if n.endswith("quant.codebook")] == ["quant.codebook"]
    assert list(ema.named_parameters()) == []
    # scripts/export_codebook.py:104 ; scripts/train_vqvae.py:197
    assert st.quant.codebook.shape == (256, 128) and st.quant.codebook_size == 256 and st.quant.emb_dim == 128
    mgr = CodebookManager(num_codes=ema.quant.codebook_size, code_dim=ema.quant.emb_dim)
    ema.quant.attach_codebook_manager(mgr)
    assert mgr.summary()["n_dead"] == 0
    sd = ema.state_dict()
    for k in ("quant.codebook", "quant.ema_cnt", "quant.ema_sum", "quant.usage_total", "quant.steps_since_used"):
        assert k in sd
    assert extract_code_usage_from_state(sd).shape == (256,)
    ema2 = Outer(VectorQuantizer(128, 256, quantizer="ema"))
    ema2.load_state_dict(sd)
    assert torch.equal(ema2.quant.codebook, ema.quant.codebook)
    with pytest.raises(RuntimeError, match="CUDA"):
        ema.quant(torch.randn(4, 128))
    with pytest.raises(ValueError):
        VectorQuantizer(8, 8, quantizer="gumbel")


def test_split_quantizer_and_host_entry_contract_on_cpu():
    """Two-codebook module (BASELINE config 3) and the host-buffer entry: object contract without a GPU."""
    from vqvae_b200 import SplitQuantizer, VectorQuantizer, extract_code_usage_from_state
    sq = SplitQuantizer(type_dim=48, phase_dim=8, codebook_size=64, quantizer="ema")        # frl_repr_model_v1.yaml:14-15
    assert sq.quant_type.codebook.shape == (64, 48) and sq.quant_phase.codebook.shape == (64, 8)
    assert sq.quant_type.channel_first and sq.quant_phase.channel_first
    sd = sq.state_dict()
    for k in ("quant_type.codebook", "quant_phase.codebook", "quant_type.usage_total", "quant_phase.ema_sum"):
        assert k in sd
    assert extract_code_usage_from_state(sd, prefix="quant_phase.").shape == (64,)
    sq2 = SplitQuantizer(48, 8, 64, quantizer="ema")
    sq2.load_state_dict(sd)
    assert torch.equal(sq2.quant_type.codebook, sq.quant_type.codebook)
    with pytest.raises(RuntimeError, match="CUDA"):
        sq(torch.randn(1, 48, 4, 4), torch.randn(1, 8, 2, 4, 4))
    # the pipelined host entry refuses to run without a device: no CPU fallback anywhere on the path
    vq = VectorQuantizer(16, 32, quantizer="ema")
    with pytest.raises(RuntimeError, match="CUDA"):
        vq.forward_host(torch.randn(10, 16))
    with pytest.raises(ValueError):
        vq.forward_host(torch.randn(10, 15))
    assert vq.cuda_graph and vq.graph_launches == 0 and not vq._graphs


def test_shard_bounds_cover_rows():
    from vqvae_b200.parallel import shard_bounds
    for n in (0, 1, 7, 4096, 4099):
        for w in (1, 2, 3, 8):
            b = [shard_bounds(n, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            assert max(e - s for s, e in b) - min(e - s for s, e in b) <= 1


def _dp_worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import vq_reference as R
        from vqvae_b200.parallel import allreduce_code_stats, assert_replicas_identical, shard_bounds
        g = torch.Generator().manual_seed(11)
        K, D, N = 16, 8, 1001
        E = torch.randn(K, D, generator=g)
        z = torch.randn(N, D, generator=g)
        a, b = shard_bounds(N, rank, world)
        idx, _ = R.assign(z[a:b], E, "fp32")
        cnt, sums = R.code_stats(z[a:b], idx, K)
        stats = torch.cat([sums.reshape(-1), cnt])
        allreduce_code_stats(stats)
        En, n, m = R.ema_update(E, torch.ones(K), E.clone(), stats[K * D:], stats[:K * D].view(K, D), 0.99, 1e-5)
        assert_replicas_identical(En)
        torch.save({"E": En, "cnt": stats[K * D:].clone()}, os.path.join(tmp, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_code_stat_allreduce_two_ranks_gloo(tmp_path):
    """world_size 2 over gloo: sharded statistics + all-reduce == the single-process full batch."""
    from oracle import vq_reference as R
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_dp_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0 = torch.load(os.path.join(tmp_path, "r0.pt"))
    r1 = torch.load(os.path.join(tmp_path, "r1.pt"))
    assert torch.equal(r0["E"], r1["E"])                      # replicas bit-identical, no broadcast needed
    g = torch.Generator().manual_seed(11)
    K, D, N = 16, 8, 1001
    E = torch.randn(K, D, generator=g)
    z = torch.randn(N, D, generator=g)
    idx, _ = R.assign(z, E, "fp32")
    cnt, sums = R.code_stats(z, idx, K)
    En, _, _ = R.ema_update(E, torch.ones(K), E.clone(), cnt, sums, 0.99, 1e-5)
    assert torch.equal(r0["cnt"], cnt)
    np.testing.assert_allclose(r0["E"].numpy(), En.numpy(), rtol=1e-5, atol=1e-6)

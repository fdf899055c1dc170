"""The two reference entry points re-created in scripts/ (reference scripts/train_vqvae.py,
scripts/export_codebook.py).  CPU part: CLI surface, annealer, export from a state dict.  GPU part: a short run."""
import json
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "scripts"))

import export_codebook  # noqa: E402
import train_vqvae  # noqa: E402
from vqvae_model import VQVAE, SyntheticPatches  # noqa: E402


def test_cli_keeps_reference_flag_names_and_defaults(tmp_path):
    a = train_vqvae.parse_args(["--run_dir", str(tmp_path)])
    # scripts/train_vqvae.py:410-414,433-436
    assert (a.codebook_size, a.emb_dim, a.beta, a.quantizer, a.ema_decay, a.ema_eps) == (256, 128, 0.25, "st", 0.99, 1e-5)
    cfg = tmp_path / "c.yaml"
    cfg.write_text("train_vqvae:\n  codebook_size: 512\n  quantizer: ema\n  beta: 0.05\n")
    a = train_vqvae.parse_args(["--run_dir", str(tmp_path), "--config", str(cfg), "--beta", "0.5"])
    assert (a.codebook_size, a.quantizer, a.beta) == (512, "ema", 0.5)          # YAML sets, CLI wins


def test_lambda_vq_annealer_profiles():
    base = {"vq": 1.0, "cont": 1.0, "cat": 1.0, "canopy": 1.0}
    assert train_vqvae.load_scheduler(base, {})(step=7)["vq"] == 1.0
    f = train_vqvae.load_scheduler(base, dict(anneal_vq_enable=True, anneal_vq_schedule="warmup_hold_decay",
                                               anneal_vq_floor=0.0, anneal_vq_ceil=0.1, anneal_vq_warmup=100,
                                               anneal_vq_hold=50, anneal_vq_decay=50, anneal_vq_final=0.08))
    assert f(step=0)["vq"] == 0.0 and abs(f(step=50)["vq"] - 0.05) < 1e-9 and f(step=120)["vq"] == 0.1
    assert abs(f(step=175)["vq"] - 0.09) < 1e-9 and f(step=10 ** 6)["vq"] == 0.08
    g = train_vqvae.load_scheduler(base, dict(anneal_vq_enable=True, anneal_vq_schedule="stepwise",
                                               anneal_vq_milestones=["10:0.01", "100:0.1"]))
    assert (g(step=5)["vq"], g(step=10)["vq"], g(step=500)["vq"]) == (0.0, 0.01, 0.1)


def test_model_contract_on_cpu_and_param_split():
    m = VQVAE(cont_dim=8, cat_vocab_sizes={"lc": 5}, emb_dim=16, codebook_size=32, hidden=16, quantizer="st")
    names = [n for n, _ in m.named_parameters()]
    assert "quant.codebook" in names                                          # scripts/train_vqvae.py:221-228
    m2 = VQVAE(cont_dim=8, cat_vocab_sizes={"lc": 5}, emb_dim=16, codebook_size=32, hidden=16, quantizer="ema")
    assert "quant.codebook" not in [n for n, _ in m2.named_parameters()] and "quant.codebook" in m2.state_dict()
    with pytest.raises(RuntimeError):                                         # no CPU fallback
        m(SyntheticPatches(cont_dim=8, cat_vocab_sizes={"lc": 5}, T=2, H=4, W=4).batch(2))


def test_export_codebook_from_cpu_checkpoint(tmp_path):
    args = dict(cont_dim=8, time_steps=3, patch=4, seed=0, emb_dim=16, codebook_size=32, beta=0.25, hidden=16,
                quantizer="ema", ema_decay=0.99, ema_eps=1e-5)
    m = VQVAE(cont_dim=8, cat_vocab_sizes={"lc": 12}, emb_dim=16, codebook_size=32, hidden=16, quantizer="ema")
    m.quant.usage_total.copy_(torch.arange(32) * 3)
    ck = tmp_path / "ckpt.pt"
    torch.save({"model": m.state_dict(), "opt": {}, "args": args, "step": 1}, ck)
    assert export_codebook.main(["--ckpt", str(ck), "--out", str(tmp_path / "dec"), "--csv"]) == 0
    z = np.load(tmp_path / "dec.npz")
    K, T = 32, 3
    assert z["cont_KT"].shape == (K * T, 8) and z["cats_KT"].shape == (K * T, 1)
    assert z["code_id"].tolist() == np.repeat(np.arange(K), T).tolist() and z["year"].shape == (K * T,)
    assert z["codes_K3"].shape == (K, 3) and z["codes_K3"][:, 1].tolist() == (np.arange(K) * 3.0).tolist()
    assert json.loads(str(z["meta"]))["K"] == K
    assert (tmp_path / "dec_codes_K3.csv").exists() and (tmp_path / "dec_cont_KT.csv").exists()


@pytest.mark.gpu
@pytest.mark.parametrize("quantizer,bf16", [("ema", True), ("st", False)])
def test_short_training_run_and_export(tmp_path, quantizer, bf16):
    argv = ["--run_dir", str(tmp_path), "--batch_size", "8", "--steps_per_epoch", "30", "--epochs", "2", "--cont_dim",
            "16", "--time_steps", "2", "--patch", "16", "--emb_dim", "32", "--codebook_size", "64", "--hidden", "32",
            "--quantizer", quantizer, "--log_every", "10", "--lr", "2e-3", "--data_init"] + (["--bf16"] if bf16 else [])
    out = train_vqvae.train(train_vqvae.parse_args(argv))
    assert out["step"] == 60 and all(np.isfinite(out["losses"])) and out["losses"][-1] < out["losses"][0]
    assert out["summary"]["used"] > 1 and out["summary"]["perplexity"] > 1.0
    ck = torch.load(tmp_path / "ckpt_epoch002.pt", map_location="cpu", weights_only=False)
    assert set(ck) == {"model", "opt", "args", "step"} and "quant.codebook" in ck["model"]
    assert int(ck["model"]["quant.usage_total"].sum()) == 60 * 8 * 2 * 16 * 16
    assert export_codebook.main(["--ckpt", str(tmp_path / "ckpt_epoch002.pt"), "--out", str(tmp_path / "dec")]) == 0
    assert np.load(tmp_path / "dec.npz")["codes_K3"][:, 1].sum() == 60 * 8 * 2 * 16 * 16


class _OracleOnCpu(torch.nn.Module):
    """The CPU oracle standing in for `model.quant` inside the GPU model: latents go to the host (a differentiable copy),
    the oracle runs there, results come back -- so that the SAME training step can be run with either quantiser."""

    def __init__(self, ref):
        super().__init__()
        self.ref = ref

    @property
    def perplexity(self):
        return self.ref.last_perplexity[0]

    def forward(self, z):
        q, loss, idx = self.ref(z.float().cpu())
        return q.to(z.device), loss.to(z.device), idx.to(z.device)


@pytest.mark.gpu
@pytest.mark.parametrize("quantizer", ["ema", "st"])
def test_training_steps_native_quantiser_vs_oracle_swapped_in(quantizer):
    """the loop of scripts/train_vqvae.py:261-347 (model(batch) -> losses -> backward -> AdamW with
    the codebook in its own weight_decay = 0 group) for 10 fp32 deterministic steps, once around the native quantiser and
    once with the oracle swapped in for `model.quant`: per-step vq_loss, perplexity and the final codebook must agree."""
    from oracle import vq_reference as R
    import torch.nn.functional as F
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    K, D = 64, 32

    def build(seed):
        torch.manual_seed(seed)
        return VQVAE(cont_dim=16, cat_vocab_sizes={"lc": 12}, emb_dim=D, codebook_size=K, hidden=32, quantizer=quantizer,
                     search="fp32", deterministic=True, cuda_graph=False)
    native = build(0).cuda().train()
    swapped = build(0)
    ref = R.VectorQuantizerOracle(D, K, beta=0.25, quantizer=quantizer).train()
    ref.load_state_dict({k: v.clone() for k, v in swapped.quant.state_dict().items()}, strict=False)
    swapped.quant = None
    swapped = swapped.cuda().train()
    swapped.quant = _OracleOnCpu(ref)                    # stays on the CPU
    ds = SyntheticPatches(cont_dim=16, T=2, H=16, W=16, seed=3, device="cuda")

    def make_opt(m, cb):
        others = [p for n, p in m.named_parameters() if "quant" not in n]
        groups = [{"params": others, "weight_decay": 1e-4}]
        if cb is not None:
            groups.append({"params": [cb], "weight_decay": 0.0})
        return torch.optim.AdamW(groups, lr=1e-3, betas=(0.9, 0.95))
    opt_n = make_opt(native, native.quant.codebook if quantizer == "st" else None)
    opt_s = make_opt(swapped, ref.codebook if quantizer == "st" else None)
    for step in range(10):
        batch = ds.batch(4)
        vals = []
        for m, opt in ((native, opt_n), (swapped, opt_s)):
            opt.zero_grad(set_to_none=True)
            cont_pred, cat_logits, canopy_pred, vq_loss, pplx = m(batch)
            loss = F.mse_loss(cont_pred, batch["cont"]) + vq_loss + F.mse_loss(canopy_pred, batch["canopy"])
            for j, name in enumerate(ds.cat_names):
                loss = loss + F.cross_entropy(cat_logits[name].reshape(-1, cat_logits[name].size(-1)),
                                              batch["cat_target"][..., j].reshape(-1))
            loss.backward()
            opt.step()
            vals.append((float(vq_loss), float(pplx), float(loss)))
        (vn, pn, ln), (vs, ps, ls) = vals
        assert abs(vn - vs) <= 5e-4 * abs(vs), (step, vn, vs)
        assert abs(pn - ps) <= 5e-4 * ps, (step, pn, ps)
        assert abs(ln - ls) <= 5e-4 * abs(ls), (step, ln, ls)
    En, Es = native.quant.codebook.detach().cpu(), ref.codebook.detach()
    assert float((En - Es).abs().max() / Es.abs().max()) <= 2e-3
    used_n, used_s = native.quant.usage_total.cpu(), ref.usage_total
    assert int(used_n.sum()) == int(used_s.sum()) and int((used_n != used_s).sum()) <= 2

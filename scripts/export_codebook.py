#!/usr/bin/env python
"""scripts/export_codebook.py -- the reference's codebook read-out, re-created for checkpoints of this repo.

Mirrors the reference CLI (scripts/export_codebook.py:76-183): `--zarr --ckpt --out [--csv]`, runs entirely on CPU
(:27), decodes the codebook `model.quant.codebook [K, D]` (:104,114,138) through the decoder as `[K, T, D]`
sequences (:93), reads per-code usage from the plain state dict with `extract_code_usage_from_state` (:124-125)
and writes ONE NPZ with the reference's keys: cont_KT [K*T, C_cont], cats_KT [K*T, C_cat], code_id [K*T],
year [K*T], codes_K3 [K, 3] = (code_id, code_usage, canopy), meta (JSON).  `--csv` adds the three CSV files.

No kernel runs here: only attribute access / state_dict loading on CPU, which the quantiser supports by contract
(SURVEY.md 8b "Device / dtype").  The decoder is the torch conv stack of scripts/vqvae_model.py; each code is
decoded as a constant 1x1-pixel sequence of length T.  `--zarr` is accepted for CLI compatibility ("synthetic").
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))

from vqvae_model import VQVAE, SyntheticPatches  # noqa: E402
from vqvae_b200 import extract_code_usage_from_state  # noqa: E402


def load_model_and_ds(zarr: str, ckpt_path: str):
    """(reference: vqvae.postprocess.load_model_and_ds, absent) rebuild the model on CPU from the checkpoint's args."""
    ckpt = torch.load(ckpt_path, map_location="cpu", weights_only=False)
    a = ckpt["args"]
    ds = SyntheticPatches(cont_dim=a.get("cont_dim", 64), T=a.get("time_steps", 5), H=a.get("patch", 32),
                          W=a.get("patch", 32), seed=a.get("seed", 0))
    model = VQVAE(cont_dim=ds.cont_dim, cat_vocab_sizes=ds.cat_vocab_sizes, emb_dim=a["emb_dim"],
                  codebook_size=a["codebook_size"], beta=a["beta"], hidden=a["hidden"], quantizer=a["quantizer"],
                  ema_decay=a["ema_decay"], ema_eps=a["ema_eps"])
    model.load_state_dict(ckpt["model"])
    return model.eval(), ds, ckpt


@torch.no_grad()
def decode_codebook_sequences(model: VQVAE, T: int):
    """[K, D] codebook -> decoder heads on [K, T, 1, 1, D]: cont [K, T, C], cat logits {name: [K, T, n]}, canopy [K]."""
    E = model.quant.codebook.detach().float()
    K, D = E.shape
    zq = E[:, None, None, None, :].expand(K, T, 1, 1, D).contiguous()
    cont, cats, canopy = model.float().decode(zq)
    return cont[:, :, 0, 0], {k: v[:, :, 0, 0] for k, v in cats.items()}, canopy


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="Decode the codebook and export arrays (NPZ, optional CSV).")
    ap.add_argument("--zarr", default="synthetic")
    ap.add_argument("--ckpt", required=True)
    ap.add_argument("--out", required=True, help="output prefix (without extension)")
    ap.add_argument("--csv", action="store_true")
    args = ap.parse_args(argv)
    out_prefix = Path(args.out)
    out_prefix.parent.mkdir(parents=True, exist_ok=True)

    model, ds, ckpt = load_model_and_ds(args.zarr, args.ckpt)
    K = int(model.quant.codebook.shape[0])
    T = ds.T
    years = np.arange(2000, 2000 + T, dtype=np.int32)             # synthetic time axis
    cont, cat_logits, canopy = decode_codebook_sequences(model, T)
    cont_KT = cont.reshape(K * T, -1).numpy().astype(np.float32)
    cats_KT = (torch.stack([cat_logits[n].argmax(-1) for n in ds.cat_names], dim=-1).reshape(K * T, -1)
               .numpy().astype(np.float32)) if ds.cat_names else np.zeros((K * T, 0), np.float32)
    code_id = np.repeat(np.arange(K, dtype=np.int32), T)
    year = np.tile(years, K)
    usage = extract_code_usage_from_state(ckpt.get("model", ckpt)).numpy()
    summary_K3 = np.stack([np.arange(K, dtype=np.float32), usage.astype(np.float32),
                           canopy.numpy().astype(np.float32)], axis=1)
    meta = {"cont_names": list(ds.cont_names), "cat_names": list(ds.cat_names), "T": T, "K": K,
            "shapes": {"cont_KT": list(cont_KT.shape), "cats_KT": list(cats_KT.shape), "code_id": list(code_id.shape),
                       "year": list(year.shape), "codes_K3": list(summary_K3.shape)},
            "notes": "synthetic data; cats_KT holds arg-max class ids; canopy is the raw head value"}
    npz_path = out_prefix.with_suffix(".npz")
    np.savez_compressed(npz_path, cont_KT=cont_KT, cats_KT=cats_KT, code_id=code_id, year=year, codes_K3=summary_K3,
                        meta=json.dumps(meta))
    if args.csv:
        import pandas as pd
        df = pd.DataFrame(cont_KT, columns=list(ds.cont_names))
        df.insert(0, "year", year.astype(int)); df.insert(0, "code_id", code_id.astype(int))
        df.to_csv(out_prefix.with_name(out_prefix.name + "_cont_KT.csv"), index=False)
        df = pd.DataFrame(cats_KT, columns=list(ds.cat_names))
        df.insert(0, "year", year.astype(int)); df.insert(0, "code_id", code_id.astype(int))
        df.to_csv(out_prefix.with_name(out_prefix.name + "_cats_KT.csv"), index=False)
        pd.DataFrame(summary_K3, columns=["code_id", "code_usage", "canopy"]).astype({"code_id": int}).to_csv(
            out_prefix.with_name(out_prefix.name + "_codes_K3.csv"), index=False)
    print(f"Wrote {npz_path}")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())

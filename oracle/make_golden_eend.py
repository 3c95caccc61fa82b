"""Freeze outputs of the reference's second model class, ``EendEdaModel(encoder_type="transformer")``
(models.py:465-700), as a golden fixture.  TEST INFRASTRUCTURE; build container only
(``python -m oracle.make_golden_eend``).  Weights: the seeded TransformerEdaModel dict with its
keys renamed to this class's names (encoder -> linear, encoder_norm -> linear_norm,
transformer_encoder -> encoder)."""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

REF_ROOT = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")


def rename(sd):
    out = {}
    for k, v in sd.items():
        if k.startswith("encoder_norm."):
            k = "linear_norm." + k[len("encoder_norm."):]
        elif k.startswith("encoder."):
            k = "linear." + k[len("encoder."):]
        elif k.startswith("transformer_encoder."):
            k = "encoder." + k[len("transformer_encoder."):]
        out[k] = v
    return out


def main():
    sys.path.insert(0, REF_ROOT)
    sys.path.insert(0, os.path.dirname(HERE))
    from oracle import librosa_stub
    from oracle.weights import EdaConfig, synth_state_dict, synth_features, synth_labels
    librosa_stub.install()
    from speaker_diarization.eend_eda.models import EendEdaModel
    name, L, lens, S, wseed, dseed = "eendeda_l2", 2, [300, 211, 64], 2, 9, 41
    cfg = EdaConfig(n_layers=L, n_speakers=S)
    model = EendEdaModel(S, cfg.in_size, cfg.n_heads, cfg.n_units, L, encoder_type="transformer")
    model.load_state_dict(rename(synth_state_dict(cfg, wseed)), strict=True)
    model.eval()
    src, ts = synth_features(lens, cfg.in_size, dseed), synth_labels(lens, S, dseed)
    out = {}
    with torch.no_grad():
        torch.manual_seed(1000 + dseed)
        for b, y in enumerate(model.infer(src, infer_num_speakers=2)):
            out[f"post2_{b}"] = y.numpy()
        torch.manual_seed(1000 + dseed)
        out["nspk_thr05"] = np.array([y.shape[1] for y in model.infer(src, attractor_threshold=0.5)], dtype=np.int64)
        torch.manual_seed(2000 + dseed)
        loss, stats = model(src, ts)
        out["loss"] = np.float32(loss.item())
        out["der"] = np.float64(float(stats["der"]))
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
    with open(os.path.join(GOLD, "manifest.json")) as f:
        manifest = json.load(f)
    manifest["cases"][name] = dict(kind="eend", n_layers=L, lengths=lens, n_speakers=S, weight_seed=wseed, data_seed=dseed)
    with open(os.path.join(GOLD, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)
    print(name, float(out["loss"]), out["nspk_thr05"], out["der"])


if __name__ == "__main__":
    main()

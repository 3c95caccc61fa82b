"""Freeze outputs of the UNMODIFIED reference ``EendEdaModel(encoder_type="conformer")`` (models.py:465-700 +
torchaudio.models.Conformer) as golden fixtures.  TEST INFRASTRUCTURE; build container only
(``python -m oracle.make_golden_conformer``).  Weights come from oracle.weights.synth_conformer_state_dict."""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

REF_ROOT = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")

CASES = {
    # name: (n_layers, lengths, n_speakers, weight_seed, data_seed)
    "conformer_l1_ragged": (1, [130, 77, 128], 2, 3, 51),
    "conformer_l2_t500": (2, [500, 377, 41], 2, 4, 52),
    "conformer_l4_s3": (4, [300, 257], 3, 5, 53),
}


def main():
    sys.path.insert(0, REF_ROOT)
    sys.path.insert(0, os.path.dirname(HERE))
    from oracle import librosa_stub
    from oracle.weights import synth_conformer_state_dict, synth_features, synth_labels
    librosa_stub.install()
    from speaker_diarization.eend_eda.models import EendEdaModel
    with open(os.path.join(GOLD, "manifest.json")) as f:
        manifest = json.load(f)
    for name, (L, lens, S, wseed, dseed) in CASES.items():
        model = EendEdaModel(S, 345, 4, 256, L, encoder_type="conformer")
        sd = synth_conformer_state_dict(L, wseed, template=model.state_dict())
        model.load_state_dict(sd, strict=True)
        model.eval()
        src, ts = synth_features(lens, 345, dseed), synth_labels(lens, S, dseed)
        out = {}
        with torch.no_grad():
            torch.manual_seed(1000 + dseed)
            _, emb, _ = model.forward_embedding(src)
            out["emb_sample"] = emb[:, ::7, ::5].numpy()
            torch.manual_seed(1000 + dseed)
            for b, y in enumerate(model.infer(src, attractor_threshold=-1.0)):
                out[f"post_all_{b}"] = y.numpy()
            torch.manual_seed(2000 + dseed)
            loss, stats = model(src, ts)
            out["loss"] = np.float32(loss.item())
            out["der"] = np.float64(float(stats["der"]))
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
        manifest["cases"][name] = dict(kind="conformer", n_layers=L, lengths=lens, n_speakers=S, weight_seed=wseed,
                                       data_seed=dseed)
        print(name, float(out["loss"]), out["der"])
    with open(os.path.join(GOLD, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()

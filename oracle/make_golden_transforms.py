"""Freeze the UNMODIFIED reference ``feature.transform`` for the variants other than 'logmel23_mn' that run under the
pinned librosa 0.10.2 (feature.py:43-47 None / 'log', :105-116 'logmel23_espnet') into tests/golden/fe_transforms.npz,
and record that the remaining variants raise TypeError there (positional librosa.filters.mel call, feature.py:53,61,79,97).

TEST INFRASTRUCTURE (see oracle/__init__.py); build container only: ``python -m oracle.make_golden_transforms``."""
import json
import os
import sys

import numpy as np

REF_ROOT = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")

CASES = {
    # name: (sample_rate, frame_size, frame_shift, n_samples, seed, dtype, transform_type)
    "tf_none_8k": (8000, 200, 80, 8000 * 2 + 13, 61, "float64", None),
    "tf_log_8k": (8000, 200, 80, 8000 * 2 + 13, 61, "float64", "log"),
    "tf_log_8k_f32": (8000, 200, 80, 8000 * 2, 62, "float32", "log"),
    "tf_espnet_8k": (8000, 200, 80, 8000 * 3 + 7, 63, "float64", "logmel23_espnet"),
    "tf_espnet_16k": (16000, 400, 160, 16000 * 2 + 301, 64, "float64", "logmel23_espnet"),
}


def main():
    sys.path.insert(0, REF_ROOT)
    sys.path.insert(0, os.path.dirname(HERE))
    from oracle import librosa_stub, feature_oracle as fo
    librosa_stub.install()
    from speaker_diarization import feature as F
    out, meta = {}, {}
    for name, (sr, fs, hop, n, seed, dt, tt) in CASES.items():
        x, _ = fo.synth_audio(n, sr, 2, seed, dtype=np.dtype(dt).type)
        Y = F.transform(F.stft(x, fs, hop), tt, sample_rate=sr)
        feats = np.ascontiguousarray(F.splice(Y, 7)[::10])
        out[name + "/Y"], out[name + "/feats"] = Y, feats
        meta[name] = dict(sample_rate=sr, frame_size=fs, frame_shift=hop, n_samples=n, seed=seed, dtype=dt, transform_type=tt,
                          input_dim=int(F.get_input_dim(fs, 7, tt)) if tt else None)
        print(name, Y.shape, Y.dtype, feats.shape)
    raised = {}
    x, _ = fo.synth_audio(4000, 8000, 2, 1, dtype=np.float64)
    Y = F.stft(x, 200, 80)
    for tt in ("logmel", "logmel23", "logmel23_swn", "logmel23_mvn"):
        try:
            F.transform(Y, tt, sample_rate=8000)
            raised[tt] = None
        except Exception as e:          # noqa: BLE001
            raised[tt] = type(e).__name__
    print("positional-mel variants:", raised)
    assert all(v == "TypeError" for v in raised.values())
    meta["_positional_mel_variants_raise"] = raised
    out["meta"] = np.array(json.dumps(meta))
    np.savez_compressed(os.path.join(GOLD, "fe_transforms.npz"), **out)


if __name__ == "__main__":
    main()

"""Freeze outputs of the UNMODIFIED reference as golden fixtures under tests/golden/.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Run in the build container only
(``python -m oracle.make_golden``): it imports the reference from /root/reference,
which does not exist on the GPU box.  Inputs and weights are not stored — they are
regenerated from seeds by ``oracle/weights.py`` / ``oracle/feature_oracle.synth_audio``;
only the reference's outputs (and checksums of the regenerated inputs) are stored.
"""
from __future__ import annotations

import hashlib
import json
import os
import sys

import numpy as np
import torch

REF_ROOT = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")

MODEL_CASES = {
    # name: (n_layers, lengths, n_speakers(for forward), weight_seed, data_seed, silent speakers)
    "c1_l2_ragged": (2, [500, 413, 257, 500], 2, 777, 1, ()),
    "c2_l4_t500": (4, [500, 500, 377], 2, 777, 2, ()),
    "c5_l4_s3_t2000": (4, [2000, 1531], 3, 0, 3, ()),
    "c3_l4_s4_silent": (4, [800, 640], 4, 1, 4, (1, 3)),
    "tiny_l1": (1, [33, 1, 64], 2, 5, 5, ()),
}
PIT_CASES = {
    # name: (B, T, S, seed, silent)
    "pit_s2": (8, 120, 2, 11, ()),
    "pit_s3_tie": (6, 97, 3, 12, (0, 2)),
    "pit_s4_tie": (5, 150, 4, 13, (1, 2)),
    "pit_s6": (3, 64, 6, 14, (5,)),
}
FEATURE_CASES = {
    # name: (sample_rate, frame_size, frame_shift, n_samples, n_speakers, seed, dtype)
    "fe_8k_f64": (8000, 200, 80, 8000 * 6 + 37, 2, 21, "float64"),
    "fe_8k_even_f32": (8000, 200, 80, 8000 * 4, 3, 22, "float32"),
    "fe_16k_f64": (16000, 400, 160, 16000 * 5 + 911, 4, 23, "float64"),
}


def sha(t) -> str:
    a = t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def main() -> None:
    sys.path.insert(0, REF_ROOT)
    sys.path.insert(0, os.path.dirname(HERE))
    from oracle import librosa_stub, feature_oracle as fo
    from oracle.weights import EdaConfig, synth_state_dict, synth_features, synth_labels
    librosa_stub.install()
    from speaker_diarization.eend_eda.models import TransformerEdaModel
    from speaker_diarization import feature as ref_feature

    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(os.cpu_count() or 1)
    manifest = {"torch": torch.__version__, "cases": {}}

    for name, (L, lens, S, wseed, dseed, silent) in MODEL_CASES.items():
        cfg = EdaConfig(n_layers=L, n_speakers=S)
        sd = synth_state_dict(cfg, wseed)
        model = TransformerEdaModel(S, cfg.in_size, cfg.n_heads, cfg.n_units, L)
        model.load_state_dict(sd, strict=True)
        model.eval()
        src = synth_features(lens, cfg.in_size, dseed)
        ts = synth_labels(lens, S, dseed, silent=silent)
        out = {}
        with torch.no_grad():
            torch.manual_seed(1000 + dseed)            # all 14 columns: nothing is < -1
            ys = model.infer(src, attractor_threshold=-1.0)
            for b, y in enumerate(ys):
                out[f"post_all_{b}"] = y.numpy()
            torch.manual_seed(1000 + dseed)
            ys = model.infer(src, attractor_threshold=0.5)
            out["nspk_thr05"] = np.array([y.shape[1] for y in ys], dtype=np.int64)
            torch.manual_seed(1000 + dseed)            # probs/attractors of that same draw
            emb_s, emb, ilens = model.forward_embedding(src)
            att, probs = model.eda(emb_s, torch.LongTensor(ilens), torch.zeros(len(lens), 15, cfg.n_units))
            out["emb_sample"] = emb[:, ::17, ::5].numpy()
            out["attractors15"] = att.numpy()
            out["probs15"] = torch.stack(probs).numpy()
            torch.manual_seed(2000 + dseed)
            loss, stats = model(src, ts)
            out["loss"] = np.float32(loss.item())
            for k, v in stats.items():
                out["stat_" + k] = np.float64(float(v))
            torch.manual_seed(2000 + dseed)            # replay to expose pred / PIT internals
            emb_s, emb, ilens = model.forward_embedding(src)
            tsp = torch.nn.utils.rnn.pad_sequence(ts, padding_value=-1, batch_first=True)
            att, probs = model.eda(emb_s, torch.LongTensor(ilens), torch.zeros(len(lens), S + 1, cfg.n_units))
            pred = torch.bmm(emb, att[:, :-1].permute(0, 2, 1))
            l_pit, idx, _, lab = model.pit_loss(pred, tsp, ilens)
            out["pred"] = pred.numpy()
            out["pit_min_idx"] = idx.numpy()
            out["label_permute"] = lab.numpy().astype(np.int8)
            out["der_counts"] = np.array(model.calc_diarization_error(pred, lab, ilens), dtype=np.float64)
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
        manifest["cases"][name] = dict(kind="model", n_layers=L, lengths=lens, n_speakers=S,
                                       weight_seed=wseed, data_seed=dseed, silent=list(silent),
                                       sha_w0=sha(sd["encoder.weight"]), sha_x0=sha(src[0]), sha_t0=sha(ts[0]))
        print(name, "loss", float(out["loss"]), "nspk", out["nspk_thr05"])

    ref = TransformerEdaModel(2, 345, 4, 256, 1)
    for name, (B, T, S, seed, silent) in PIT_CASES.items():
        g = torch.Generator().manual_seed(seed)
        lens = [int(x) for x in torch.randint(T // 3, T + 1, (B,), generator=g)]
        lens[0] = T
        pred = torch.randn((B, T, S), generator=g) * 2.0
        ts = synth_labels(lens, S, seed, silent=silent)
        label = torch.nn.utils.rnn.pad_sequence(ts, padding_value=-1, batch_first=True)
        loss, idx, _, lab = ref.pit_loss(pred, label, lens)
        cnt = ref.calc_diarization_error(pred, lab, lens)
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), loss=np.float32(loss.item()),
                            min_idx=idx.numpy(), label_permute=lab.numpy().astype(np.int8),
                            der_counts=np.array(cnt, dtype=np.float64))
        manifest["cases"][name] = dict(kind="pit", B=B, T=T, S=S, seed=seed, silent=list(silent),
                                       lengths=lens, sha_pred=sha(pred))
        print(name, float(loss), idx.tolist())

    for name, (sr, fs, sh, n, nspk, seed, dt) in FEATURE_CASES.items():
        x, segs = fo.synth_audio(n, sr, nspk, seed, dtype=np.dtype(dt).type)
        Y = ref_feature.stft(x, fs, sh)
        lm = ref_feature.transform(Y, "logmel23_mn", sample_rate=sr)
        feats = np.ascontiguousarray(ref_feature.splice(lm, 7)[::10])
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), logmel_mn=lm, feats=feats,
                            n_frames=np.int64(Y.shape[0]))
        manifest["cases"][name] = dict(kind="feature", sample_rate=sr, frame_size=fs, frame_shift=sh,
                                       n_samples=n, n_speakers=nspk, seed=seed, dtype=dt, sha_audio=sha(x))
        print(name, Y.shape, feats.shape)

    with open(os.path.join(GOLD, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()

"""Full-size goldens from the UNMODIFIED reference (tests/golden/full_*.npz).

TEST INFRASTRUCTURE (see oracle/__init__.py).  Run in the build container only
(``python -m oracle.make_golden_full``): it imports /root/reference, which does not exist on the
GPU box.  Inputs are regenerated from seeds at test time; only reference OUTPUTS are stored.

Cases (BASELINE.json:configs order):
* ``full_c2``  configs[1] at full size: ``TransformerEdaModel.infer`` / ``forward`` on 64 x 500 frames,
  4 layers - the batch regime bench.py runs (persistent GEMM tiles, 8 LSTM clusters).
* ``full_c5``  configs[4]: ``forward`` (PIT, S = 3) on 32 x 2000 frames + ``infer`` posteriors of 3 sequences.
* ``full_c3_e2e``  configs[2]: 16 kHz audio (frame 400 / hop 160) -> the reference's ``feature.py`` (on the
  librosa stub) -> the reference's own DRIVER ``eend_eda/infer_eda.py:infer`` run unmodified (per-recording
  loop, chunks 2000/2000/800, one ``model.infer([chunk])`` each) -> ``T_hat`` per recording.
* ``full_c4_counts``  configs[3]: 8 kHz recordings (one of 30 min = 9 chunks) through the same unmodified
  driver with ``attractor_threshold=0.5``; the existence head ``eda.linear`` is FITTED (hard-margin linear
  SVM on the attractors of a first pass) so that the speaker count the reference decides is
  1..6 depending on the recording, every decision at least 0.2 away from the threshold.

The driver needs h5py / soundfile / a Kaldi data directory: stand-ins are registered for exactly
those (an in-memory ``h5py.File``, an empty ``soundfile``, a ``KaldiData`` handing out seeded
synthetic audio); everything between them - seeding, model construction, ``load_state_dict``,
``feature.*``, chunking, ``model.infer``, ``np.vstack`` - is the reference's own code.
"""
from __future__ import annotations

import json
import os
import sys
import tempfile
import types

import numpy as np
import torch

REF_ROOT = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")

SEED = 777                          # infer.py / train.py default --seed
C3 = dict(rate=16000, frame_size=400, frame_shift=160, n_layers=4, weight_seed=777, n_speakers=4,
          recs={"sess_a": (480.0, 31), "sess_b": (480.0 + 0.73, 32)})           # name: (seconds, audio seed)
C4 = dict(rate=8000, frame_size=200, frame_shift=80, n_layers=4, weight_seed=777,
          # name: (seconds, audio seed, true speakers = the count the fitted head must yield)
          recs={"r30min": (1800.0, 41, 3), "r500": (500.0, 42, 1), "r430": (430.3, 43, 2), "r620": (620.0, 44, 4),
                "r380": (379.9, 45, 5), "r710": (710.0, 46, 6)})


def pcm_audio(seconds: float, rate: int, n_spk: int, seed: int) -> np.ndarray:
    """Seeded synthetic mixture as 16-bit PCM (what a wav file stores)."""
    from oracle import feature_oracle as fo
    x, _ = fo.synth_audio(int(round(seconds * rate)), rate, n_spk, seed, dtype=np.float64)
    return np.clip(np.round(x * 32768.0), -32768, 32767).astype(np.int16)


# ------------------------------------------------------------------ stand-ins for the driver's I/O
H5_STORE = {}
RNG_AT_LOOP = {}


class _H5File:
    def __init__(self, path, mode="r"):
        self.path = path

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False

    def create_dataset(self, name, data=None):
        H5_STORE[(os.path.basename(self.path), name)] = np.array(data)


def make_fake_kaldi(recs: dict, rate: int):
    class FakeKaldi:
        """wavs + load_wav of kaldi_data.KaldiData (kaldi_data.py:146-163); load_wav returns what
        soundfile returns for a 16-bit wav: float64 = pcm / 32768 (kaldi_data.py:82)."""

        def __init__(self, data_dir):
            self.wavs = {k: k for k in recs}
            RNG_AT_LOOP["state"] = torch.get_rng_state().clone()     # generator state when the loop starts

        def load_wav(self, recid, start=0, end=None):
            return recs[recid].astype(np.float64) / 32768.0, rate
    return FakeKaldi


def run_driver(sd, recs_pcm, rate, frame_size, frame_shift, n_layers, attractor_threshold):
    """speaker_diarization/eend_eda/infer_eda.py:infer(args), unmodified."""
    from speaker_diarization.eend_eda import infer_eda
    from speaker_diarization import kaldi_data
    kaldi_data.KaldiData = make_fake_kaldi(recs_pcm, rate)
    H5_STORE.clear()
    with tempfile.TemporaryDirectory() as tmp:
        mf = os.path.join(tmp, "model.pt")
        torch.save(sd, mf)
        args = types.SimpleNamespace(
            seed=SEED, frame_size=frame_size, frame_shift=frame_shift, context_size=7, input_transform="logmel23_mn",
            model_type="TransformerEda", num_speakers=None, hidden_size=256, transformer_encoder_n_heads=4,
            transformer_encoder_n_layers=n_layers, model_file=mf, data_dir="fake", sampling_rate=rate, subsampling=10,
            chunk_size=2000, attractor_threshold=attractor_threshold, out_dir=tmp, label_delay=0)
        infer_eda.infer(args)
    return {k[0][:-3]: v for k, v in H5_STORE.items()}, RNG_AT_LOOP["state"].numpy().copy()


def attractors_of_driver_pass(sd, recs_pcm, rate, frame_size, frame_shift, n_layers, rng_state):
    """The driver's loop (infer_eda.py:92-113) replayed from the generator state it had at the loop's
    start, keeping the attractors the existence head sees (the head does not influence them)."""
    from speaker_diarization.eend_eda.models import TransformerEdaModel
    from speaker_diarization import feature as F
    model = TransformerEdaModel(None, 345, n_heads=4, n_units=256, n_layers=n_layers, has_pos=False)
    model.load_state_dict(sd)
    model.eval()
    torch.set_rng_state(torch.from_numpy(rng_state))
    out = []
    with torch.no_grad():
        for rec, pcm in recs_pcm.items():
            Y = F.stft(pcm.astype(np.float64) / 32768.0, frame_size, frame_shift)
            Y = F.transform(Y, transform_type="logmel23_mn", sample_rate=rate)
            Y = F.splice(Y, context_size=7)[::10]
            for s in range(0, len(Y), 2000):
                chunk = torch.from_numpy(Y[s:s + 2000].copy())
                emb_s, emb, ilens = model.forward_embedding([chunk])
                att, _ = model.eda(emb_s, torch.LongTensor(ilens), torch.zeros(1, 15, 256))
                out.append((rec, att[0].numpy().astype(np.float64)))
    return out


def main() -> None:
    sys.path.insert(0, REF_ROOT)
    sys.path.insert(0, os.path.dirname(HERE))
    from oracle import librosa_stub
    from oracle.weights import EdaConfig, synth_features, synth_labels, synth_state_dict
    librosa_stub.install()
    sys.modules["h5py"] = types.ModuleType("h5py")
    sys.modules["h5py"].File = _H5File
    sys.modules.setdefault("soundfile", types.ModuleType("soundfile"))
    from speaker_diarization.eend_eda.models import TransformerEdaModel

    torch.set_num_threads(os.cpu_count() or 1)
    with open(os.path.join(GOLD, "manifest.json")) as f:
        manifest = json.load(f)

    # ---------------------------------------------------------------- full_c2 / full_c5 -------
    for name, L, lens, S, wseed, dseed, keep in (("full_c2", 4, [500] * 64, 2, 777, 12, None),
                                                 ("full_c5", 4, [2000] * 32, 3, 777, 15, (0, 13, 31))):
        cfg = EdaConfig(n_layers=L, n_speakers=S)
        sd = synth_state_dict(cfg, wseed)
        model = TransformerEdaModel(S, cfg.in_size, cfg.n_heads, cfg.n_units, L)
        model.load_state_dict(sd, strict=True)
        model.eval()
        src = synth_features(lens, cfg.in_size, dseed)
        ts = synth_labels(lens, S, dseed)
        out = {}
        with torch.no_grad():
            torch.manual_seed(1000 + dseed)
            ys = model.infer(src, attractor_threshold=-1.0)
            post = torch.stack(ys).numpy()
            out["post_all"] = post if keep is None else post[list(keep)]
            out["labels_packed"] = np.packbits(post > 0.5)                    # every frame's decisions
            torch.manual_seed(2000 + dseed)
            loss, stats = model(src, ts)
            out["loss"] = np.float32(loss.item())
            for k, v in stats.items():
                out["stat_" + k] = np.float64(float(v))
            torch.manual_seed(2000 + dseed)
            emb_s, emb, ilens = model.forward_embedding(src)
            tsp = torch.nn.utils.rnn.pad_sequence(ts, padding_value=-1, batch_first=True)
            att, probs = model.eda(emb_s, torch.LongTensor(ilens), torch.zeros(len(lens), S + 1, cfg.n_units))
            pred = torch.bmm(emb, att[:, :-1].permute(0, 2, 1))
            _, idx, _, lab = model.pit_loss(pred, tsp, ilens)
            out["pred"] = pred.numpy()
            out["pit_min_idx"] = idx.numpy()
            out["label_permute_packed"] = np.packbits(lab.numpy() > 0.5)
            out["der_counts"] = np.array(model.calc_diarization_error(pred, lab, ilens), dtype=np.float64)
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
        manifest["cases"][name] = dict(kind="model_full", n_layers=L, lengths=lens, n_speakers=S, weight_seed=wseed,
                                       data_seed=dseed, silent=[], keep=None if keep is None else list(keep))
        print(name, "loss", float(out["loss"]), "der", float(out["stat_der"]))

    # ---------------------------------------------------------------- full_c3_e2e -------------
    c = C3
    sd = synth_state_dict(EdaConfig(n_layers=c["n_layers"]), c["weight_seed"])
    recs = {k: pcm_audio(sec, c["rate"], c["n_speakers"], seed) for k, (sec, seed) in c["recs"].items()}
    t_hat, rng = run_driver(sd, recs, c["rate"], c["frame_size"], c["frame_shift"], c["n_layers"], -1.0)
    out = {"rng_state": rng}
    for k, v in t_hat.items():
        out["T_hat_" + k] = v.astype(np.float32)
        print("full_c3_e2e", k, v.shape)
    np.savez_compressed(os.path.join(GOLD, "full_c3_e2e.npz"), **out)
    manifest["cases"]["full_c3_e2e"] = dict(kind="driver", seed=SEED, **{k: v for k, v in c.items() if k != "recs"},
                                            recs={k: list(v) for k, v in c["recs"].items()}, rec_order=list(c["recs"]),
                                            attractor_threshold=-1.0)

    # ---------------------------------------------------------------- full_c4_counts ----------
    c = C4
    sd = synth_state_dict(EdaConfig(n_layers=c["n_layers"]), c["weight_seed"])
    recs = {k: pcm_audio(sec, c["rate"], max(n, 1), seed) for k, (sec, seed, n) in c["recs"].items()}
    want = {k: n for k, (_, _, n) in c["recs"].items()}
    _, rng = run_driver(sd, recs, c["rate"], c["frame_size"], c["frame_shift"], c["n_layers"], -1.0)
    atts = attractors_of_driver_pass(sd, recs, c["rate"], c["frame_size"], c["frame_shift"], c["n_layers"], rng)
    rows, y = [], []
    for rec, a in atts:
        n = want[rec]
        for k in range(n + 1):
            rows.append(np.concatenate([a[k], [1.0]]))
            y.append(1.0 if k < n else -1.0)                      # logit +-1: p = 0.73 / 0.27
    A, y = np.array(rows), np.array(y)
    # hard-margin linear SVM = the smallest head weights with y * logit >= 1 (p >= 0.73 / <= 0.27): the decoder's
    # attractor trajectory contracts (a_k - a_{k+1} shrinks ~0.6x per step), so an exact interpolation would
    # need weights ~1e5 and amplify any kernel rounding; the max-margin solution stays ~1e2
    from sklearn.svm import SVC
    svm = SVC(kernel="linear", C=1e7, tol=1e-7).fit(A[:, :-1], y)
    wb = np.concatenate([svm.coef_[0], svm.intercept_])
    margin = (y * (A @ wb)).min()
    print("full_c4_counts: constraints", A.shape, "min y*logit", margin, "|w|_2", np.linalg.norm(wb[:-1]))
    assert margin > 0.99
    sd["eda.linear.weight"] = torch.from_numpy(wb[:-1].astype(np.float32)).reshape(1, 256)
    sd["eda.linear.bias"] = torch.from_numpy(wb[-1:].astype(np.float32))
    t_hat, rng2 = run_driver(sd, recs, c["rate"], c["frame_size"], c["frame_shift"], c["n_layers"], 0.5)
    assert np.array_equal(rng, rng2)
    out = {"rng_state": rng, "head_w": sd["eda.linear.weight"].numpy(), "head_b": sd["eda.linear.bias"].numpy()}
    for k, v in t_hat.items():
        assert v.shape[1] == want[k], (k, v.shape, want[k])
        out["T_hat_" + k] = v.astype(np.float32)
        print("full_c4_counts", k, v.shape)
    # margin of every existence decision the driver made (all 15 attractors of every chunk)
    w64, b64 = wb[:-1], wb[-1]
    margins = []
    for rec, a in atts:
        p = 1.0 / (1.0 + np.exp(-(a @ w64 + b64)))
        margins.append(np.abs(p[: want[rec] + 1] - 0.5).min())
    out["min_margin"] = np.float64(min(margins))
    print("full_c4_counts: min |p - 0.5| over the deciding attractors", out["min_margin"])
    np.savez_compressed(os.path.join(GOLD, "full_c4_counts.npz"), **out)
    manifest["cases"]["full_c4_counts"] = dict(kind="driver", seed=SEED, **{k: v for k, v in c.items() if k != "recs"},
                                               recs={k: list(v) for k, v in c["recs"].items()}, rec_order=list(c["recs"]),
                                               attractor_threshold=0.5)

    with open(os.path.join(GOLD, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()

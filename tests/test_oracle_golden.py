"""Pin the CPU oracle to the outputs the unmodified reference produced
(tests/golden/*.npz, written by oracle/make_golden.py)."""
import hashlib

import numpy as np
import pytest
import torch

from conftest import cases_of, load_golden
from oracle import feature_oracle as fo
from oracle.eend_eda_oracle import OracleEda, diarization_counts, pit_loss
from oracle.weights import EdaConfig, synth_features, synth_labels, synth_state_dict


def sha(a):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else np.asarray(a)
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


def build(case):
    cfg = EdaConfig(n_layers=case["n_layers"], n_speakers=case["n_speakers"])
    sd = synth_state_dict(cfg, case["weight_seed"])
    src = synth_features(case["lengths"], cfg.in_size, case["data_seed"])
    ts = synth_labels(case["lengths"], case["n_speakers"], case["data_seed"], silent=tuple(case["silent"]))
    return cfg, sd, src, ts


@pytest.mark.parametrize("name", cases_of("model"))
@pytest.mark.parametrize("library_lstm", [False, True])
def test_model_oracle_matches_reference(name, library_lstm, manifest):
    case = manifest["cases"][name]
    if not library_lstm and max(case["lengths"]) > 600:
        pytest.skip("python-loop LSTM only on the short cases")
    cfg, sd, src, ts = build(case)
    assert sha(sd["encoder.weight"]) == case["sha_w0"]
    assert sha(src[0]) == case["sha_x0"] and sha(ts[0]) == case["sha_t0"]
    g = load_golden(name)
    o = OracleEda(sd, library_lstm=library_lstm)
    d = case["data_seed"]
    torch.manual_seed(1000 + d)
    ys = o.infer(src, attractor_threshold=-1.0)
    for b, y in enumerate(ys):
        np.testing.assert_allclose(y.numpy(), g[f"post_all_{b}"], atol=2e-5, rtol=0)
    np.testing.assert_allclose(o.trace["attractors"].numpy(), g["attractors15"], atol=2e-5, rtol=0)
    np.testing.assert_allclose(o.trace["probs"].numpy(), g["probs15"], atol=2e-5, rtol=0)
    np.testing.assert_allclose(o.trace["emb"][:, ::17, ::5].numpy(), g["emb_sample"], atol=5e-5, rtol=0)
    torch.manual_seed(1000 + d)
    ys = o.infer(src, attractor_threshold=0.5)
    assert [y.shape[1] for y in ys] == g["nspk_thr05"].tolist()
    torch.manual_seed(2000 + d)
    loss, stats = o.forward(src, ts)
    assert abs(float(loss) - float(g["loss"])) < 2e-5
    np.testing.assert_allclose(o.trace["pred"].numpy(), g["pred"], atol=1e-4, rtol=0)
    assert o.trace["min_idx"].tolist() == g["pit_min_idx"].tolist()
    assert np.array_equal(o.trace["label_permute"].numpy().astype(np.int8), g["label_permute"])
    assert list(o.trace["counts"]) == g["der_counts"].tolist()
    for k in ("sad_mr", "sad_fr", "mi", "fa", "cf", "acc", "der"):
        assert abs(float(stats[k]) - float(g["stat_" + k])) < 1e-12, k


@pytest.mark.parametrize("name", cases_of("pit"))
def test_pit_oracle_matches_reference(name, manifest):
    case = manifest["cases"][name]
    B, T, S, seed = case["B"], case["T"], case["S"], case["seed"]
    gen = torch.Generator().manual_seed(seed)
    lens = [int(x) for x in torch.randint(T // 3, T + 1, (B,), generator=gen)]
    lens[0] = T
    assert lens == case["lengths"]
    pred = torch.randn((B, T, S), generator=gen) * 2.0
    assert sha(pred) == case["sha_pred"]
    ts = synth_labels(lens, S, seed, silent=tuple(case["silent"]))
    label = torch.nn.utils.rnn.pad_sequence(ts, padding_value=-1, batch_first=True)
    g = load_golden(name)
    loss, idx, perms, lab = pit_loss(pred, label, lens)
    assert abs(float(loss) - float(g["loss"])) < 1e-6
    assert idx.tolist() == g["min_idx"].tolist()
    assert np.array_equal(lab.numpy().astype(np.int8), g["label_permute"])
    assert list(diarization_counts(pred, lab, lens)) == g["der_counts"].tolist()


@pytest.mark.parametrize("name", cases_of("feature"))
def test_feature_oracle_matches_reference(name, manifest):
    case = manifest["cases"][name]
    x, _ = fo.synth_audio(case["n_samples"], case["sample_rate"], case["n_speakers"], case["seed"],
                          dtype=np.dtype(case["dtype"]).type)
    assert sha(x) == case["sha_audio"]
    g = load_golden(name)
    Y = fo.stft(x, case["frame_size"], case["frame_shift"])
    assert Y.shape[0] == int(g["n_frames"]) == fo.n_stft_frames(case["n_samples"], case["frame_shift"])
    lm = fo.transform_logmel23_mn(Y, case["sample_rate"])
    np.testing.assert_allclose(lm, g["logmel_mn"], atol=1e-6, rtol=0)
    feats = fo.frontend(x, case["sample_rate"], case["frame_size"], case["frame_shift"])
    np.testing.assert_allclose(feats, g["feats"], atol=1e-6, rtol=0)
    assert feats.shape[1] == fo.get_input_dim(case["frame_size"], 7, "logmel23_mn") == 345


def _transform_cases():
    import json
    g = load_golden("fe_transforms")
    return g, json.loads(str(g["meta"]))


@pytest.mark.parametrize("name", ["tf_none_8k", "tf_log_8k", "tf_log_8k_f32", "tf_espnet_8k", "tf_espnet_16k"])
def test_transform_variants_oracle_matches_reference(name):
    g, meta = _transform_cases()
    c = meta[name]
    x, _ = fo.synth_audio(c["n_samples"], c["sample_rate"], 2, c["seed"], dtype=np.dtype(c["dtype"]).type)
    Y = fo.transform(fo.stft(x, c["frame_size"], c["frame_shift"]), c["transform_type"], c["sample_rate"])
    assert Y.dtype == np.float32 and Y.shape == g[name + "/Y"].shape
    np.testing.assert_allclose(Y, g[name + "/Y"], atol=1e-6, rtol=1e-6)
    feats = np.ascontiguousarray(fo.splice(Y, 7)[::10])
    np.testing.assert_allclose(feats, g[name + "/feats"], atol=1e-6, rtol=1e-6)
    if c["transform_type"]:
        assert feats.shape[1] == fo.get_input_dim(c["frame_size"], 7, c["transform_type"]) == c["input_dim"]
    assert set(meta["_positional_mel_variants_raise"].values()) == {"TypeError"}
    with pytest.raises(TypeError):
        fo.transform(fo.stft(x, c["frame_size"], c["frame_shift"]), "logmel23_mvn", c["sample_rate"])

"""End-to-end parity of the drop-in model and front end against the golden outputs of the
unmodified reference (tests/golden/, made by oracle/make_golden.py) and the CPU oracle.

north_star tolerances: posteriors max-abs <= 1e-3; PIT assignments and thresholded labels
bit-exact; DER within 0.1 % absolute.  The fp32 path is held to much tighter bounds.
"""
import numpy as np
import pytest
import torch

from conftest import cases_of, load_golden

pytestmark = pytest.mark.gpu

POST_TOL = {"fp32": 5e-5, "bf16x3": 1e-3}
PRECISIONS = ["fp32", "bf16x3"]


def build(case, precision):
    from oracle.weights import EdaConfig, synth_features, synth_labels, synth_state_dict
    from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
    cfg = EdaConfig(n_layers=case["n_layers"], n_speakers=case["n_speakers"])
    sd = synth_state_dict(cfg, case["weight_seed"])
    model = TransformerEdaModel(case["n_speakers"], cfg.in_size, cfg.n_heads, cfg.n_units, cfg.n_layers,
                                precision=precision)
    model.load_state_dict(sd, strict=True)
    model = model.cuda().eval()
    src = synth_features(case["lengths"], cfg.in_size, case["data_seed"])
    ts = synth_labels(case["lengths"], case["n_speakers"], case["data_seed"], silent=tuple(case["silent"]))
    return model, sd, src, ts


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("name", cases_of("model"))
def test_infer_matches_reference_golden(name, precision, manifest):
    case = manifest["cases"][name]
    model, sd, src, ts = build(case, precision)
    g = load_golden(name)
    d = case["data_seed"]
    torch.manual_seed(1000 + d)
    ys = model.infer(src, attractor_threshold=-1.0, activity_threshold=0.5)   # host (CPU) inputs
    worst, flips, total = 0.0, 0, 0
    for b, y in enumerate(ys):
        ref = g[f"post_all_{b}"]
        assert tuple(y.shape) == ref.shape
        worst = max(worst, float(np.abs(y.cpu().numpy() - ref).max()))
        n = case["lengths"][b]
        mine = model.last["labels"][b, :n].cpu().numpy().astype(bool)
        flips += int((mine != (ref[:n] > 0.5)).sum())
        total += mine.size
    assert worst <= POST_TOL[precision], f"posterior max-abs-diff {worst}"
    assert flips == 0, f"{flips}/{total} thresholded speaker-activity labels differ from the reference"
    np.testing.assert_allclose(model.last["attractors"].cpu().numpy(), g["attractors15"], atol=POST_TOL[precision], rtol=0)
    np.testing.assert_allclose(model.last["probs"].cpu().numpy(), g["probs15"], atol=POST_TOL[precision], rtol=0)
    np.testing.assert_allclose(model.last["emb"][:, ::17, ::5].cpu().numpy(), g["emb_sample"],
                               atol=10 * POST_TOL[precision], rtol=0)
    # speaker counting by attractor threshold, device-resident inputs this time
    torch.manual_seed(1000 + d)
    ys = model.infer([s.cuda() for s in src], attractor_threshold=0.5)
    assert [int(y.shape[1]) for y in ys] == g["nspk_thr05"].tolist()


@pytest.mark.parametrize("precision", PRECISIONS)
@pytest.mark.parametrize("name", cases_of("model"))
def test_forward_matches_reference_golden(name, precision, manifest):
    case = manifest["cases"][name]
    model, sd, src, ts = build(case, precision)
    g = load_golden(name)
    torch.manual_seed(2000 + case["data_seed"])
    loss, stats = model(src, ts)
    assert loss.is_cuda and loss.dim() == 0
    assert abs(loss.item() - float(g["loss"])) < 20 * POST_TOL[precision]
    assert np.abs(model.last["pred"].cpu().numpy() - g["pred"]).max() < 20 * POST_TOL[precision]
    assert model.last["min_idx"].cpu().tolist() == g["pit_min_idx"].tolist()                 # bit-exact
    assert np.array_equal(model.last["label_permute"].cpu().numpy().astype(np.int8), g["label_permute"])
    S = case["n_speakers"]
    want = g["der_counts"].tolist()
    got = model.last["counts"].cpu().tolist()
    assert got[0] == round(want[0] * S) and [float(c) for c in got[1:]] == want[1:]
    for k in ("sad_mr", "sad_fr", "mi", "fa", "cf", "acc", "der"):
        assert abs(float(stats[k]) - float(g["stat_" + k])) <= 1e-3, k                        # DER within 0.1 % abs
    for k in ("loss", "loss_att", "loss_pit"):
        assert abs(float(stats[k]) - float(g["stat_" + k])) < 20 * POST_TOL[precision], k


def test_public_helpers_match_oracle():
    """pit_loss / calc_diarization_error / attractor_loss / eda() / forward_embedding called
    the way a user of the reference class would call them."""
    from oracle import eend_eda_oracle as O
    from oracle.weights import EdaConfig, synth_features, synth_labels, synth_state_dict
    from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
    cfg = EdaConfig(n_layers=1)
    sd = synth_state_dict(cfg, 9)
    m = TransformerEdaModel(2, 345, 4, 256, 1)
    m.load_state_dict(sd)
    m = m.cuda().eval()
    lens = [50, 31]
    g = torch.Generator().manual_seed(4)
    pred = torch.randn((2, 50, 3), generator=g)
    label = torch.nn.utils.rnn.pad_sequence(synth_labels(lens, 3, 4), padding_value=-1, batch_first=True)
    loss, idx, plist, lab = m.pit_loss(pred.cuda(), label.cuda(), lens)
    lo, io, po, labo = O.pit_loss(pred, label, lens)
    assert abs(loss.item() - lo.item()) < 1e-6 and idx.cpu().tolist() == io.tolist()
    assert not lab.is_cuda and lab.dtype == torch.float32 and torch.equal(lab, labo)
    assert all(np.array_equal(a, b) for a, b in zip(plist, po))
    assert m.calc_diarization_error(pred.cuda(), lab, lens) == O.diarization_counts(pred, labo, lens)
    probs = torch.rand((2, 4), generator=g)
    la = m.attractor_loss([p for p in probs], label)
    assert abs(la.item() - O.attractor_loss(probs, 3).item()) < 1e-6
    d = m.pit_loss_backward(pred.cuda(), lab, lens)
    assert d.shape == pred.shape and float(d[1, 31:].abs().max()) == 0.0
    # eda() on already-shuffled input, reference signature
    src = synth_features(lens, 345, 4)
    torch.manual_seed(3)
    shuf, emb, ilens = m.forward_embedding(src)
    assert ilens == lens and shuf.shape == emb.shape == (2, 50, 256)
    att, plist2 = m.eda(shuf, torch.LongTensor(ilens), torch.zeros(2, 3, 256).cuda())
    torch.manual_seed(3)
    o = O.OracleEda(sd)
    emb_o, _, perms = o.forward_embedding(src)
    att_o, p_o = O.eda(emb_o, lens, perms, 3, o.sd)
    assert (emb.cpu() - emb_o).abs().max().item() < 5e-5
    assert (att.cpu() - att_o).abs().max().item() < 5e-5
    assert (torch.stack(plist2).cpu() - p_o).abs().max().item() < 5e-5


def test_infer_num_speakers_paths():
    from oracle.weights import EdaConfig, synth_features, synth_state_dict
    from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
    m = TransformerEdaModel(2, 345, 4, 256, 1)
    m.load_state_dict(synth_state_dict(EdaConfig(n_layers=1), 9))
    m = m.cuda().eval()
    src = synth_features([40, 25], 345, 1)
    torch.manual_seed(0)
    try:
        ys = m.infer(src, infer_num_speakers=2)
    except IndexError:                      # reference raises too when the order hits index 14 (models.py:339)
        return
    assert [tuple(y.shape) for y in ys] == [(40, 2), (40, 2)]
    with pytest.raises(NotImplementedError):
        m.train()
        m(src, [torch.zeros(40, 2), torch.zeros(25, 2)])


@pytest.mark.parametrize("name", cases_of("feature"))
def test_frontend_matches_reference_golden(name, manifest):
    from oracle import feature_oracle as fo
    from speaker_diarization_b200 import feature
    case = manifest["cases"][name]
    x, _ = fo.synth_audio(case["n_samples"], case["sample_rate"], case["n_speakers"], case["seed"],
                          dtype=np.dtype(case["dtype"]).type)
    g = load_golden(name)
    fe = feature.FrontEnd(case["sample_rate"], case["frame_size"], case["frame_shift"], 7, 10)
    lm, colsum, _ = fe.logmel([x])
    assert lm.shape[0] == int(g["n_frames"])
    mean = (colsum[0] / lm.shape[0]).cpu().numpy()
    tol = 2e-5 if case["dtype"] == "float64" else 2e-4     # float32 audio: the reference itself runs in fp32
    assert np.abs((lm.cpu().numpy().astype(np.float64) - mean) - g["logmel_mn"]).max() < tol
    feats = fe.features([x])[0]
    assert tuple(feats.shape) == g["feats"].shape
    assert np.abs(feats.cpu().numpy() - g["feats"]).max() < tol
    one = feature.extract(x, case["frame_size"], case["frame_shift"], "logmel23_mn", case["sample_rate"], 7, 10)
    assert torch.equal(one, feats)


def test_frontend_batch_of_ragged_recordings():
    from oracle import feature_oracle as fo
    from speaker_diarization_b200 import feature
    fe = feature.FrontEnd(8000, 200, 80, 7, 10)
    recs = [fo.synth_audio(n, 8000, 2, 40 + i, dtype=np.float64)[0] for i, n in enumerate([8000, 79, 80 * 33, 12345, 1])]
    outs = fe.features(recs, ld=352)
    for r, o in zip(recs, outs):
        want = fo.frontend(r, 8000, 200, 80)
        assert tuple(o.shape) == want.shape
        if want.size:
            assert np.abs(o.cpu().numpy() - want).max() < 2e-5


def test_frontend_int16_pcm_matches_soundfile_scaling():
    """int16 PCM fed straight to the kernel == the float64 samples soundfile would hand the
    reference (x / 32768), both 8 kHz (n_fft 256) and 16 kHz (n_fft 512) fast paths."""
    from oracle import feature_oracle as fo
    from speaker_diarization_b200 import feature
    for rate, fs, hop in ((8000, 200, 80), (16000, 400, 160)):
        x, _ = fo.synth_audio(rate * 3 + 17, rate, 2, 91, dtype=np.float64)
        pcm = np.clip(np.round(x * 32768.0), -32768, 32767).astype(np.int16)
        want = fo.frontend(pcm.astype(np.float64) / 32768.0, rate, fs, hop)
        fe = feature.FrontEnd(rate, fs, hop, 7, 10)
        got = fe.features([pcm])[0]
        assert tuple(got.shape) == want.shape
        assert np.abs(got.cpu().numpy() - want).max() < 2e-5


def test_pipeline_matches_infer():
    """InferencePipeline (3 streams, 2 batches in flight, int16 PCM from pinned host memory) ==
    feature.extract + TransformerEdaModel.infer batch by batch with the same randperm stream."""
    from oracle import feature_oracle as fo
    from oracle.weights import EdaConfig, synth_state_dict
    from speaker_diarization_b200 import feature
    from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
    from speaker_diarization_b200.pipeline import InferencePipeline
    cfg = EdaConfig(n_layers=2)
    m = TransformerEdaModel(2, cfg.in_size, cfg.n_heads, cfg.n_units, cfg.n_layers, precision="bf16x3")
    m.load_state_dict(synth_state_dict(cfg, 777))
    m = m.cuda().eval()
    fe = feature.FrontEnd(8000, 200, 80, 7, 10)
    batches = []
    for k in range(4):
        n = 8000 * (4 + k)
        pcm = np.stack([np.clip(np.round(fo.synth_audio(n, 8000, 2, 10 * k + i, dtype=np.float64)[0] * 32768), -32768, 32767)
                        for i in range(3)]).astype(np.int16)
        batches.append((torch.from_numpy(pcm).pin_memory(), [n] * 3))
    torch.manual_seed(5)
    want = []
    for audio, counts in batches:
        feats = fe.features([a for a in audio])
        ys = m.infer(feats, attractor_threshold=0.5, activity_threshold=0.5)
        want.append((m.last["posteriors"].cpu(), [int(y.shape[1]) for y in ys], m.last["labels"].cpu()))
    torch.manual_seed(5)
    pipe = InferencePipeline(m, fe, depth=2)
    got = list(pipe.run(batches))
    assert len(got) == 4
    for (wp, wn, wl), g in zip(want, got):
        B, T, n_att = wp.shape
        # recordings lie back to back on the time axis (np.vstack of infer_eda.py:121); equal lengths here
        assert torch.equal(g["posteriors"].view(B, T, n_att), wp) and torch.equal(g["labels"].view(B, T, n_att), wl)
        assert [min(x, 14) for x in g["n_speakers"]] == wn
        assert g["lengths"] == [T] * 3 and g["rec_offsets"].tolist() == [0, T, 2 * T, 3 * T]


def _full_size_model(n_layers, precision):
    from oracle.weights import EdaConfig, synth_state_dict
    from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
    cfg = EdaConfig(n_layers=n_layers)
    m = TransformerEdaModel(2, cfg.in_size, cfg.n_heads, cfg.n_units, cfg.n_layers, precision=precision)
    m.load_state_dict(synth_state_dict(cfg, 777))
    return m.cuda().eval(), cfg


def test_full_size_c2_batch_invariance():
    """configs[1] at full size (64 x 500 frames, 4 layers); the comparison with the reference itself is
    tests/test_gpu_full.py.  Here a size-independent property: a sequence's posteriors do not depend on what
    else is in the batch.  The full batch must agree with single-sequence runs (same frame shuffle) to well
    inside the parity bar, and the two precisions must agree with each other within it."""
    from oracle.weights import synth_features
    m, cfg = _full_size_model(4, "bf16x3")
    lens = [500] * 64
    src = [s.cuda() for s in synth_features(lens, cfg.in_size, 11)]
    g = torch.Generator().manual_seed(3)
    perms = [torch.randperm(n, generator=g) for n in lens]
    m.infer(src, attractor_threshold=0.5, activity_threshold=0.5, perms=perms)
    post, labels = m.last["posteriors"].clone(), m.last["labels"].clone()
    for b in (0, 17, 63):
        m.infer([src[b]], attractor_threshold=0.5, activity_threshold=0.5, perms=[perms[b]])
        assert (m.last["posteriors"][0] - post[b]).abs().max().item() < 2e-5
    m32, _ = _full_size_model(4, "fp32")
    m32.infer(src, attractor_threshold=0.5, activity_threshold=0.5, perms=perms)
    d = (m32.last["posteriors"] - post).abs()
    assert d.max().item() < 1e-3
    # thresholded labels may only differ where the posterior sits within the tolerance of 0.5
    flip = m32.last["labels"] != labels
    assert bool(((m32.last["posteriors"] - 0.5).abs()[flip] < 1e-3).all())


def test_full_size_c4_long_form_chunks():
    """configs[3]: a 30-minute recording = 9 chunks of 2000 frames, speaker count decided by the
    attractor threshold.  Chunks in one batch == chunks one at a time (the reference's loop,
    infer_eda.py:102-113), and ragged tails do not disturb the full chunks."""
    from oracle.weights import synth_features
    m, cfg = _full_size_model(4, "bf16x3")
    lens = [2000] * 8 + [800]
    src = [s.cuda() for s in synth_features(lens, cfg.in_size, 23)]
    g = torch.Generator().manual_seed(9)
    perms = [torch.randperm(n, generator=g) for n in lens]
    ys = m.infer(src, attractor_threshold=0.5, perms=perms)
    post = m.last["posteriors"].clone()
    probs = m.last["probs"].clone()
    for b in (0, 7):                 # full chunks: independent of the rest of the batch
        y1 = m.infer([src[b]], attractor_threshold=0.5, perms=[perms[b]])
        assert (m.last["posteriors"][0] - post[b]).abs().max().item() < 2e-5
        assert (m.last["probs"][0] - probs[b]).abs().max().item() < 2e-5
        assert y1[0].shape[1] == ys[b].shape[1]
    # the 800-frame tail attends to its own -1 padding (no key-padding mask in the reference,
    # models.py:223), so it depends on Tmax only: same result next to ONE full chunk
    y2 = m.infer([src[3], src[8]], attractor_threshold=0.5, perms=[perms[3], perms[8]])
    assert (m.last["posteriors"][1] - post[8]).abs().max().item() < 2e-5
    assert y2[1].shape[1] == ys[8].shape[1]


def test_eend_eda_model_transformer_variant(manifest):
    dev = torch.device("cuda:0")
    """The reference's second class EendEdaModel(encoder_type='transformer') (models.py:465-700):
    same kernels under its own state_dict names; infer keeps the FIRST n columns (models.py:644)."""
    from oracle.make_golden_eend import rename
    from oracle.weights import EdaConfig, synth_features, synth_labels, synth_state_dict
    from speaker_diarization_b200.eend_eda.models import EendEdaModel
    case = manifest["cases"]["eendeda_l2"]
    g = load_golden("eendeda_l2")
    cfg = EdaConfig(n_layers=case["n_layers"], n_speakers=case["n_speakers"])
    m = EendEdaModel(case["n_speakers"], cfg.in_size, cfg.n_heads, cfg.n_units, cfg.n_layers, precision="bf16x3")
    m.load_state_dict(rename(synth_state_dict(cfg, case["weight_seed"])), strict=True)
    m = m.to(dev).eval()
    src = synth_features(case["lengths"], cfg.in_size, case["data_seed"])
    ts = synth_labels(case["lengths"], case["n_speakers"], case["data_seed"])
    torch.manual_seed(1000 + case["data_seed"])
    for b, y in enumerate(m.infer(src, infer_num_speakers=2)):
        assert (y.cpu().numpy() - g[f"post2_{b}"]).__abs__().max() < 1e-3
    torch.manual_seed(1000 + case["data_seed"])
    assert [y.shape[1] for y in m.infer(src, attractor_threshold=0.5)] == g["nspk_thr05"].tolist()
    torch.manual_seed(2000 + case["data_seed"])
    loss, stats = m(src, ts)
    assert abs(loss.item() - float(g["loss"])) < 1e-3 and abs(stats["der"] - float(g["der"])) < 1e-3
    with pytest.raises(NotImplementedError):
        EendEdaModel(2, cfg.in_size, 4, 256, 2, encoder_type="rnn")          # models.py:503

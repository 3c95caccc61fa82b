"""Per-kernel parity: each C-ABI entry point against the CPU oracle on seeded inputs.

Tolerances are written per test.  PIT assignments, permuted labels, DER counters and
thresholded activity labels are compared bit-exact.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import cases_of, load_golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


PREC = ["fp32"]
LIN_PREC = ["fp32", "bf16x3", "bf16"]


def _prec(name):
    from speaker_diarization_b200 import _lib as L
    return L.PRECISIONS[name]


@pytest.mark.parametrize("prec", LIN_PREC)
@pytest.mark.parametrize("M,N,K,epi", [
    (300, 768, 256, "bias"), (1000, 2048, 256, "relu"), (257, 256, 352, "ln"),
    (515, 256, 2048, "res_ln"), (64, 1024, 256, "bias"), (1, 256, 256, "res_ln"), (129, 40, 64, "bias"),
])
def test_linear(dev, prec, M, N, K, epi):
    from speaker_diarization_b200 import _lib as L, ops
    g = torch.Generator().manual_seed(M * 7 + N)
    a = torch.randn((M, K), generator=g)
    w = (torch.rand((N, K), generator=g) * 2 - 1) / K ** 0.5
    b = torch.randn((N,), generator=g) * 0.1
    r = torch.randn((M, N), generator=g)
    gm, bt = 1 + 0.1 * torch.randn((N,), generator=g), 0.1 * torch.randn((N,), generator=g)
    ref = F.linear(a, w, b)
    code = {"bias": L.EPI_BIAS, "relu": L.EPI_BIAS_RELU, "ln": L.EPI_BIAS_LN, "res_ln": L.EPI_BIAS_RES_LN}[epi]
    if epi == "relu":
        ref = ref.relu()
    if epi == "ln":
        ref = F.layer_norm(ref, (N,), gm, bt, 1e-5)
    if epi == "res_ln":
        ref = F.layer_norm(ref + r, (N,), gm, bt, 1e-5)
    out = ops.linear(a.to(dev), w.to(dev), b.to(dev), code, _prec(prec),
                     resid=r.to(dev) if epi == "res_ln" else None,
                     gamma=gm.to(dev) if "ln" in epi else None, beta=bt.to(dev) if "ln" in epi else None)
    # fp32: FFMA reorder only; bf16x3: dropped lo*lo term ~2^-18 rel; bf16: 2^-9 operand rounding
    tol = {"fp32": 2e-5, "bf16x3": 1e-4, "bf16": 6e-2}[prec]
    err = (out.cpu() - ref).abs().max().item()
    assert err < tol, err
    if prec == "bf16":
        assert err > 1e-5      # the single-pass mode really is the cheaper arithmetic


@pytest.mark.parametrize("prec", ["fp32", "bf16x3"])
@pytest.mark.parametrize("B,T", [(2, 130), (1, 64), (3, 500), (1, 1), (1, 2000), (2, 128), (5, 257)])
def test_mha(dev, prec, B, T):
    from speaker_diarization_b200 import ops
    g = torch.Generator().manual_seed(B * 1000 + T)
    qkv = torch.randn((B, T, 768), generator=g) * 1.5
    q, k, v = (t.reshape(B, T, 4, 64).transpose(1, 2) for t in qkv.split(256, dim=-1))
    ref = (torch.softmax(q @ k.transpose(-1, -2) * 0.125, dim=-1) @ v).transpose(1, 2).reshape(B * T, 256)
    out = ops.mha(qkv.reshape(B * T, 768).to(dev), B, T, 4, _prec(prec))
    tol = 2e-5 if prec == "fp32" else 1e-4   # logits reach |20|: 1 ulp there is 2e-6 before exp
    assert (out.cpu() - ref).abs().max().item() < tol


@pytest.mark.parametrize("lens,n_dec", [([40, 17, 1, 40, 33], 3), ([300] * 9, 15), ([5], 15), ([64, 0, 20], 4),
                                        ([37 + (i * 7) % 23 for i in range(66)], 3),      # 5 sequences per cluster
                                        ([20 + (i * 5) % 11 for i in range(117)], 2),     # 8 per cluster
                                        ([10 + i % 7 for i in range(130)], 2)])           # more clusters than fit at once
@pytest.mark.parametrize("lstm_prec", ["fp32", "bf16x3"])     # CUDA-core fp32 / tensor-core split-bf16 recurrence
def test_lstm_attractors(dev, lens, n_dec, lstm_prec):
    from oracle.eend_eda_oracle import eda as oracle_eda
    from oracle.weights import EdaConfig, synth_state_dict
    from speaker_diarization_b200 import _lib as L, ops
    sd = synth_state_dict(EdaConfig(n_layers=1), 3)
    B, T = len(lens), max(lens)
    g = torch.Generator().manual_seed(sum(lens))
    emb = torch.randn((B, T, 256), generator=g)
    perms = [torch.randperm(n, generator=g) for n in lens]
    # torch's packed LSTM rejects empty sequences; the python-loop oracle handles them (h=c=0)
    att_ref, p_ref = oracle_eda(emb, lens, perms, n_dec, sd, library_lstm=(0 not in lens))
    perm = torch.arange(T, dtype=torch.int32).repeat(B, 1)
    for i, p in enumerate(perms):
        perm[i, : lens[i]] = p.to(torch.int32)
    be = (sd["eda.encoder.bias_ih_l0"] + sd["eda.encoder.bias_hh_l0"]).to(dev)
    xproj = ops.linear(emb.reshape(B * T, 256).to(dev), sd["eda.encoder.weight_ih_l0"].to(dev), be, L.EPI_BIAS, L.PREC_FP32)
    lens_d = torch.tensor(lens, dtype=torch.int32, device=dev)
    order = torch.tensor(sorted(range(B), key=lambda b: -lens[b]), dtype=torch.int32, device=dev)
    bd = (sd["eda.decoder.bias_ih_l0"] + sd["eda.decoder.bias_hh_l0"]).to(dev)
    att, _ = ops.lstm_attractors(xproj, perm.to(dev), lens_d, order, sd["eda.encoder.weight_hh_l0"].to(dev),
                                 sd["eda.decoder.weight_hh_l0"].to(dev), bd, B, T, n_dec, prec=_prec(lstm_prec))
    probs, _ = ops.attractor_head(att, sd["eda.linear.weight"].reshape(-1).to(dev), sd["eda.linear.bias"].to(dev))
    # split-bf16 products carry ~2^-17 relative error per term through up to 300 recurrent steps
    assert (att.cpu() - att_ref).abs().max().item() < (2e-5 if lstm_prec == "fp32" else 1e-4)
    assert (probs.cpu() - p_ref).abs().max().item() < (2e-6 if lstm_prec == "fp32" else 2e-5)
    # identity `order` (NULL) must give the same result
    att2, _ = ops.lstm_attractors(xproj, perm.to(dev), lens_d, None, sd["eda.encoder.weight_hh_l0"].to(dev),
                                  sd["eda.decoder.weight_hh_l0"].to(dev), bd, B, T, n_dec, prec=_prec(lstm_prec))
    assert torch.equal(att, att2)


@pytest.mark.parametrize("B,T,n_att", [(3, 77, 14), (2, 500, 2), (1, 1, 1), (70, 40, 4)])
def test_posterior_and_threshold(dev, B, T, n_att):
    from speaker_diarization_b200 import ops
    g = torch.Generator().manual_seed(B + T + n_att)
    emb = torch.randn((B, T, 256), generator=g)
    att = torch.randn((B, n_att + 1, 256), generator=g) * 0.2
    ref_logit = emb @ att[:, :n_att].transpose(1, 2)
    logits, post, labels = ops.posterior(emb.reshape(B * T, 256).to(dev), att.to(dev), B, T, n_att, 0.5,
                                         want_logits=True, want_post=True, want_labels=True)
    assert (logits.cpu() - ref_logit).abs().max().item() < 2e-5
    assert (post.cpu() - torch.sigmoid(ref_logit)).abs().max().item() < 2e-6
    # the fused decision is exactly `posterior > threshold` on the posteriors it wrote
    assert torch.equal(labels.cpu(), (post.cpu() > 0.5).to(torch.uint8))
    # and agrees with the oracle wherever the oracle posterior is not within 1e-5 of 0.5
    ref_lab = (torch.sigmoid(ref_logit) > 0.5)
    far = (torch.sigmoid(ref_logit) - 0.5).abs() > 1e-5
    assert torch.equal(labels.cpu().bool()[far], ref_lab[far])


@pytest.mark.parametrize("name", cases_of("pit"))
def test_pit_loss_and_der_golden(dev, name, manifest):
    from oracle.weights import synth_labels
    from speaker_diarization_b200 import ops
    case = manifest["cases"][name]
    B, T, S, seed = case["B"], case["T"], case["S"], case["seed"]
    gen = torch.Generator().manual_seed(seed)
    lens = [int(x) for x in torch.randint(T // 3, T + 1, (B,), generator=gen)]
    lens[0] = T
    pred = torch.randn((B, T, S), generator=gen) * 2.0
    ts = synth_labels(lens, S, seed, silent=tuple(case["silent"]))
    label = torch.nn.utils.rnn.pad_sequence(ts, padding_value=-1, batch_first=True)
    gold = load_golden(name)
    lens_d = torch.tensor(lens, dtype=torch.int32, device=dev)
    loss, idx, lab, perm = ops.pit_loss(pred.to(dev), label.to(dev), lens_d)
    assert idx.cpu().tolist() == gold["min_idx"].tolist()                       # bit-exact assignment
    assert np.array_equal(lab.cpu().numpy().astype(np.int8), gold["label_permute"])
    assert abs(loss.item() - float(gold["loss"])) < 2e-6
    counts = ops.der_counters(pred.to(dev), lab, lens_d).cpu().tolist()
    want = gold["der_counts"].tolist()
    assert counts[0] == round(want[0] * S) and [float(c) for c in counts[1:]] == want[1:]
    # gradient of the winning permutation vs autograd on the oracle formula
    p = pred.clone().requires_grad_(True)
    mask = (torch.arange(T)[None, :] < torch.tensor(lens)[:, None]).float()[:, :, None]
    l = (F.binary_cross_entropy_with_logits(p, lab.cpu().clamp(min=0), reduction="none") * mask).mean(2).sum() / sum(lens)
    l.backward()
    d = ops.pit_loss_bwd(pred.to(dev), lab, lens_d)
    assert (d.cpu() - p.grad).abs().max().item() < 1e-7


def test_pit_edge_cases(dev):
    from oracle.eend_eda_oracle import pit_loss as oracle_pit
    from speaker_diarization_b200 import ops
    g = torch.Generator().manual_seed(99)
    for B, T, S in [(1, 1, 1), (2, 5, 2), (4, 300, 5), (2, 50, 7)]:
        lens = [T] + [max(1, T // 2)] * (B - 1)
        pred = torch.randn((B, T, S), generator=g)
        label = (torch.rand((B, T, S), generator=g) < 0.4).float()
        label[:, :, S // 2] = label[:, :, 0]          # duplicate column -> exact ties
        for b in range(B):
            label[b, lens[b]:] = -1
        loss_o, idx_o, _, lab_o = oracle_pit(pred, label, lens)
        loss, idx, lab, _ = ops.pit_loss(pred.to(dev), label.to(dev), torch.tensor(lens, dtype=torch.int32, device=dev))
        assert idx.cpu().tolist() == idx_o.tolist(), (B, T, S)
        assert torch.equal(lab.cpu(), lab_o)
        assert abs(loss.item() - loss_o.item()) < 1e-5


@pytest.mark.parametrize("lens,n_spk,median", [([7], 2, 1), ([1, 500, 33], 3, 1), ([500, 17, 2000], 4, 11),
                                               ([64, 64], 1, 3), ([4097], 2, 5), ([3, 1, 2], 2, 7)])
def test_rttm_segments_match_make_rttm(dev, lens, n_spk, median):
    """eda_rttm_segments vs oracle/rttm_oracle.py (bin/make_rttm.py:29-35), bit-exact segment lists."""
    import io
    import numpy as np
    from oracle import rttm_oracle as R
    from speaker_diarization_b200 import rttm
    g = torch.Generator().manual_seed(sum(lens) + n_spk + median)
    recs = []
    for n in lens:                      # smooth random posteriors so that segments have realistic lengths
        x = torch.randn((n + 16, n_spk), generator=g).cumsum(0)
        x = (x[16:] - x[:-16]) / 4.0
        recs.append(torch.sigmoid(x).contiguous())
    recs[0][0, 0] = 0.5                 # exactly at the threshold: strict '>'
    want = [R.segments(r.numpy(), 0.5, median) for r in recs]
    got = rttm.segments([r.to(dev) for r in recs], 0.5, median)
    assert got == want
    # uint8 decisions as input (the fused output of eda_posterior) give the same segments
    got8 = rttm.segments([(r > 0.5).to(torch.uint8).to(dev) for r in recs], 0.5, median)
    assert got8 == want
    # RTTM text through the public helper
    names = [f"rec{len(lens) - i}" for i in range(len(lens))]
    wf = io.StringIO()
    rttm.make_rttm(wf, names, [r.to(dev) for r in recs], 0.5, 80, 10, median, 8000)
    assert wf.getvalue() == R.rttm_text(names, [r.numpy() for r in recs], 0.5, 80, 10, median, 8000)


def test_rttm_ragged_speaker_counts(dev):
    from oracle import rttm_oracle as R
    from speaker_diarization_b200 import rttm
    g = torch.Generator().manual_seed(5)
    recs = [torch.rand((50, 3), generator=g), torch.rand((20, 1), generator=g), torch.rand((31, 2), generator=g)]
    got = rttm.segments([r.to(dev) for r in recs], 0.6, 3)
    assert got == [R.segments(r.numpy(), 0.6, 3) for r in recs]


@pytest.mark.parametrize("subsampling,dtype", [(1, torch.int32), (10, torch.float32), (7, torch.int32)])
def test_frame_labels_match_reference_golden(dev, subsampling, dtype):
    """eda_frame_labels vs outputs of the unmodified feature.get_frame_labels (tests/golden/labels.npz),
    all chunks of all cases in ONE launch; subsampled rows = T[::subsampling] (feature.py:126)."""
    import json, os
    import numpy as np
    from speaker_diarization_b200 import labels as LB
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "labels.npz"))
    meta = json.loads(str(G["meta"]))
    for name, m in meta.items():
        if not isinstance(m, dict):
            continue
        segs = list(zip(G[f"{name}/st"].tolist(), G[f"{name}/et"].tolist(), G[f"{name}/spk"].tolist()))
        want = [G[f"{name}/T{i}"] for i in range(len(m["ranges"]))]
        starts = [r[0] for r in m["ranges"]]
        ends = [w.shape[0] if r[1] is None else r[1] for r, w in zip(m["ranges"], want)]
        out, rows = LB.frame_labels([segs] * len(want), starts, ends, [w.shape[0] for w in want], m["n_speakers"],
                                    m["rate"], m["frame_shift"], subsampling, dtype, dev)
        assert out.dtype == dtype
        row = 0
        for w, r in zip(want, rows):
            assert r == w[::subsampling].shape[0]
            assert np.array_equal(out[row:row + r].cpu().numpy().astype(np.int32), w[::subsampling]), name
            row += r


def test_frame_labels_edge_cases(dev):
    from oracle import label_oracle as LO
    from speaker_diarization_b200 import labels as LB
    import numpy as np
    # no segments at all; a chunk with zero frames; half-way rounding (0.005 s * 8000 / 80 = 0.5 -> 0, 1.5 -> 2)
    out, rows = LB.frame_labels([[], [(0.005, 0.015, 1)], [(0.0, 9.0, 0)]], [0, 0, 100], [50, 50, 100], [50, 50, 0], 2,
                                8000, 80, 1, torch.int32, dev)
    assert rows == [50, 50, 0] and out.shape == (100, 2)
    want = np.concatenate([LO.frame_labels([], 0, 50, 50, 2, 8000, 80), LO.frame_labels([(0.005, 0.015, 1)], 0, 50, 50, 2, 8000, 80)])
    assert np.array_equal(out.cpu().numpy(), want) and want[50:53, 1].tolist() == [1, 1, 0]

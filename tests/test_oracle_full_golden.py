"""Pin the CPU oracle at FULL size to what the unmodified reference produced (tests/golden/full_*.npz,
oracle/make_golden_full.py): configs[1] 64 x 500 frames, configs[4] 32 x 2000 frames, and the
reference's own driver loop (infer_eda.py:92-124) on configs[2]-like 16 kHz sessions and
configs[3]-like long recordings whose speaker count (1..6) the attractor threshold decides."""
import numpy as np
import pytest
import torch

from full_cases import driver_case, model_full_case, unpack_bits
from oracle import feature_oracle as fo
from oracle.eend_eda_oracle import OracleEda


@pytest.mark.parametrize("name", ["full_c2", "full_c5"])
def test_oracle_full_batches(name, manifest):
    case, g, cfg, sd, src, ts = model_full_case(name, manifest)
    o = OracleEda(sd, library_lstm=True)
    d = case["data_seed"]
    torch.manual_seed(1000 + d)
    post = torch.stack(o.infer(src, attractor_threshold=-1.0)).numpy()
    want = g["post_all"]
    got = post if case["keep"] is None else post[case["keep"]]
    np.testing.assert_allclose(got, want, atol=2e-5, rtol=0)
    ref_lab = unpack_bits(g["labels_packed"], post.shape)
    flips = (post > 0.5) != ref_lab
    assert int(flips.sum()) <= 2 and bool((np.abs(post[flips] - 0.5) < 2e-5).all())     # fp32 summation order only
    torch.manual_seed(2000 + d)
    loss, stats = o.forward(src, ts)
    assert abs(float(loss) - float(g["loss"])) < 2e-5
    np.testing.assert_allclose(o.trace["pred"].numpy(), g["pred"], atol=2e-4, rtol=0)
    assert o.trace["min_idx"].tolist() == g["pit_min_idx"].tolist()
    assert np.array_equal(o.trace["label_permute"].numpy() > 0.5, unpack_bits(g["label_permute_packed"], g["pred"].shape))
    assert abs(float(stats["der"]) - float(g["stat_der"])) < 1e-4


def _oracle_driver(case, g, sd, recs, attractor_threshold):
    """infer_eda.py:92-124 restated on the oracle: per recording features, per chunk one infer([chunk])."""
    o = OracleEda(sd, library_lstm=True)
    torch.set_rng_state(torch.from_numpy(g["rng_state"]))
    out = {}
    for rec, pcm in recs.items():
        Y = fo.frontend(pcm.astype(np.float64) / 32768.0, case["rate"], case["frame_size"], case["frame_shift"], 7, 10)
        chunks = []
        for s in range(0, len(Y), 2000):
            ys = o.infer([torch.from_numpy(Y[s:s + 2000].copy())], attractor_threshold=attractor_threshold)
            chunks.append(ys[0].numpy())
        out[rec] = np.vstack(chunks)
    return out


@pytest.mark.parametrize("name", ["full_c3_e2e", "full_c4_counts"])
def test_oracle_driver_loop(name, manifest):
    case, g, sd, recs = driver_case(name, manifest)
    out = _oracle_driver(case, g, sd, recs, case["attractor_threshold"])
    for rec, t_hat in out.items():
        want = g["T_hat_" + rec]
        assert t_hat.shape == want.shape, rec            # speaker count per recording included
        np.testing.assert_allclose(t_hat, want, atol=3e-5, rtol=0)
    if name == "full_c4_counts":
        assert sorted(v.shape[1] for v in out.values()) == [1, 2, 3, 4, 5, 6]
        assert float(g["min_margin"]) > 0.2


@pytest.mark.parametrize("name", ["full_c3_e2e", "full_c4_counts"])
def test_drop_in_model_construction_consumes_the_generator_like_the_reference(name, manifest):
    """The driver seeds torch, BUILDS the model (random init draws), loads the checkpoint and only then draws
    the frame shuffles (infer_eda.py:39-113).  The golden stores the generator state the reference had at the
    start of its loop: the drop-in class must leave the generator in that very state, or the same --seed
    would shuffle differently."""
    from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
    from conftest import load_golden
    case, g = manifest["cases"][name], load_golden(name)
    torch.manual_seed(case["seed"])
    TransformerEdaModel(n_speakers=None, in_size=345, n_units=256, n_heads=4, n_layers=case["n_layers"], has_pos=False)
    assert np.array_equal(torch.get_rng_state().numpy(), g["rng_state"])

"""Pin oracle/conformer_oracle.py to the unmodified reference EendEdaModel(encoder_type='conformer')
(tests/golden/conformer_*.npz, oracle/make_golden_conformer.py)."""
import numpy as np
import pytest
import torch

from conftest import cases_of, load_golden
from oracle.conformer_oracle import OracleEendConformer
from oracle.weights import synth_conformer_state_dict, synth_features, synth_labels


@pytest.mark.parametrize("name", cases_of("conformer"))
def test_conformer_oracle_matches_reference(name, manifest):
    case = manifest["cases"][name]
    sd = synth_conformer_state_dict(case["n_layers"], case["weight_seed"])
    src = synth_features(case["lengths"], 345, case["data_seed"])
    ts = synth_labels(case["lengths"], case["n_speakers"], case["data_seed"])
    g = load_golden(name)
    o = OracleEendConformer(sd)
    torch.manual_seed(1000 + case["data_seed"])
    ys = o.infer(src, attractor_threshold=-1.0)
    np.testing.assert_allclose(o.trace["emb"][:, ::7, ::5].numpy(), g["emb_sample"], atol=5e-5, rtol=0)
    for b, y in enumerate(ys):
        np.testing.assert_allclose(y.numpy(), g[f"post_all_{b}"], atol=3e-5, rtol=0)
    torch.manual_seed(2000 + case["data_seed"])
    loss, stats = o.forward(src, ts)
    assert abs(float(loss) - float(g["loss"])) < 3e-5
    assert abs(stats["der"] - float(g["der"])) < 1e-12

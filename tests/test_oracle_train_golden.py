"""Pin the oracle's training step (autograd through the restatement + restated Adam/Noam) to the
gradients / updated parameters the unmodified reference produced (oracle/make_golden_train.py)."""
import numpy as np
import pytest
import torch

from conftest import cases_of, load_golden
from oracle.eend_eda_oracle import adam_step, forward_backward, noam_lr
from oracle.make_golden_train import summary
from oracle.weights import EdaConfig, synth_features, synth_labels, synth_state_dict


def check_summary(got: torch.Tensor, want: np.ndarray, key: str, rtol: float, what: str):
    s = summary(got, key)
    scale = max(want[2], 1e-12) / np.sqrt(max(got.numel(), 1))          # rms magnitude of the tensor
    assert abs(s[2] - want[2]) <= rtol * max(want[2], 1e-9), (what, key, "L2", s[2], want[2])
    assert abs(s[1] - want[1]) <= rtol * max(want[1], 1e-9), (what, key, "abs-sum", s[1], want[1])
    np.testing.assert_allclose(s[3:], want[3:], atol=rtol * 30 * scale + 1e-9, rtol=rtol, err_msg=f"{what} {key}")


@pytest.mark.parametrize("name", cases_of("train"))
def test_oracle_gradients_match_reference(name, manifest):
    case = manifest["cases"][name]
    cfg = EdaConfig(n_layers=case["n_layers"], n_speakers=case["n_speakers"])
    sd = synth_state_dict(cfg, case["weight_seed"])
    src = synth_features(case["lengths"], cfg.in_size, case["data_seed"])
    ts = synth_labels(case["lengths"], case["n_speakers"], case["data_seed"])
    g = load_golden(name)
    torch.manual_seed(3000 + case["data_seed"])
    loss, grads, _ = forward_backward(sd, src, ts)
    assert abs(float(loss) - float(g["loss"])) < 2e-5
    for k in sd:
        check_summary(grads[k], g["g." + k], k, 2e-3, "grad")


def test_oracle_adam_noam_steps_match_reference(manifest):
    (name,) = cases_of("train_steps")
    case = manifest["cases"][name]
    cfg = EdaConfig(n_layers=case["n_layers"], n_speakers=case["n_speakers"])
    p0 = synth_state_dict(cfg, case["weight_seed"])
    params = {k: v.clone() for k, v in p0.items()}
    g = load_golden(name)
    state = {}
    for it in range(case["n_steps"]):
        src = synth_features(case["lengths"], cfg.in_size, case["data_seed"] + it)
        ts = synth_labels(case["lengths"], case["n_speakers"], case["data_seed"] + it)
        torch.manual_seed(4000 + it)
        loss, grads, _ = forward_backward(params, src, ts)
        lr = noam_lr(it, cfg.n_units, case["warmup"])
        assert abs(lr - float(g[f"lr_{it}"])) < 1e-12
        assert abs(float(loss) - float(g[f"loss_{it}"])) < 5e-4
        adam_step(params, grads, state, it + 1, lr)
    for k in params:
        if k.endswith("in_proj_bias"):
            # the key bias has an analytically ZERO gradient (softmax is invariant to it): Adam turns
            # its rounding noise into +-lr steps, so only the size of the update is comparable
            got = summary(params[k] - p0[k], k)
            assert abs(got[2] - g["d." + k][2]) <= 0.05 * g["d." + k][2]
            continue
        check_summary(params[k] - p0[k], g["d." + k], k, 5e-3, "update")

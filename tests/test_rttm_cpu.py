"""Known-answer checks of the RTTM oracle (oracle/rttm_oracle.py restates bin/make_rttm.py:29-42)."""
import numpy as np

from oracle import rttm_oracle as R


def test_segments_known_answer():
    # speaker 0: frames 1-2 and 5; speaker 1: frame 0 and frames 3-6 (end of the recording)
    a = np.array([[0, 1], [1, 0], [1, 0], [0, 1], [0, 1], [1, 1], [0, 1]], dtype=np.float32)
    assert R.segments(a * 0.9, 0.5, 1) == [(0, 1, 3), (0, 5, 6), (1, 0, 1), (1, 3, 7)]
    # strict '>' at the threshold (make_rttm.py:29)
    assert R.segments(np.full((3, 1), 0.5, np.float32), 0.5, 1) == []


def test_median_filter_zero_padding():
    # a single active frame disappears under median 3; two adjacent ones at the edge survive partly
    a = np.zeros((8, 1), np.float32); a[3] = 1
    assert R.segments(a, 0.5, 3) == []
    a = np.zeros((8, 1), np.float32); a[0:2] = 1
    assert R.segments(a, 0.5, 3) == [(0, 0, 2)]         # window at t=0 is (pad, 1, 1) -> 1
    a = np.zeros((8, 1), np.float32); a[0] = 1
    assert R.segments(a, 0.5, 3) == []                  # (pad, 1, 0) -> 0


def test_rttm_text_format_and_order():
    a = np.zeros((10, 2), np.float32); a[2:5, 0] = 1; a[7:, 1] = 1
    txt = R.rttm_text(["b", "a"], [a, a], 0.5, frame_shift=80, subsampling=10, sampling_rate=8000)
    lines = txt.splitlines()
    assert lines[0] == "SPEAKER a 1    0.20    0.30 <NA> <NA> a_0 <NA> <NA>"
    assert lines[1] == "SPEAKER a 1    0.70    0.30 <NA> <NA> a_1 <NA> <NA>"
    assert lines[2].startswith("SPEAKER b 1") and len(lines) == 4


# ---- pinned to the UNMODIFIED bin/make_rttm.py (tests/golden/rttm.json, oracle/make_golden_rttm.py) ----------
import json
import os

import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _rttm_golden():
    with open(os.path.join(GOLD, "rttm.json")) as f:
        return json.load(f)


def _kw(args):
    kw = dict(threshold=0.5, frame_shift=256, subsampling=1, median=1, sampling_rate=16000)
    for k, v in zip(args[::2], args[1::2]):
        kw[k.lstrip("-")] = float(v) if k == "--threshold" else int(v)
    return kw


@pytest.mark.parametrize("name", ["smooth_default", "smooth_med11_8k", "smooth_thr_med5"])
def test_oracle_text_equals_unmodified_make_rttm(name):
    c = _rttm_golden()[name]
    arrs = R.smooth_posteriors(c["lens"], c["n_spk"], c["seed"])
    assert R.rttm_text(c["sessions"], arrs, **_kw(c["args"])) == c["text"]


def test_oracle_text_on_the_reference_drivers_t_hat():
    c = _rttm_golden()["c4_driver"]
    g = np.load(os.path.join(GOLD, "full_c4_counts.npz"))
    assert R.rttm_text(c["sessions"], [g["T_hat_" + s] for s in c["sessions"]], **_kw(c["args"])) == c["text"]


def test_md_eval_der_loop():
    """make_rttm -> md-eval.pl -c 0.25.  The scorer is the reference's vendored NIST perl
    script (third party, not copied into this repo): runs only where /root/reference exists."""
    from oracle import make_golden_rttm as M
    if not os.path.exists(M.MD_EVAL):
        pytest.skip("md-eval.pl lives in /root/reference (build container only)")
    c = _rttm_golden()["c4_driver"]
    with open(os.path.join(GOLD, "manifest.json")) as f:
        truth = M.c4_truth(json.load(f))
    der = M.md_eval(truth, c["text"])
    assert der == c["md_eval_der_percent"]
    # an identical RTTM scores identically; one segment moved by one 0.1 s frame stays far inside 0.1 % abs
    lines = c["text"].splitlines()
    f = lines[5].split()
    f[3] = f"{float(f[3]) + 0.1:.2f}"
    moved = "\n".join(lines[:5] + [" ".join(f)] + lines[6:]) + "\n"
    assert abs(M.md_eval(truth, moved) - der) < 0.1

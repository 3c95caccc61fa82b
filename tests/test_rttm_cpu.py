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

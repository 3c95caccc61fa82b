"""oracle/label_oracle.py and the integer chunking helpers against outputs of the UNMODIFIED
reference (feature.get_frame_labels, feature._count_frames, diarization_dataset._gen_frame_indices)
frozen in tests/golden/labels.npz by oracle/make_golden_labels.py."""
import json
import os

import numpy as np

from oracle import label_oracle as LO
from speaker_diarization_b200 import labels as LB

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "labels.npz"))
META = json.loads(str(G["meta"]))


def chunk_cases():
    for name, m in META.items():
        if not isinstance(m, dict):
            continue
        segs = list(zip(G[f"{name}/st"].tolist(), G[f"{name}/et"].tolist(), G[f"{name}/spk"].tolist()))
        for i, (start, end) in enumerate(m["ranges"]):
            T = G[f"{name}/T{i}"]
            yield name, m, segs, start, (T.shape[0] if end is None else end), T


def test_label_oracle_matches_reference_golden():
    n = 0
    for name, m, segs, start, end, T in chunk_cases():
        got = LO.frame_labels(segs, start, end, T.shape[0], m["n_speakers"], m["rate"], m["frame_shift"])
        assert np.array_equal(got, T), (name, start, end)
        n += 1
    assert n >= 10


def test_spanning_segment_is_dropped_like_the_reference():
    # a segment covering the whole chunk sets nothing (rel_start and rel_end both None, feature.py:239)
    T = LO.frame_labels([(0.0, 100.0, 0)], 300, 600, 300, 1, 8000, 80)
    assert T.sum() == 0
    T = LO.frame_labels([(0.0, 100.0, 0)], 0, 600, 600, 1, 8000, 80)       # starts inside: kept
    assert T.sum() == 600


def test_chunking_helpers_match_reference_golden():
    for c in META["gen_frame_indices"]:
        assert [list(p) for p in LB.gen_frame_indices(*c["args"])] == c["out"], c
    for c in META["count_frames"]:
        assert LB.count_frames(*c["args"]) == c["out"], c

"""TEST INFRASTRUCTURE ONLY: CPU restatement of the labelling loop of feature.get_frame_labels
(feature.py:228-245) and of feature.subsample's label half (:122-127), on plain lists instead
of a KaldiData object.  Pinned: oracle/make_golden_labels.py runs the UNMODIFIED reference
function on a stand-in KaldiData and freezes its outputs in tests/golden/labels.npz.
"""
import numpy as np


def frame_labels(segs, start, end, n_frames, n_speakers, rate, frame_shift, subsampling=1):
    """segs: [(st_sec, et_sec, speaker_index)]; returns (ceil(n_frames / subsampling), n_speakers) int32."""
    T = np.zeros((n_frames, n_speakers), dtype=np.int32)                     # :224
    for st, et, speaker_index in segs:                                      # :228
        start_frame = np.rint(st * rate / frame_shift).astype(int)          # :230-231
        end_frame = np.rint(et * rate / frame_shift).astype(int)            # :232-233
        rel_start = rel_end = None                                          # :234
        if start <= start_frame and start_frame < end:                      # :235-236
            rel_start = start_frame - start
        if start < end_frame and end_frame <= end:                          # :237-238
            rel_end = end_frame - start
        if rel_start is not None or rel_end is not None:                    # :239-240
            T[rel_start:rel_end, speaker_index] = 1
    return T[::subsampling]                                                 # feature.subsample :126

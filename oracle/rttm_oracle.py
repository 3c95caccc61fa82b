"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): CPU restatement of bin/make_rttm.py:29-42.

The reference script reads ``T_hat`` from h5 files (h5py is not installed here); everything after
the read is restated line for line on a numpy array, using the same ``scipy.signal.medfilt`` the
reference calls, so the restatement is pinned to the reference's own arithmetic.
"""
import io

import numpy as np
from scipy.signal import medfilt


def segments(t_hat: np.ndarray, threshold: float = 0.5, median: int = 1):
    """make_rttm.py:29-35 -> [(spkid, start_frame, end_frame)] in the script's print order."""
    a = np.where(t_hat > threshold, 1, 0)                                  # :29
    if median > 1:
        a = medfilt(a, (median, 1))                                        # :30-31
    out = []
    for spkid, frames in enumerate(a.T):                                   # :32
        frames = np.pad(frames, (1, 1), 'constant')                        # :33
        changes, = np.where(np.diff(frames, axis=0) != 0)                  # :34
        for s, e in zip(changes[::2], changes[1::2]):                      # :37
            out.append((spkid, int(s), int(e)))
    return out


def rttm_text(sessions, t_hats, threshold=0.5, frame_shift=256, subsampling=1, median=1, sampling_rate=16000) -> str:
    """make_rttm.py:22-42 with (session, T_hat) pairs instead of h5 files."""
    wf = io.StringIO()
    for session, t_hat in sorted(zip(sessions, t_hats), key=lambda p: p[0]):        # :23
        fmt = "SPEAKER {:s} 1 {:7.2f} {:7.2f} <NA> <NA> {:s} <NA> <NA>"            # :36
        for spkid, s, e in segments(np.asarray(t_hat), threshold, median):
            print(fmt.format(session, s * frame_shift * subsampling / sampling_rate,
                             (e - s) * frame_shift * subsampling / sampling_rate,
                             session + "_" + str(spkid)), file=wf)                  # :38-42
    return wf.getvalue()


def smooth_posteriors(lens, n_spk: int, seed: int):
    """Seeded random posteriors with realistic segment lengths (moving sums of white noise through a sigmoid),
    one float32 (T, n_spk) array per recording; shared by the golden generator and the tests."""
    import torch
    g = torch.Generator().manual_seed(0x577A + seed)
    out = []
    for n in lens:
        x = torch.randn((n + 16, n_spk), generator=g).cumsum(0)
        out.append(torch.sigmoid((x[16:] - x[:-16]) / 4.0).contiguous().numpy())
    out[0][0, 0] = 0.5                  # exactly at the threshold: make_rttm.py:29 uses a strict '>'
    return out


def truth_rttm(session: str, segs) -> str:
    """Ground-truth RTTM of a synthetic mixture from feature_oracle.synth_audio's (speaker, start s, end s)."""
    fmt = "SPEAKER {:s} 1 {:7.2f} {:7.2f} <NA> <NA> {:s} <NA> <NA>\n"
    return "".join(fmt.format(session, st, et - st, f"{session}_true{spk}") for spk, st, et in segs if et > st)

"""Restatement of the two librosa==0.10.2 functions the reference front end calls.

TEST INFRASTRUCTURE (see oracle/__init__.py).

The reference's ``speaker_diarization/feature.py`` does ``import librosa`` (feature.py:7)
and uses exactly ``librosa.stft`` (feature.py:180,183) and ``librosa.filters.mel``
(feature.py:69 for ``logmel23_mn``).  librosa is pinned at 0.10.2 in the reference's
requirements.txt:15, is not vendored under /root/reference and cannot be installed
here, so its published algorithm is restated below in numpy:

* ``stft``: periodic Hann window of ``win_length`` zero-padded (centred) to ``n_fft``;
  ``center=True`` pads the signal with ``n_fft//2`` zeros on both sides
  (``pad_mode="constant"``, the 0.10 default); frame ``t`` starts at padded sample
  ``t*hop``; ``1 + len(y)//hop`` frames; one-sided FFT per frame; result
  ``(1 + n_fft/2, n_frames)``, complex64 for float32 input and complex128 for float64.
* ``filters.mel``: Slaney mel scale (linear below 1 kHz with 200/3 Hz per mel, log above
  with step ln(6.4)/27), ``n_mels + 2`` equally spaced mel points between fmin and fmax,
  triangular filters built from the ramps between neighbouring centre frequencies,
  Slaney area normalisation ``2 / (f[i+2] - f[i])``, float32 output.

``install()`` registers the stub as ``librosa`` in ``sys.modules`` so the reference's
own ``feature.py`` can be imported unmodified by ``oracle/make_golden.py``.
An independent second statement (``torch.stft`` + torchaudio's Slaney filterbank) is
compared with this one in ``tests/test_feature_oracle.py``.
"""
from __future__ import annotations

import sys
import types

import numpy as np


def hann_periodic(win_length: int) -> np.ndarray:
    n = np.arange(win_length, dtype=np.float64)
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * n / win_length)


def padded_window(n_fft: int, win_length: int) -> np.ndarray:
    w = np.zeros(n_fft, dtype=np.float64)
    lpad = (n_fft - win_length) // 2
    w[lpad:lpad + win_length] = hann_periodic(win_length)
    return w


def stft(y, n_fft=2048, hop_length=None, win_length=None, window="hann", center=True,
         dtype=None, pad_mode="constant"):
    if window != "hann" or not center or pad_mode != "constant":
        raise NotImplementedError("stub covers only the reference's call (feature.py:180-184)")
    y = np.asarray(y)
    win_length = n_fft if win_length is None else win_length
    hop_length = win_length // 4 if hop_length is None else hop_length
    real_t = np.float32 if y.dtype == np.float32 else np.float64
    cplx_t = np.complex64 if real_t == np.float32 else np.complex128
    w = padded_window(n_fft, win_length).astype(real_t)
    yp = np.pad(y.astype(real_t), n_fft // 2, mode="constant")
    n_frames = 1 + (len(yp) - n_fft) // hop_length
    idx = np.arange(n_fft)[None, :] + hop_length * np.arange(n_frames)[:, None]
    frames = yp[idx] * w[None, :]
    return np.fft.rfft(frames, axis=1).astype(cplx_t).T


def hz_to_mel(f):
    f = np.asarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    lin = f / f_sp
    with np.errstate(divide="ignore", invalid="ignore"):
        log = min_log_mel + np.log(np.maximum(f, 1e-300) / min_log_hz) / logstep
    return np.where(f >= min_log_hz, log, lin)


def mel_to_hz(m):
    m = np.asarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def mel(*, sr, n_fft, n_mels=128, fmin=0.0, fmax=None, htk=False, norm="slaney", dtype=np.float32):
    # keyword-only like librosa>=0.10: the reference's positional call sites
    # (feature.py:53,61,79,97) raise TypeError there too; only :69 and :115 are valid.
    if htk or norm != "slaney":
        raise NotImplementedError
    fmax = sr / 2.0 if fmax is None else fmax
    fft_f = np.fft.rfftfreq(n_fft, 1.0 / sr)
    mel_f = mel_to_hz(np.linspace(hz_to_mel(fmin), hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fft_f[None, :]
    w = np.zeros((n_mels, 1 + n_fft // 2), dtype=dtype)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        w[i] = np.maximum(0.0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    w *= enorm[:, None]
    return w


def install() -> types.ModuleType:
    """Register this stub as ``librosa`` (idempotent)."""
    if "librosa" in sys.modules and getattr(sys.modules["librosa"], "__eda_stub__", False):
        return sys.modules["librosa"]
    mod = types.ModuleType("librosa")
    mod.__eda_stub__ = True
    mod.__version__ = "0.10.2-restated"
    mod.stft = stft
    filt = types.ModuleType("librosa.filters")
    filt.mel = mel
    mod.filters = filt
    sys.modules["librosa"] = mod
    sys.modules["librosa.filters"] = filt
    return mod

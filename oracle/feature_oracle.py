"""CPU restatement (numpy fp64) of the reference feature front end.

TEST INFRASTRUCTURE (see oracle/__init__.py).
``R`` = /root/reference/speaker_diarization/feature.py.
"""
from __future__ import annotations

import numpy as np

from . import librosa_stub


def get_input_dim(frame_size: int, context_size: int, transform_type: str) -> int:
    """R:10-21."""
    if transform_type.startswith("logmel23"):
        width = 23
    else:
        width = (1 << (frame_size - 1).bit_length()) // 2 + 1
    return (2 * context_size + 1) * width


def stft(data: np.ndarray, frame_size: int = 400, frame_shift: int = 160) -> np.ndarray:
    """R:155-184 — n_fft is the next power of two ≥ frame_size; the last librosa frame is
    dropped iff ``len(data) % frame_shift == 0``.  Returns (frames, bins) complex."""
    n_fft = 1 << (frame_size - 1).bit_length()
    Y = librosa_stub.stft(data, n_fft=n_fft, win_length=frame_size, hop_length=frame_shift).T
    return Y[:-1] if len(data) % frame_shift == 0 else Y


def n_stft_frames(n_samples: int, frame_shift: int) -> int:
    """R:186-193 (``_count_frames``)."""
    n = 1 + n_samples // frame_shift
    return n - 1 if n_samples % frame_shift == 0 else n


def logmel23(Y: np.ndarray, sample_rate: int) -> np.ndarray:
    """R:64-71 — |Y|² · melᵀ → log10(max(·, 1e-10)); float64, before mean removal."""
    n_fft = 2 * (Y.shape[1] - 1)
    basis = librosa_stub.mel(sr=sample_rate, n_fft=n_fft, n_mels=23)
    return np.log10(np.maximum(np.dot(np.abs(Y) ** 2, basis.T), 1e-10))


def transform_logmel23_mn(Y: np.ndarray, sample_rate: int) -> np.ndarray:
    """R:64-73,119 — log-mel minus its per-bin mean over ALL frames passed in."""
    lm = logmel23(Y, sample_rate)
    return (lm - np.mean(lm, axis=0)).astype(np.float32)


def transform(Y: np.ndarray, transform_type, sample_rate: int) -> np.ndarray:
    """R:24-119 for the variants that run under the pinned librosa 0.10.2: None (R:44-45), 'log' (R:46-47),
    'logmel23_mn' (R:64-73), 'logmel23_espnet' (R:105-116); the others call librosa.filters.mel positionally and
    raise TypeError there (R:53,61,79,97)."""
    if transform_type in ("logmel", "logmel23", "logmel23_swn", "logmel23_mvn"):
        raise TypeError("mel() takes 0 positional arguments but 3 were given")
    A = np.abs(Y)
    if not transform_type:
        out = A
    elif transform_type == "log":
        out = np.log(np.maximum(A, 1e-10))
    elif transform_type == "logmel23_mn":
        return transform_logmel23_mn(Y, sample_rate)
    elif transform_type == "logmel23_espnet":
        n_fft = 2 * (Y.shape[1] - 1)
        basis = librosa_stub.mel(sr=sample_rate, n_fft=n_fft, n_mels=23, fmin=0, fmax=sample_rate / 2)
        out = np.log10(np.maximum(1e-10, np.dot(A, basis.T)))
    else:
        raise ValueError("Unknown transform_type: %s" % transform_type)
    return out.astype(np.float32)


def splice(Y: np.ndarray, context_size: int = 0) -> np.ndarray:
    """R:130-152 — row t = concat(rows t-c … t+c) with zero rows beyond the edges."""
    T, F = Y.shape
    pad = np.zeros((T + 2 * context_size, F), dtype=Y.dtype)
    pad[context_size:context_size + T] = Y
    return np.concatenate([pad[k:k + T] for k in range(2 * context_size + 1)], axis=1)


def subsample(Y: np.ndarray, T: np.ndarray, subsampling: int = 1):
    """R:122-127."""
    return Y[::subsampling], T[::subsampling]


def frontend(data: np.ndarray, sample_rate: int, frame_size: int, frame_shift: int,
             context_size: int = 7, subsampling: int = 10) -> np.ndarray:
    """The chain infer_eda.py:94-98 runs per recording → (ceil(frames/sub), 345) float32."""
    Y = transform_logmel23_mn(stft(data, frame_size, frame_shift), sample_rate)
    return np.ascontiguousarray(splice(Y, context_size)[::subsampling])


def frame_labels(segments, n_frames: int, n_speakers: int, rate: int, frame_shift: int,
                 start: int = 0) -> np.ndarray:
    """R:296-322 — rasterise (speaker_index, t_start_s, t_end_s) segments with the
    ``np.rint(t*rate/shift)`` rule into (n_frames, S) int32."""
    end = start + n_frames
    T = np.zeros((n_frames, n_speakers), dtype=np.int32)
    for spk, st, et in segments:
        sf = int(np.rint(st * rate / frame_shift))
        ef = int(np.rint(et * rate / frame_shift))
        rs = sf - start if start <= sf < end else None
        re = ef - start if start < ef <= end else None
        if rs is not None or re is not None:
            T[rs:re, spk] = 1
    return T


def synth_audio(n_samples: int, sample_rate: int, n_speakers: int, seed: int,
                dtype=np.float32):
    """Seeded synthetic mixture: per speaker a 2-state talk/pause Markov
    chain gating noise shaped by a random 2-pole resonator.  Returns (audio, segments)."""
    rng = np.random.default_rng(0xA0D10 + seed)
    mix = np.zeros(n_samples, dtype=np.float64)
    segs = []
    for s in range(n_speakers):
        f0 = rng.uniform(150.0, 0.4 * sample_rate)
        r = rng.uniform(0.90, 0.98)
        a1, a2 = -2.0 * r * np.cos(2 * np.pi * f0 / sample_rate), r * r
        x = rng.standard_normal(n_samples)
        y = np.zeros(n_samples)
        y1 = y2 = 0.0
        # vectorised IIR via scipy if present, else a python loop on short clips
        try:
            from scipy.signal import lfilter
            y = lfilter([1.0], [1.0, a1, a2], x)
        except Exception:  # pragma: no cover
            for i in range(n_samples):
                y[i] = x[i] - a1 * y1 - a2 * y2
                y2, y1 = y1, y[i]
        y /= (np.abs(y).max() + 1e-9)
        gate = np.zeros(n_samples)
        t, talking = 0.0, rng.random() < 0.5
        dur = n_samples / sample_rate
        while t < dur:
            length = rng.exponential(2.0 if talking else 1.5)
            if talking:
                a, b = int(t * sample_rate), min(n_samples, int((t + length) * sample_rate))
                gate[a:b] = 1.0
                segs.append((s, t, min(dur, t + length)))
            t += length
            talking = not talking
        mix += rng.uniform(0.05, 0.5) * gate * y
    mix += 1e-3 * rng.standard_normal(n_samples)
    mix = np.clip(mix, -1.0, 1.0)
    return mix.astype(dtype), segs

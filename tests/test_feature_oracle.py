"""The librosa restatement (oracle/librosa_stub.py) against an INDEPENDENT second statement of the same two
functions: ``torch.stft`` and torchaudio's Slaney filterbank.  librosa==0.10.2 itself (requirements.txt:15 of
the reference) is not installable here, so with respect to librosa proper the front-end oracle stays
"parity unpinned"; this dual restatement is the mitigation DESIGN.md section 4 names."""
import numpy as np
import pytest
import torch

from oracle import librosa_stub as LS

torchaudio = pytest.importorskip("torchaudio")


@pytest.mark.parametrize("sr,frame_size,hop,n", [(8000, 200, 80, 8000 * 2 + 37), (8000, 200, 80, 80 * 50),
                                                  (16000, 400, 160, 16000 + 911), (8000, 200, 80, 79)])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_stft_stub_equals_torch_stft(sr, frame_size, hop, n, dtype):
    rng = np.random.default_rng(n)
    y = rng.standard_normal(n).astype(dtype) * 0.3
    n_fft = 1 << (frame_size - 1).bit_length()
    mine = LS.stft(y, n_fft=n_fft, hop_length=hop, win_length=frame_size)
    ref = torch.stft(torch.from_numpy(y), n_fft, hop_length=hop, win_length=frame_size,
                     window=torch.hann_window(frame_size, periodic=True, dtype=torch.from_numpy(y).dtype),
                     center=True, pad_mode="constant", return_complex=True).numpy()
    assert mine.shape == ref.shape == (n_fft // 2 + 1, 1 + n // hop)
    assert mine.dtype == (np.complex128 if dtype == np.float64 else np.complex64)
    assert np.abs(mine - ref).max() < (1e-12 if dtype == np.float64 else 2e-4)


@pytest.mark.parametrize("sr,n_fft", [(8000, 256), (16000, 512)])
def test_mel_stub_equals_torchaudio_slaney(sr, n_fft):
    mine = LS.mel(sr=sr, n_fft=n_fft, n_mels=23)
    ref = torchaudio.functional.melscale_fbanks(n_fft // 2 + 1, 0.0, sr / 2.0, 23, sr, norm="slaney", mel_scale="slaney").numpy().T
    assert mine.shape == ref.shape == (23, n_fft // 2 + 1) and mine.dtype == np.float32
    assert np.abs(mine - ref).max() < 1e-7
    assert (mine >= 0).all() and (mine.sum(axis=1) > 0).all()

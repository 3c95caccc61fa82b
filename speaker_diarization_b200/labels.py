"""Frame-level speaker labels on the device.

``frame_labels`` is the loop of ``feature.get_frame_labels`` (feature.py:228-245; the same loop
labels ``get_labeledSTFT`` chunks at :310-330) for a whole batch of chunks in one launch, with
``feature.subsample``'s ``T[::subsampling]`` (:122-127) folded in, so training labels never
round-trip through CPU workers.  Kaldi file parsing stays with the caller: it passes, per
chunk, the segments as (start_sec, end_sec, speaker_index) exactly as the reference derives
them from ``kaldi_obj.segments`` / ``utt2spk`` / ``np.unique``.

``count_frames`` and ``gen_frame_indices`` restate the two integer helpers that decide the
chunking (feature.py:186-191, diarization_dataset.py:14-30).
"""
from __future__ import annotations

from typing import Iterator, List, Optional, Sequence, Tuple

import torch

from . import _lib as L


def count_frames(data_len: int, size: int, shift: int) -> int:
    """feature._count_frames (feature.py:186-191): frames of librosa.stft(center=True) with the
    reference's drop-last rule."""
    n_frames = 1 + int(data_len / shift)
    if data_len % shift == 0:
        n_frames = n_frames - 1
    return n_frames


def gen_frame_indices(data_length: int, size: int = 2000, step: int = 2000, use_last_samples: bool = False,
                      label_delay: int = 0, subsampling: int = 1) -> Iterator[Tuple[int, int]]:
    """diarization_dataset._gen_frame_indices (:14-30)."""
    i = -1
    for i in range(int((data_length - size + step) / step)):          # _count_frames (:14-16)
        yield i * step, i * step + size
    if use_last_samples and i * step + size < data_length:
        if data_length - (i + 1) * step - subsampling * label_delay > 0:
            yield (i + 1) * step, data_length


def frame_labels(chunks: Sequence[Sequence[Tuple[float, float, int]]], starts: Sequence[int], ends: Sequence[int],
                 n_frames: Sequence[int], n_speakers: int, rate: int, frame_shift: int, subsampling: int = 1,
                 dtype: torch.dtype = torch.float32, device=None) -> Tuple[torch.Tensor, List[int]]:
    """Labels of a batch of chunks.  ``chunks[c]`` = [(st_sec, et_sec, speaker_index), ...] of the
    recording chunk c comes from, ``starts/ends`` its frame range, ``n_frames[c]`` the rows of its
    label matrix before subsampling (``Y.shape[0]`` / ``_count_frames``).  Returns the stacked
    (sum rows, n_speakers) device tensor (int32 or float32) and the rows per chunk."""
    dev = torch.device("cuda") if device is None else torch.device(device)
    if dev.type != "cuda":
        raise L.EdaError("speaker_diarization_b200 has no CPU path")
    assert dtype in (torch.float32, torch.int32)
    n_chunks = len(chunks)
    rows = [(int(n) + subsampling - 1) // subsampling for n in n_frames]
    row_off = [0]
    for r in rows:
        row_off.append(row_off[-1] + r)
    st, et, spk, cid = [], [], [], []
    for c, segs in enumerate(chunks):
        for a, b, k in segs:
            st.append(float(a)); et.append(float(b)); spk.append(int(k)); cid.append(c)
    n_seg = len(st)
    meta = torch.tensor([list(map(int, starts)), list(map(int, ends)), list(map(int, n_frames))], dtype=torch.int64).to(dev)
    off_d = torch.tensor(row_off, dtype=torch.int64).to(dev)
    out = torch.empty((row_off[-1], n_speakers), dtype=dtype, device=dev)
    if n_seg:
        t_d = torch.tensor([st, et], dtype=torch.float64).to(dev)
        i_d = torch.tensor([spk, cid], dtype=torch.int32).to(dev)
        args = (L.ptr(t_d[0]), L.ptr(t_d[1]), L.ptr(i_d[0]), L.ptr(i_d[1]))
    else:
        args = (None, None, None, None)
    L.call("eda_frame_labels", *args, n_seg, L.ptr(meta[0]), L.ptr(meta[1]), L.ptr(meta[2]), L.ptr(off_d), n_chunks,
           row_off[-1], n_speakers, float(rate), int(frame_shift), int(subsampling), L.ptr(out), n_speakers,
           1 if dtype == torch.float32 else 0, L.stream())
    return out, rows

"""RTTM post-processing, the B200 side of ``bin/make_rttm.py`` (reference lines 22-42).

The threshold (:29), the optional median filter (:30-31) and the run-length step (:32-35) run
in one kernel on the posteriors while they are still in HBM (``eda_rttm_segments``); only the
segment list crosses to the host, where the RTTM text is formatted exactly like :36-42.
"""
from __future__ import annotations

from typing import IO, List, Optional, Sequence, Tuple

import torch

from . import _lib as L, ops

RTTM_FMT = "SPEAKER {:s} 1 {:7.2f} {:7.2f} <NA> <NA> {:s} <NA> <NA>"      # make_rttm.py:36


def segments(t_hat, threshold: float = 0.5, median: int = 1, n_speakers: Optional[Sequence[int]] = None
             ) -> List[List[Tuple[int, int, int]]]:
    """``t_hat``: one (T, n_spk) CUDA tensor or a list of them (one per recording; fp32
    posteriors or uint8 decisions).  Returns, per recording, [(spkid, start_frame, end_frame)]
    ordered by speaker then time - the order make_rttm.py prints them in."""
    recs = [t_hat] if isinstance(t_hat, torch.Tensor) else list(t_hat)
    if not recs:
        return []
    dev = recs[0].device
    if dev.type != "cuda":
        raise L.EdaError("speaker_diarization_b200 has no CPU path")
    n_spk = max(int(r.shape[1]) for r in recs)
    if all(int(r.shape[1]) == n_spk for r in recs):
        flat = recs[0] if len(recs) == 1 else torch.cat(recs, 0)
    else:                                      # ragged speaker counts: pad columns, mask them in the kernel
        flat = torch.zeros((sum(int(r.shape[0]) for r in recs), n_spk), dtype=recs[0].dtype, device=dev)
        row = 0
        for r in recs:
            flat[row:row + r.shape[0], : r.shape[1]] = r
            row += int(r.shape[0])
    flat = flat.contiguous()
    lens = [int(r.shape[0]) for r in recs]
    off = torch.tensor([0] + list(torch.tensor(lens).cumsum(0).tolist()), dtype=torch.int64).to(dev)
    n_rec_spk = torch.tensor([int(r.shape[1]) for r in recs] if n_speakers is None else list(n_speakers),
                             dtype=torch.int32).to(dev)
    cap = max(lens) // 2 + 1
    seg, counts = ops.rttm_segments(flat, off, n_spk, threshold, median, cap, n_rec_spk)
    counts_h = counts.cpu()
    kmax = int(counts_h.max()) if counts_h.numel() else 0
    seg_h = seg[:, :, : max(kmax, 1)].cpu()
    out = []
    for r in range(len(recs)):
        cur = []
        for s in range(n_spk):
            for i in range(int(counts_h[r, s])):
                cur.append((s, int(seg_h[r, s, i, 0]), int(seg_h[r, s, i, 1])))
        out.append(cur)
    return out


def write_rttm(wf: IO[str], session: str, segs: Sequence[Tuple[int, int, int]], frame_shift: int = 256,
               subsampling: int = 1, sampling_rate: int = 16000) -> None:
    """make_rttm.py:36-42 for one session."""
    for spkid, s, e in segs:
        print(RTTM_FMT.format(session, s * frame_shift * subsampling / sampling_rate,
                              (e - s) * frame_shift * subsampling / sampling_rate,
                              session + "_" + str(spkid)), file=wf)


def make_rttm(wf: IO[str], sessions: Sequence[str], t_hats: Sequence[torch.Tensor], threshold: float = 0.5,
              frame_shift: int = 256, subsampling: int = 1, median: int = 1, sampling_rate: int = 16000) -> None:
    """The whole of make_rttm.py for posteriors that are already on the device; sessions are
    written in sorted order like the reference's ``filepaths.sort()`` (:23)."""
    order = sorted(range(len(sessions)), key=lambda i: sessions[i])
    segs = segments([t_hats[i] for i in order], threshold, median)
    for i, sg in zip(order, segs):
        write_rttm(wf, sessions[i], sg, frame_shift, subsampling, sampling_rate)

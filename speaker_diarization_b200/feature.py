"""B200-native feature front end.

Mirrors the part of the reference's ``speaker_diarization/feature.py`` that the EEND-EDA
drivers use (infer_eda.py:46-49,94-98; diarization_dataset.py:93-115):

    Y = feature.stft(data, frame_size, frame_shift)            # feature.py:155-184
    Y = feature.transform(Y, 'logmel23_mn', sample_rate=rate)  # feature.py:64-73
    Y = feature.splice(Y, context_size)                        # feature.py:130-152
    Y = Y[::subsampling]                                       # feature.py:122-127

On the GPU the first two calls are one kernel (audio -> log-mel + per-bin sums) and the last
two another (mean-norm + splice + subsample), so the public surface here is the fused chain:
``extract`` for one recording (reference argument names) and ``FrontEnd`` for batches.
There is no CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence, Union

import numpy as np
import torch

from . import _lib as L

Array = Union[np.ndarray, torch.Tensor]

# audio sample types the kernels read directly (include/eda_b200.h: EDA_AUDIO_*).  int16 is the
# wav file's own PCM: the kernel scales by 1/32768 exactly like soundfile does for the
# reference (kaldi_data.py:82), so shipping int16 halves / quarters the host->device bytes.
_AUDIO_DTYPES = {torch.float32: 0, torch.float64: 1, torch.int16: 2}

# feature.transform variants (feature.py:43-47, 64-73, 105-116) -> EDA_TRANSFORM_* of include/eda_b200.h.  The
# other names ('logmel', 'logmel23', 'logmel23_swn', 'logmel23_mvn') call librosa.filters.mel positionally, which
# raises TypeError under the librosa==0.10.2 the reference pins (requirements.txt:15): they raise here as well.
_TRANSFORMS = {"logmel23_mn": 0, None: 1, "": 1, "log": 2, "logmel23_espnet": 3}
_POSITIONAL_MEL = ("logmel", "logmel23", "logmel23_swn", "logmel23_mvn")


def _transform_id(transform_type):
    if transform_type in _POSITIONAL_MEL:
        raise TypeError(f"transform_type={transform_type!r}: mel() takes 0 positional arguments but 3 were given "
                        "(feature.py calls librosa.filters.mel positionally for this variant; librosa>=0.10 rejects it)")
    if transform_type not in _TRANSFORMS:
        raise ValueError("Unknown transform_type: %s" % transform_type)           # feature.py:118
    return _TRANSFORMS[transform_type]


def get_input_dim(frame_size: int, context_size: int, transform_type: str) -> int:
    """feature.py:10-21."""
    if transform_type and transform_type.startswith("logmel23"):
        width = 23
    else:
        width = (1 << (frame_size - 1).bit_length()) // 2 + 1
    return (2 * context_size + 1) * width


def count_frames(n_samples: int, frame_shift: int) -> int:
    """feature.py:186-193 (``_count_frames``): frames ``stft`` returns."""
    return int(L.load().eda_frontend_num_frames(int(n_samples), int(frame_shift)))


class FrontEnd:
    """Batched ``stft -> logmel23_mn -> splice -> subsample`` on one CUDA device."""

    def __init__(self, sample_rate: int = 8000, frame_size: int = 200, frame_shift: int = 80,
                 context_size: int = 7, subsampling: int = 10, n_mels: int = 23, device="cuda",
                 transform_type="logmel23_mn"):
        self.sample_rate, self.frame_size, self.frame_shift = sample_rate, frame_size, frame_shift
        self.context_size, self.subsampling, self.n_mels = context_size, subsampling, n_mels
        self.transform_type, self.transform = transform_type, _transform_id(transform_type)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise L.EdaError("speaker_diarization_b200 has no CPU path")
        self._plan = C.c_void_p()
        self._off_cache = {}
        with torch.cuda.device(self.device):
            L.call("eda_frontend_plan_create", sample_rate, frame_size, frame_shift, n_mels, C.byref(self._plan))
        self.width = int(L.load().eda_frontend_width(self._plan, self.transform))     # columns of the pass-1 matrix

    def __del__(self):
        try:
            if getattr(self, "_plan", None):
                L.load().eda_frontend_plan_destroy(self._plan)
                self._plan = None
        except Exception:
            pass

    @property
    def feat_dim(self) -> int:
        return (2 * self.context_size + 1) * self.width

    # ------------------------------------------------------------------ batching ------
    def _stage(self, recs, counts=None):
        """Concatenate recordings into one device buffer (one H2D copy) + prefix offsets.
        ``recs`` may also be ONE flat/2-D tensor holding the recordings back to back, with
        ``counts`` giving each recording's sample count (no concatenation copy then)."""
        if counts is not None:
            flat = recs.reshape(-1)
            if flat.dtype not in _AUDIO_DTYPES:
                flat = flat.float()
            audio = flat if flat.is_cuda else flat.to(self.device, non_blocking=True)
            return (audio.contiguous(), _AUDIO_DTYPES[flat.dtype]) + self._offsets([int(c) for c in counts])
        first = torch.as_tensor(recs[0])
        tdt = first.dtype if first.dtype in _AUDIO_DTYPES else torch.float32
        is_f64 = _AUDIO_DTYPES[tdt]
        ns = [int(r.shape[0]) for r in recs]
        if isinstance(first, torch.Tensor) and first.is_cuda:
            audio = torch.cat([r.to(tdt) for r in recs]).contiguous()
        else:
            stage = torch.empty((sum(ns),), dtype=tdt, pin_memory=True)
            pos = 0
            for r, n in zip(recs, ns):
                stage[pos:pos + n] = torch.as_tensor(r).to(tdt)
                pos += n
            audio = stage.to(self.device, non_blocking=True)
        return (audio, is_f64) + self._offsets(ns)

    def _offsets(self, ns):
        key = tuple(ns)
        hit = self._off_cache.get(key)
        if hit is None:
            if len(self._off_cache) > 64:
                self._off_cache.clear()
            hit = self._off_cache[key] = self._offsets_uncached(ns)
        return hit

    def rows_of(self, counts) -> List[int]:
        """Model frames (rows after subsampling) of recordings with these sample counts."""
        return self._offsets([int(c) for c in counts])[2]

    def _offsets_uncached(self, ns):
        frames = [count_frames(n, self.frame_shift) for n in ns]
        rows = [(f + self.subsampling - 1) // self.subsampling for f in frames]
        off = np.zeros((3, len(ns) + 1), dtype=np.int64)
        np.cumsum(ns, out=off[0, 1:]); np.cumsum(frames, out=off[1, 1:]); np.cumsum(rows, out=off[2, 1:])
        offs = torch.from_numpy(off).to(self.device, non_blocking=True)
        return offs, frames, rows

    def logmel(self, recs, counts=None, sum_ranges=None):
        """Pass 1 = stft + transform before the mean removal.  Returns (matrix (total_frames, width) fp32,
        colsum (n_rec, width) fp64 - per-recording column sums, 'logmel23_mn' only, else None -, staging tuple)."""
        with torch.cuda.device(self.device):
            audio, is_f64, offs, frames, rows = self._stage(recs, counts)
            n_rec = len(frames)
            total = int(sum(frames))
            lm = torch.empty((total, self.width), dtype=torch.float32, device=self.device)
            colsum = torch.empty((n_rec, self.width), dtype=torch.float64, device=self.device) if self.transform == 0 else None
            if sum_ranges is not None:
                # slices of recordings (chunk-level sharding): only frames [lo, hi) of each slice enter colsum
                assert self.transform == 0 and len(sum_ranges) == n_rec
                rng = torch.tensor([[int(a), int(b)] for a, b in sum_ranges], dtype=torch.int64).to(self.device)
                L.call("eda_frontend_logmel_ranges", self._plan, L.ptr(audio), int(is_f64), L.ptr(offs[0]), L.ptr(offs[1]),
                       L.ptr(rng), n_rec, total, L.ptr(lm), L.ptr(colsum), L.stream())
            else:
                L.call("eda_frontend_transform", self._plan, L.ptr(audio), int(is_f64), L.ptr(offs[0]), L.ptr(offs[1]),
                       n_rec, total, self.transform, L.ptr(lm), L.ptr(colsum), L.stream())
        return lm, colsum, (offs, frames, rows)

    def splice(self, lm: torch.Tensor, colsum, offs, rows, ld: int = 0) -> torch.Tensor:
        """Pass 2 on a pass-1 matrix: mean removal (colsum / frames per recording; None = none), +-context splice,
        every ``subsampling``-th row.  Returns (sum rows, feat_dim) fp32 (row stride ``ld`` if given)."""
        ld = ld or self.feat_dim
        total_rows = int(sum(rows))
        with torch.cuda.device(self.device):
            feats = torch.empty((total_rows, ld), dtype=torch.float32, device=self.device)
            L.call("eda_frontend_splice_subsample_w", L.ptr(lm), self.width, L.ptr(colsum), L.ptr(offs[1]),
                   L.ptr(offs[2]), len(rows), total_rows, self.context_size, self.subsampling,
                   L.ptr(feats), ld, L.stream())
        return feats[:, : self.feat_dim]

    def features(self, recs, ld: int = 0, counts=None) -> List[torch.Tensor]:
        """The whole chain for a batch of recordings: list of (ceil(frames/sub), feat_dim)
        fp32 CUDA tensors (views into one buffer; row stride ``ld`` if given, e.g. 352)."""
        feats, rows = self.features_flat(recs, ld, counts)
        out, pos = [], 0
        for r in rows:
            out.append(feats[pos:pos + r])
            pos += r
        return out

    def features_flat(self, recs, ld: int = 0, counts=None):
        """Same, as ONE (sum rows, feat_dim) view (row stride ``ld``) plus the row count of each
        recording -- the form ``TransformerEdaModel.infer_device`` consumes without a copy."""
        lm, colsum, (offs, frames, rows) = self.logmel(recs, counts)
        return self.splice(lm, colsum, offs, rows, ld), rows


_frontends = {}


def extract(data: Array, frame_size: int = 200, frame_shift: int = 80, transform_type: str = "logmel23_mn",
            sample_rate: int = 8000, context_size: int = 7, subsampling: int = 10, device="cuda") -> torch.Tensor:
    """One recording through ``stft -> transform -> splice -> [::subsampling]``
    (infer_eda.py:94-98) -> (T, 345) fp32 CUDA tensor."""
    key = (sample_rate, frame_size, frame_shift, context_size, subsampling, str(device), transform_type)
    if key not in _frontends:
        _frontends[key] = FrontEnd(sample_rate, frame_size, frame_shift, context_size, subsampling, device=device,
                                   transform_type=transform_type)
    return _frontends[key].features([data])[0]

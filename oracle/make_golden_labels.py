"""Freeze outputs of the UNMODIFIED reference ``feature.get_frame_labels`` (feature.py:194-246)
into tests/golden/labels.npz.  Run in the build container only (needs /root/reference):
    python oracle/make_golden_labels.py
A stand-in KaldiData supplies ``segments`` (structured array), ``utt2spk`` and ``load_wav``; the
reference's own code does the speaker indexing (np.unique order) and the labelling."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = "/root/reference"
GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")

# name: (rate, frame_size, frame_shift, n_samples, n_speakers, seed, [(start, end or None)])
CASES = {
    "lab_8k": (8000, 200, 80, 8000 * 30 + 17, 3, 1, [(0, None), (0, 1000), (500, 1500), (2900, 3000)]),
    "lab_16k": (16000, 400, 160, 16000 * 20, 4, 2, [(0, None), (100, 700), (1999, 2000)]),
    "lab_span": (8000, 200, 80, 8000 * 12, 2, 3, [(0, None), (300, 600), (0, 1)]),
}


class FakeKaldi:
    def __init__(self, rec, segs, n_samples, rate):
        self.segments = np.array([(rec, f"utt{i}", st, et) for i, (st, et, _) in enumerate(segs)],
                                 dtype=[("rec", "U16"), ("utt", "U16"), ("st", "f8"), ("et", "f8")])
        self.utt2spk = {f"utt{i}": spk for i, (_, _, spk) in enumerate(segs)}
        self.n_samples, self.rate = n_samples, rate

    def load_wav(self, rec, start=0, end=None):
        end = self.n_samples if end is None else min(end, self.n_samples)
        return np.zeros(max(0, end - start), dtype=np.float32), self.rate


def make_segments(rng, n_spk, dur):
    segs = []
    for k in range(n_spk):
        t = rng.uniform(0, 2)
        while t < dur:
            length = rng.exponential(2.0) + 0.05
            segs.append((round(t, 3), round(min(t + length, dur), 3), f"spk{k:02d}"))
            t += length + rng.exponential(1.5)
    # half-way cases for the round-half-to-even rule and one segment spanning everything
    segs.append((0.005, 0.015, "spk00"))
    segs.append((0.0, dur, f"spk{n_spk - 1:02d}"))
    rng.shuffle(segs)
    return [tuple(s) for s in segs]


def main():
    sys.path.insert(0, REF_ROOT)
    sys.path.insert(0, os.path.dirname(HERE))
    from oracle import librosa_stub
    librosa_stub.install()
    from speaker_diarization import feature as ref_feature
    out = {}
    meta = {}
    for name, (rate, fsize, fshift, n_samples, n_spk, seed, ranges) in CASES.items():
        rng = np.random.default_rng(seed)
        segs = make_segments(rng, n_spk, n_samples / rate)
        speakers = sorted({s[2] for s in segs})
        ko = FakeKaldi("rec1", segs, n_samples, rate)
        out[f"{name}/st"] = np.array([s[0] for s in segs]); out[f"{name}/et"] = np.array([s[1] for s in segs])
        out[f"{name}/spk"] = np.array([speakers.index(s[2]) for s in segs], dtype=np.int32)
        for i, (start, end) in enumerate(ranges):
            T = ref_feature.get_frame_labels(ko, "rec1", start, end, fsize, fshift, n_speakers=None)
            out[f"{name}/T{i}"] = T.astype(np.int32)
        meta[name] = dict(rate=rate, frame_size=fsize, frame_shift=fshift, n_samples=n_samples, n_speakers=len(speakers),
                          ranges=[[s, e] for s, e in ranges])
    # diarization_dataset._gen_frame_indices (:19-30), unmodified (soundfile is only needed by the
    # module's file readers: an empty stand-in module lets the import through)
    import types
    sys.modules.setdefault("soundfile", types.ModuleType("soundfile"))
    from speaker_diarization.eend_eda import diarization_dataset as dd
    idx_cases = [(4800, 2000, 2000, True, 0, 10), (4800, 2000, 2000, False, 0, 10), (1999, 2000, 2000, True, 0, 1),
                 (500, 2000, 2000, True, 0, 1), (6001, 2000, 2000, True, 2, 10), (6001, 2000, 2000, True, 0, 1),
                 (18000, 2000, 2000, True, 0, 1), (500, 500, 500, False, 0, 10), (0, 2000, 2000, True, 0, 1)]
    meta["gen_frame_indices"] = [dict(args=list(a), out=[list(p) for p in dd._gen_frame_indices(*a)]) for a in idx_cases]
    meta["count_frames"] = [dict(args=[n, 200, 80], out=int(ref_feature._count_frames(n, 200, 80)))
                            for n in (0, 1, 79, 80, 81, 160, 8000, 8017, 480037)]
    out["meta"] = np.array(json.dumps(meta))
    np.savez_compressed(os.path.join(GOLD, "labels.npz"), **out)
    print("wrote", os.path.join(GOLD, "labels.npz"), {k: v.shape for k, v in out.items() if k.endswith("T0")})


if __name__ == "__main__":
    main()

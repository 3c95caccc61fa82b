"""Recipe for ``oracle/_ref``: the reference's OWN Python files for the hot path, taken unmodified from where they
lie under /root/reference, so that ``bench.py --impl reference`` and ``cpu_baseline`` can time the reference
itself on the GPU box's host cores (``cpu_baseline.kind = "reference"``) instead of the oracle port.

TEST / MEASUREMENT INFRASTRUCTURE (see oracle/__init__.py).  ``oracle/_ref/`` is git-ignored (nothing of the
reference enters the history) but not gpurun-ignored (it travels to the box like the built .so).  Run in the build
container: ``python -m oracle.build_ref``; ``__graft_entry__.build()`` calls it when /root/reference exists.

The reference is pure Python with no setup.py / pyproject.toml, so there is nothing to compile or pip-install
(DESIGN.md section 4); the path needs exactly these files:
    speaker_diarization/feature.py                                  stft / transform / splice / subsample
    speaker_diarization/eend_eda/models.py                          TransformerEdaModel
    speaker_diarization/eend_eda/encoder_decoder_attractor.py       LstmEncoderDedecoderAttractor
``feature.py`` imports librosa (absent): the timed run registers oracle/librosa_stub first, and says so in its JSON line.
"""
from __future__ import annotations

import os
import shutil

REF_ROOT = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
FILES = ["speaker_diarization/__init__.py", "speaker_diarization/feature.py",
         "speaker_diarization/eend_eda/__init__.py", "speaker_diarization/eend_eda/models.py",
         "speaker_diarization/eend_eda/encoder_decoder_attractor.py"]


def available() -> bool:
    return all(os.path.exists(os.path.join(DST, f)) for f in FILES)


def build(verbose: bool = False) -> bool:
    """Copy the files (only when /root/reference exists).  Returns whether oracle/_ref is usable afterwards."""
    if os.path.isdir(REF_ROOT):
        for f in FILES:
            src, dst = os.path.join(REF_ROOT, f), os.path.join(DST, f)
            os.makedirs(os.path.dirname(dst), exist_ok=True)
            shutil.copyfile(src, dst)
            if verbose:
                print("oracle/_ref <-", src)
    return available()


def import_reference():
    """(feature module, TransformerEdaModel) of the unmodified reference in oracle/_ref, with the librosa stand-in."""
    import sys
    from . import librosa_stub
    librosa_stub.install()
    if DST not in sys.path:
        sys.path.insert(0, DST)
    from speaker_diarization import feature
    from speaker_diarization.eend_eda.models import TransformerEdaModel
    return feature, TransformerEdaModel


if __name__ == "__main__":
    print("oracle/_ref usable:", build(verbose=True))

"""Seeded synthetic weights / features / labels (re-exported from the neutral module
``speaker_diarization_b200.synth`` so that bench.py's product arm needs nothing from ``oracle/``).

TEST INFRASTRUCTURE (see oracle/__init__.py)."""
from speaker_diarization_b200.synth import (EdaConfig, state_dict_spec, synth_conformer_state_dict,  # noqa: F401
                                             synth_features, synth_labels, synth_state_dict)

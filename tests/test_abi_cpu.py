"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol that
include/eda_b200.h declares, and the ctypes table covers exactly that set."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "eda_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(eda_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_path():
    syms = header_symbols()
    for need in ("eda_frontend_logmel", "eda_frontend_splice_subsample", "eda_linear", "eda_mha_fwd",
                 "eda_lstm_attractors", "eda_attractor_head", "eda_posterior", "eda_pit_loss",
                 "eda_pit_loss_bwd", "eda_der_counters"):
        assert need in syms


def test_library_exports_every_declared_symbol():
    from speaker_diarization_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        from speaker_diarization_b200.build import build
        build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for s in header_symbols():
        assert hasattr(lib, s), f"libeda_b200.so does not export {s}"
    assert sorted(_lib.SIGNATURES) == header_symbols()
    assert _lib.load().eda_abi_version() == 1


def test_ctypes_table_matches_header_argument_counts():
    """Every prototype in the header has as many parameters as its ctypes signature."""
    from speaker_diarization_b200 import _lib
    text = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "eda_b200.h")).read(), flags=re.S)
    seen = 0
    for m in re.finditer(r"\b(eda_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
        name, args = m.group(1), m.group(2).strip()
        n = 0 if args in ("void", "") else len(args.split(","))
        assert n == len(_lib.SIGNATURES[name][1]), (name, n, len(_lib.SIGNATURES[name][1]))
        seen += 1
    assert seen == len(_lib.SIGNATURES)


def test_noam_schedule_matches_reference_formula():
    """lr_scheduler.py:29-33 with last_epoch = number of scheduler.step() calls so far."""
    from speaker_diarization_b200.eend_eda.training import noam_lr
    assert noam_lr(0, 256, 25000) == noam_lr(1, 256, 25000) == 256 ** -0.5 * 25000 ** -1.5
    assert abs(noam_lr(25000, 256, 25000) - 256 ** -0.5 * 25000 ** -0.5) < 1e-15
    assert noam_lr(100000, 256, 25000) == 256 ** -0.5 * 100000 ** -0.5


def test_host_helpers_without_gpu():
    from speaker_diarization_b200 import feature
    # feature.py:186-193 frame counting incl. the dropped last frame when n % shift == 0
    assert feature.count_frames(24000, 80) == 300
    assert feature.count_frames(24037, 80) == 301
    assert feature.count_frames(79, 80) == 1
    assert feature.get_input_dim(200, 7, "logmel23_mn") == 345
    assert feature.get_input_dim(200, 7, "log") == 15 * 129


def test_model_has_reference_state_dict_keys_and_no_cpu_path():
    import torch
    from oracle.weights import EdaConfig, state_dict_spec
    from speaker_diarization_b200._lib import EdaError
    from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
    m = TransformerEdaModel(2, 345, 4, 256, 2)
    spec = state_dict_spec(EdaConfig(n_layers=2))
    sd = m.state_dict()
    assert list(sd.keys()) == [k for k, *_ in spec]
    for k, shape, *_ in spec:
        assert tuple(sd[k].shape) == tuple(shape), k
    m.eval()
    with pytest.raises(EdaError):      # CPU model: must fail loudly, never fall back
        m.infer([torch.zeros(10, 345)])

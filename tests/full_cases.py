"""Inputs of the full-size golden cases (tests/golden/full_*.npz, oracle/make_golden_full.py), regenerated
from the seeds in the manifest.  Shared by the CPU oracle tests and the GPU parity tests."""
import numpy as np
import torch

from conftest import load_golden


def driver_case(name, manifest):
    """(case dict, golden dict, state_dict, {recording: int16 PCM}) of a 'driver' case."""
    from oracle.make_golden_full import pcm_audio
    from oracle.weights import EdaConfig, synth_state_dict
    case = manifest["cases"][name]
    g = load_golden(name)
    sd = synth_state_dict(EdaConfig(n_layers=case["n_layers"]), case["weight_seed"])
    if "head_w" in g:                      # the fitted existence head of full_c4_counts
        sd["eda.linear.weight"] = torch.from_numpy(g["head_w"]).reshape(1, 256)
        sd["eda.linear.bias"] = torch.from_numpy(g["head_b"]).reshape(1)
    recs = {}
    for rec in case["rec_order"]:          # the driver's processing order (the frame shuffles depend on it)
        spec = case["recs"][rec]
        seconds, seed = spec[0], spec[1]
        n_spk = case["n_speakers"] if "n_speakers" in case else max(int(spec[2]), 1)
        recs[rec] = pcm_audio(seconds, case["rate"], n_spk, seed)
    return case, g, sd, recs


def model_full_case(name, manifest):
    from oracle.weights import EdaConfig, synth_features, synth_labels, synth_state_dict
    case = manifest["cases"][name]
    cfg = EdaConfig(n_layers=case["n_layers"], n_speakers=case["n_speakers"])
    sd = synth_state_dict(cfg, case["weight_seed"])
    src = synth_features(case["lengths"], cfg.in_size, case["data_seed"])
    ts = synth_labels(case["lengths"], case["n_speakers"], case["data_seed"])
    return case, load_golden(name), cfg, sd, src, ts


def unpack_bits(packed, shape):
    return np.unpackbits(packed)[: int(np.prod(shape))].reshape(shape).astype(bool)

"""Seeded synthetic weights / features / labels with the reference's ``state_dict`` key set.

Neutral support code for tests, ``bench.py`` and ``__graft_entry__.smoke()``: it generates INPUTS only
(nothing here computes any part of the hot path, and nothing in the product path imports it).
``oracle/weights.py`` re-exports these names so that the golden generators keep their imports.

The key names and shapes are the ones ``TransformerEdaModel`` exposes
(reference: speaker_diarization/eend_eda/models.py:185-199 for the modules,
encoder_decoder_attractor.py:11-14 for the EDA); ``make_golden.py`` loads the dict
into the unmodified reference model with ``strict=True``, which proves the key set.

Values are "trained-like": ranges follow PyTorch's default initialisers, but every
tensor (including biases, LayerNorm affine terms and each encoder layer) gets its own
draw, so layer-indexing or bias bugs cannot hide behind the reference's identical
deep-copied layers / zero biases.  A private ``torch.Generator`` is used so the global
RNG (which the model's ``torch.randperm`` shuffle consumes, models.py:229-232) is not
touched.
"""
from __future__ import annotations

import math
from collections import OrderedDict
from dataclasses import dataclass

import torch


@dataclass(frozen=True)
class EdaConfig:
    in_size: int = 345
    n_units: int = 256
    n_heads: int = 4
    n_layers: int = 4
    dim_feedforward: int = 2048
    n_speakers: int = 2


def state_dict_spec(cfg: EdaConfig):
    """[(key, shape, kind, scale)] in the order PyTorch reports them."""
    D, F, FF = cfg.n_units, cfg.in_size, cfg.dim_feedforward
    spec = [
        ("encoder.weight", (D, F), "uniform", 0.1),           # models.py:209 initrange
        ("encoder.bias", (D,), "uniform", 0.05),
        ("encoder_norm.weight", (D,), "ln_w", 0.1),
        ("encoder_norm.bias", (D,), "normal", 0.1),
    ]
    for l in range(cfg.n_layers):
        p = f"transformer_encoder.layers.{l}."
        spec += [
            (p + "self_attn.in_proj_weight", (3 * D, D), "uniform", math.sqrt(6.0 / (4 * D))),
            (p + "self_attn.in_proj_bias", (3 * D,), "normal", 0.02),
            (p + "self_attn.out_proj.weight", (D, D), "uniform", 1.0 / math.sqrt(D)),
            (p + "self_attn.out_proj.bias", (D,), "normal", 0.02),
            (p + "linear1.weight", (FF, D), "uniform", 1.0 / math.sqrt(D)),
            (p + "linear1.bias", (FF,), "uniform", 1.0 / math.sqrt(D)),
            (p + "linear2.weight", (D, FF), "uniform", 1.0 / math.sqrt(FF)),
            (p + "linear2.bias", (D,), "uniform", 1.0 / math.sqrt(FF)),
            (p + "norm1.weight", (D,), "ln_w", 0.1),
            (p + "norm1.bias", (D,), "normal", 0.1),
            (p + "norm2.weight", (D,), "ln_w", 0.1),
            (p + "norm2.bias", (D,), "normal", 0.1),
        ]
    k = 1.0 / math.sqrt(D)
    for m in ("encoder", "decoder"):
        spec += [
            (f"eda.{m}.weight_ih_l0", (4 * D, D), "uniform", k),
            (f"eda.{m}.weight_hh_l0", (4 * D, D), "uniform", k),
            (f"eda.{m}.bias_ih_l0", (4 * D,), "uniform", k),
            (f"eda.{m}.bias_hh_l0", (4 * D,), "uniform", k),
        ]
    spec += [("eda.linear.weight", (1, D), "uniform", k), ("eda.linear.bias", (1,), "uniform", k)]
    return spec


def synth_state_dict(cfg: EdaConfig, seed: int) -> "OrderedDict[str, torch.Tensor]":
    g = torch.Generator(device="cpu")
    g.manual_seed(0x5EED0000 + int(seed))
    sd = OrderedDict()
    for key, shape, kind, scale in state_dict_spec(cfg):
        if kind == "uniform":
            t = (torch.rand(shape, generator=g, dtype=torch.float32) * 2.0 - 1.0) * scale
        elif kind == "normal":
            t = torch.randn(shape, generator=g, dtype=torch.float32) * scale
        elif kind == "ln_w":
            t = 1.0 + torch.randn(shape, generator=g, dtype=torch.float32) * scale
        else:  # pragma: no cover
            raise ValueError(kind)
        sd[key] = t.contiguous()
    return sd


def synth_features(lengths, in_size: int, seed: int):
    """Seeded stand-in for spliced log-mel features: list of (T_i, in_size) float32."""
    g = torch.Generator(device="cpu")
    g.manual_seed(0xFEA70000 + int(seed))
    return [torch.randn((int(t), in_size), generator=g, dtype=torch.float32) for t in lengths]


def synth_labels(lengths, n_speakers: int, seed: int, p_active: float = 0.3, silent=()):
    """Seeded 0/1 speaker-activity labels, list of (T_i, S) float32.

    Activity is a run-length process (so labels look like talk spurts rather than
    white noise); speakers listed in ``silent`` are all-zero, which produces the exact
    PIT ties the reference resolves by lowest permutation index.
    """
    g = torch.Generator(device="cpu")
    g.manual_seed(0x1ABE1000 + int(seed))
    out = []
    for t in lengths:
        t = int(t)
        lab = torch.zeros((t, n_speakers), dtype=torch.float32)
        for s in range(n_speakers):
            if s in silent:
                continue
            pos, state = 0, bool(torch.rand((), generator=g) < p_active)
            while pos < t:
                run = int(torch.randint(5, 60, (), generator=g))
                if state:
                    lab[pos:pos + run, s] = 1.0
                pos += run
                state = bool(torch.rand((), generator=g) < (p_active if not state else 0.5))
        out.append(lab)
    return out


def synth_conformer_state_dict(n_layers: int, seed: int, template=None):
    """Seeded "trained-like" weights with the key set of EendEdaModel(encoder_type="conformer")
    (models.py:487-502 + torchaudio ConformerLayer): every tensor gets its own draw; LayerNorm / BatchNorm affine
    terms near (1, 0), running_var in [0.5, 1.5], running_mean ~ N(0, 0.1).  ``template`` (a state_dict of the real
    class, build container only) fixes key order and shapes; without it the shapes are derived here."""
    D, FF, F_IN, K = 256, 2048, 345, 31
    shapes = OrderedDict()
    shapes["linear.weight"], shapes["linear.bias"] = (D, F_IN), (D,)
    shapes["linear_norm.weight"], shapes["linear_norm.bias"] = (D,), (D,)
    for l in range(n_layers):
        p = f"encoder.conformer_layers.{l}."
        for f in ("ffn1", "ffn2"):
            shapes[p + f + ".sequential.0.weight"], shapes[p + f + ".sequential.0.bias"] = (D,), (D,)
            shapes[p + f + ".sequential.1.weight"], shapes[p + f + ".sequential.1.bias"] = (FF, D), (FF,)
            shapes[p + f + ".sequential.4.weight"], shapes[p + f + ".sequential.4.bias"] = (D, FF), (D,)
        shapes[p + "self_attn_layer_norm.weight"], shapes[p + "self_attn_layer_norm.bias"] = (D,), (D,)
        shapes[p + "self_attn.in_proj_weight"], shapes[p + "self_attn.in_proj_bias"] = (3 * D, D), (3 * D,)
        shapes[p + "self_attn.out_proj.weight"], shapes[p + "self_attn.out_proj.bias"] = (D, D), (D,)
        c = p + "conv_module."
        shapes[c + "layer_norm.weight"], shapes[c + "layer_norm.bias"] = (D,), (D,)
        shapes[c + "sequential.0.weight"], shapes[c + "sequential.0.bias"] = (2 * D, D, 1), (2 * D,)
        shapes[c + "sequential.2.weight"], shapes[c + "sequential.2.bias"] = (D, 1, K), (D,)
        shapes[c + "sequential.3.weight"], shapes[c + "sequential.3.bias"] = (D,), (D,)
        shapes[c + "sequential.3.running_mean"], shapes[c + "sequential.3.running_var"] = (D,), (D,)
        shapes[c + "sequential.3.num_batches_tracked"] = ()
        shapes[c + "sequential.5.weight"], shapes[c + "sequential.5.bias"] = (D, D, 1), (D,)
        shapes[p + "final_layer_norm.weight"], shapes[p + "final_layer_norm.bias"] = (D,), (D,)
    k = 1.0 / math.sqrt(D)
    for m in ("encoder", "decoder"):
        for n, s in (("weight_ih_l0", (4 * D, D)), ("weight_hh_l0", (4 * D, D)), ("bias_ih_l0", (4 * D,)), ("bias_hh_l0", (4 * D,))):
            shapes[f"eda.{m}.{n}"] = s
    shapes["eda.linear.weight"], shapes["eda.linear.bias"] = (1, D), (1,)
    if template is not None:
        assert {kk: tuple(v.shape) for kk, v in template.items()} == dict(shapes), "conformer key set / shapes differ"
    g = torch.Generator(device="cpu")
    g.manual_seed(0xC0F00000 + int(seed))
    sd = OrderedDict()
    for key, shape in shapes.items():
        if key.endswith("num_batches_tracked"):
            sd[key] = torch.tensor(100, dtype=torch.long)
        elif key.endswith("running_var"):
            sd[key] = 0.5 + torch.rand(shape, generator=g)
        elif key.endswith("running_mean"):
            sd[key] = torch.randn(shape, generator=g) * 0.1
        elif key.endswith("norm.weight") or key.endswith("sequential.0.weight") and len(shape) == 1 or key.endswith("sequential.3.weight"):
            sd[key] = 1.0 + torch.randn(shape, generator=g) * 0.1
        elif len(shape) == 1:
            sd[key] = torch.randn(shape, generator=g) * 0.05
        else:
            fan_in = shape[1] * (shape[2] if len(shape) == 3 else 1)
            sd[key] = (torch.rand(shape, generator=g) * 2.0 - 1.0) / math.sqrt(fan_in)
        sd[key] = sd[key].contiguous()
    return sd

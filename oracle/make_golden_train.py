"""Freeze gradients of the UNMODIFIED reference model (train mode, dropout 0) as fixtures.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Run in the build container only
(``python -m oracle.make_golden_train``): it imports /root/reference.  For every case the
reference ``TransformerEdaModel`` is built with ``dropout=0`` (R/models.py:161), put in
``train()`` mode, run as ``loss, stats = model(src, ts); loss.backward()``
(R/train_eda.py:182-184), and for every parameter the fixture keeps the gradient's sum,
absolute sum, L2 norm (float64) and 24 entries at seeded positions - a few hundred bytes per
tensor instead of megabytes.  One case also takes three Adam + Noam steps
(R/train_eda.py:131,137,188-192) and stores the same summaries of the accumulated parameter updates.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

REF_ROOT = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")

TRAIN_CASES = {
    # name: (n_layers, lengths, n_speakers, weight_seed, data_seed)
    "train_l1_ragged": (1, [130, 77, 128], 2, 3, 31),
    "train_l2_s3": (2, [300, 211], 3, 777, 32),
    "train_l4_t500": (4, [500, 500, 377], 2, 777, 33),
}
STEP_CASE = ("train_steps_l2", 2, [200, 150, 64], 2, 4, 34, 3, 100)   # + n_steps, warmup


def summary(t: torch.Tensor, key: str):
    """(sum, abs-sum, L2 norm) in float64 + 24 sampled entries at positions seeded by the key."""
    a = t.detach().double().reshape(-1)
    g = torch.Generator().manual_seed(abs(hash_key(key)) % (2 ** 31))
    idx = torch.randint(0, a.numel(), (24,), generator=g)
    return np.concatenate([[a.sum().item(), a.abs().sum().item(), a.pow(2).sum().sqrt().item()], a[idx].numpy()])


def hash_key(key: str) -> int:
    h = 0
    for ch in key:
        h = (h * 131 + ord(ch)) % 1000000007
    return h


def main() -> None:
    sys.path.insert(0, REF_ROOT)
    sys.path.insert(0, os.path.dirname(HERE))
    from oracle import librosa_stub
    from oracle.weights import EdaConfig, synth_state_dict, synth_features, synth_labels
    librosa_stub.install()
    from speaker_diarization.eend_eda.models import TransformerEdaModel
    from speaker_diarization.eend_eda.lr_scheduler import NoamScheduler

    torch.set_num_threads(os.cpu_count() or 1)
    with open(os.path.join(GOLD, "manifest.json")) as f:
        manifest = json.load(f)
    for name, (L, lens, S, wseed, dseed) in TRAIN_CASES.items():
        cfg = EdaConfig(n_layers=L, n_speakers=S)
        sd = synth_state_dict(cfg, wseed)
        model = TransformerEdaModel(S, cfg.in_size, cfg.n_heads, cfg.n_units, L, dropout=0.0)
        model.load_state_dict(sd, strict=True)
        model.train()
        src, ts = synth_features(lens, cfg.in_size, dseed), synth_labels(lens, S, dseed)
        torch.manual_seed(3000 + dseed)
        loss, stats = model(src, ts)
        loss.backward()
        out = {"loss": np.float64(loss.item())}
        for k, p in model.named_parameters():
            g = p.grad if p.grad is not None else torch.zeros_like(p)
            out["g." + k] = summary(g, k)
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
        manifest["cases"][name] = dict(kind="train", n_layers=L, lengths=lens, n_speakers=S,
                                       weight_seed=wseed, data_seed=dseed)
        print(name, "loss", loss.item(), "grad L2 enc.w", out["g.encoder.weight"][2])

    name, L, lens, S, wseed, dseed, n_steps, warmup = STEP_CASE
    cfg = EdaConfig(n_layers=L, n_speakers=S)
    model = TransformerEdaModel(S, cfg.in_size, cfg.n_heads, cfg.n_units, L, dropout=0.0)
    model.load_state_dict(synth_state_dict(cfg, wseed), strict=True)
    model.train()
    p0 = {k: p.detach().clone() for k, p in model.named_parameters()}
    opt = torch.optim.Adam(model.parameters(), lr=1.0, betas=(0.9, 0.98), eps=1e-9)
    sched = NoamScheduler(opt, cfg.n_units, warmup_steps=warmup)
    out = {}
    opt.zero_grad()
    for it in range(n_steps):
        src, ts = synth_features(lens, cfg.in_size, dseed + it), synth_labels(lens, S, dseed + it)
        torch.manual_seed(4000 + it)
        loss, _ = model(src, ts)
        loss.backward()
        out[f"lr_{it}"] = np.float64(opt.param_groups[0]["lr"])
        opt.step()
        opt.zero_grad()
        sched.step()
        out[f"loss_{it}"] = np.float64(loss.item())
    for k, p in model.named_parameters():
        out["d." + k] = summary(p.detach() - p0[k], k)     # the accumulated UPDATE, not the parameter
    np.savez_compressed(os.path.join(GOLD, name + ".npz"), **out)
    manifest["cases"][name] = dict(kind="train_steps", n_layers=L, lengths=lens, n_speakers=S, weight_seed=wseed,
                                   data_seed=dseed, n_steps=n_steps, warmup=warmup)
    print(name, [out[f"loss_{i}"] for i in range(n_steps)], [out[f"lr_{i}"] for i in range(n_steps)])
    with open(os.path.join(GOLD, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()

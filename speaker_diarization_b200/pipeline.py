"""Overlapped EEND-EDA inference over a stream of recording batches.

What the reference's ``infer_eda.py:92-124`` loop does one chunk at a time on the default
stream (load -> feature.stft/transform/splice/subsample on the CPU -> H2D -> model.infer ->
D2H), restructured for a B200: CUDA streams for audio H2D, compute and result D2H chained
with events, ``depth`` batches in flight, results landing in pinned host buffers.  The host
never blocks on the batch it has just submitted, so its own work (drawing the reference's
``torch.randperm`` frame shuffles, staging, kernel launches) overlaps the GPU's.

Consecutive batches alternate between ``compute_streams`` compute streams (each with its own
model workspace): the attractor LSTM is a latency-bound recurrence that occupies 104 of the
148 SMs at low utilisation for about a third of a batch, and with two streams the next
batch's front end and attention CTAs run on the SMs it leaves idle.

    pipe = InferencePipeline(model, frontend)
    for result in pipe.run(batches):          # batches: iterable of (audio, counts)
        result["posteriors"]                  # (B, Tmax, 14) pinned host tensor
        result["n_speakers"]                  # list[int], attractor-threshold speaker counts

Semantics per batch are those of ``feature.extract`` + ``TransformerEdaModel.infer`` on the
batch's recordings taken as one padded batch (models.py:297-347); the permutation draws happen
at ``submit`` time, in submission order, from torch's global CPU generator.
"""
from __future__ import annotations

from collections import deque
from typing import Iterable, List, Optional, Sequence

import numpy as np
import torch

from . import _lib as L


class InferencePipeline:
    def __init__(self, model, frontend, max_n_speakers: int = 15, attractor_threshold: float = 0.5,
                 activity_threshold: float = 0.5, depth: int = 3, ld: int = 352, compute_streams: int = 3):
        self.model, self.fe = model, frontend
        self.max_n_speakers, self.attractor_threshold = max_n_speakers, attractor_threshold
        self.activity_threshold, self.depth, self.ld = activity_threshold, max(1, depth), ld
        self.device = frontend.device
        if self.device.type != "cuda":
            raise L.EdaError("speaker_diarization_b200 has no CPU path")
        with torch.cuda.device(self.device):
            self.s_in, self.s_out = torch.cuda.Stream(), torch.cuda.Stream()
            self.s_computes = [torch.cuda.Stream() for _ in range(max(1, min(compute_streams, self.depth)))]
        self.s_compute = self.s_computes[0]
        self._n_submitted = 0
        self._slots = [dict(busy=None) for _ in range(self.depth + 1)]
        self._next = 0
        self._pending = deque()
        self.gather = None          # optional callable(post_dev) run on the compute stream (multi-GPU collection)

    # ------------------------------------------------------------------------------------
    def submit(self, audio, counts: Optional[Sequence[int]] = None) -> None:
        """Enqueue one batch.  ``audio``: one (n_rec, n) / flat tensor with ``counts`` samples per
        recording, or a list of 1-D arrays/tensors.  Host tensors should be pinned."""
        if len(self._pending) >= self.depth:
            raise RuntimeError("pipeline full: call collect() first")
        slot = self._slots[self._next]
        self._next = (self._next + 1) % len(self._slots)
        if slot["busy"] is not None:
            slot["busy"].synchronize()            # its previous results have been handed out long ago
        dev = self.device
        cur = torch.cuda.current_stream(dev)
        k = self._n_submitted % len(self.s_computes)
        self._n_submitted += 1
        s_compute = self.s_computes[k]
        self.model.workspace_slot = k
        if isinstance(audio, torch.Tensor) and counts is not None:
            if audio.is_cuda:
                s_compute.wait_stream(cur)        # whoever produced it
                dev_audio = audio
            else:
                with torch.cuda.stream(self.s_in):
                    dev_audio = audio.to(dev, non_blocking=True)
                s_compute.wait_stream(self.s_in)
        else:
            dev_audio = audio                     # list of recordings: FrontEnd stages them itself
        with torch.cuda.stream(s_compute), L.use_stream(s_compute):
            feats, rows = self.fe.features_flat(dev_audio, ld=self.ld, counts=counts)
            post, probs, labels = self.model.infer_device(feats, rows, self.max_n_speakers, self.activity_threshold)
            if self.gather is not None:
                self.gather(post)
            if isinstance(dev_audio, torch.Tensor):
                dev_audio.record_stream(s_compute)
        B, Tmax, n_att = post.shape
        key = (B, Tmax, n_att)
        if slot.get("key") != key:
            slot.update(key=key,
                        h_post=torch.empty((B, Tmax, n_att), dtype=torch.float32, pin_memory=True),
                        h_probs=torch.empty((B, self.max_n_speakers), dtype=torch.float32, pin_memory=True),
                        h_labels=torch.empty((B, Tmax, n_att), dtype=torch.uint8, pin_memory=True))
        self.s_out.wait_stream(s_compute)
        with torch.cuda.stream(self.s_out):
            slot["h_post"].copy_(post, non_blocking=True)
            slot["h_probs"].copy_(probs, non_blocking=True)
            slot["h_labels"].copy_(labels, non_blocking=True)
            for t in (post, probs, labels):
                t.record_stream(self.s_out)
            done = torch.cuda.Event()
            done.record(self.s_out)
        slot["busy"] = done
        self._pending.append((slot, done, list(rows)))

    def collect(self) -> dict:
        """Block until the OLDEST submitted batch is on the host and return it."""
        slot, done, rows = self._pending.popleft()
        done.synchronize()
        below = slot["h_probs"].numpy() < self.attractor_threshold          # models.py:340-343, all rows at once
        first = below.argmax(axis=1)
        n_spk = [int(f) if hit else self.max_n_speakers - 1 for f, hit in zip(first, below.any(axis=1))]
        return dict(posteriors=slot["h_post"], probs=slot["h_probs"], labels=slot["h_labels"],
                    n_speakers=n_spk, lengths=rows)

    def run(self, batches: Iterable):
        """Generator over results, keeping ``depth`` batches in flight."""
        for item in batches:
            audio, counts = item if isinstance(item, tuple) else (item, None)
            if len(self._pending) >= self.depth:
                yield self.collect()
            self.submit(audio, counts)
        while self._pending:
            yield self.collect()

    def drain(self) -> List[dict]:
        out = []
        while self._pending:
            out.append(self.collect())
        return out

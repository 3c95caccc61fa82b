"""Overlapped EEND-EDA inference over a stream of recording batches.

What the reference's ``infer_eda.py:92-124`` loop does one chunk at a time on the default
stream (load -> feature.stft/transform/splice/subsample on the CPU -> H2D -> model.infer([chunk])
-> D2H -> np.vstack), restructured for a B200: CUDA streams for audio H2D, compute and result D2H
chained with events, ``depth`` batches in flight, results landing in pinned host buffers.  The host
never blocks on the batch it has just submitted, so its own work (drawing the reference's
``torch.randperm`` frame shuffles, staging, kernel launches) overlaps the GPU's.

Semantics are the reference DRIVER's: every recording goes through the front end as a whole
(recording-wide mean normalisation, feature.py:72-73), its frames are cut into ``chunk_size``-frame
chunks (infer_eda.py:21-27), and every chunk is diarized as if it were alone in the model call
(``independent=True``: attention sees only the chunk's own frames, which is what a batch of one
computes).  The frame shuffles are drawn from torch's global CPU generator at ``submit`` time in the
driver's order (recording by recording, chunk by chunk), so the same ``torch.manual_seed`` gives the
reference driver's ``T_hat``.  On the device the chunks of ALL recordings of a batch run together in at
most two model passes: the full ``chunk_size`` chunks, and the ragged tails padded to the longest tail.

    pipe = InferencePipeline(model, frontend, chunk_size=2000)
    for result in pipe.run(batches):          # batches: iterable of (audio, counts)
        result["posteriors"]                  # (sum of frames, 14) pinned host tensor, recordings back to back
        result["rec_offsets"]                 # recording r = rows [rec_offsets[r], rec_offsets[r+1])
        result["n_speakers"]                  # per chunk, attractor-threshold speaker counts (models.py:340-343)
        InferencePipeline.t_hats(result)      # per recording: the (T, n_spk) matrix infer_eda.py:121 writes

Result tensors are freshly allocated (torch's pinned caching allocator) for every batch and belong
to the caller: they stay valid however many batches are submitted afterwards.

Consecutive batches alternate between ``compute_streams`` compute streams (each with its own
model workspace): the attractor LSTM is a latency-bound recurrence that occupies part of the SMs
at low utilisation for about a quarter of a batch, and with several streams the next batch's front
end and attention CTAs run on the SMs it leaves idle.
"""
from __future__ import annotations

import time
from collections import deque
from typing import Iterable, List, Optional, Sequence

import numpy as np
import torch

from . import _lib as L


def chunk_plan(rows: Sequence[int], chunk_size: Optional[int]):
    """The driver's chunking (infer_eda.py:21-27, 102) over a batch of recordings whose features lie back
    to back: [(recording, first row in the flat matrix, length)] in the driver's processing order."""
    out, base = [], 0
    for r, n in enumerate(rows):
        step = chunk_size if chunk_size and chunk_size > 0 else max(n, 1)
        for s in range(0, n, step):
            out.append((r, base + s, min(step, n - s)))
        base += n
    return out


class InferencePipeline:
    def __init__(self, model, frontend, max_n_speakers: int = 15, attractor_threshold: float = 0.5,
                 activity_threshold: float = 0.5, depth: int = 3, ld: int = 352, compute_streams: int = 3,
                 chunk_size: Optional[int] = 2000, independent: bool = True):
        self.model, self.fe = model, frontend
        self.max_n_speakers, self.attractor_threshold = max_n_speakers, attractor_threshold
        self.activity_threshold, self.depth, self.ld = activity_threshold, max(1, depth), ld
        self.chunk_size, self.independent = chunk_size, independent
        self.device = frontend.device
        if self.device.type != "cuda":
            raise L.EdaError("speaker_diarization_b200 has no CPU path")
        with torch.cuda.device(self.device):
            self.s_in, self.s_out = torch.cuda.Stream(), torch.cuda.Stream()
            self.s_computes = [torch.cuda.Stream() for _ in range(max(1, min(compute_streams, self.depth)))]
            # kernel-side copies of the weights are made here, once, and every stream is ordered after them
            model.prepare_weights()
            torch.cuda.synchronize(self.device)
        self.s_compute = self.s_computes[0]
        self._n_submitted = 0
        # host seconds spent inside submit() (launch work, shuffle draws), of which in the gather hook / result staging
        self.stats = dict(submit_s=0.0, submits=0, gather_s=0.0, results_s=0.0)
        self._pending = deque()
        self._plans = {}
        # optional callable(post_dev, mask_dev) run on the compute stream (multi-GPU result collection): posteriors
        # (rows, 14) fp32 and the thresholded decisions as (rows,) int16 bit masks (2 bytes per frame)
        self.gather = None

    # ------------------------------------------------------------------------------------
    def _plan(self, rows: Sequence[int]):
        """Chunk list, length buckets and device-side offsets of a batch geometry (cached per geometry)."""
        key = tuple(rows)
        p = self._plans.get(key)
        if p is None:
            if len(self._plans) > 64:
                self._plans.clear()
            chunks = chunk_plan(rows, self.chunk_size)
            full = [i for i, c in enumerate(chunks) if self.chunk_size and c[2] == self.chunk_size]
            tails = [i for i, c in enumerate(chunks) if not (self.chunk_size and c[2] == self.chunk_size)]
            if not self.independent:
                full, tails = list(range(len(chunks))), []        # one padded batch, like model.infer on a list
            buckets = [b for b in (full, tails) if b]
            where = {}
            for g, b in enumerate(buckets):
                for j, i in enumerate(b):
                    where[i] = (g, j)
            rec_off = np.zeros(len(rows) + 1, dtype=np.int64)
            np.cumsum(rows, out=rec_off[1:])
            p = dict(chunks=chunks, buckets=buckets, order=[where[i] for i in range(len(chunks))],
                     lens=[[chunks[i][2] for i in b] for b in buckets], offs=[[chunks[i][1] for i in b] for b in buckets],
                     inv=np.concatenate([np.asarray(b, dtype=np.int64) for b in buckets]).argsort(),
                     rec_off=rec_off, chunk_rec=[c[0] for c in chunks], lengths=[c[2] for c in chunks])
            self._plans[key] = p
        return p

    def submit(self, audio, counts: Optional[Sequence[int]] = None) -> None:
        """Enqueue one batch.  ``audio``: one (n_rec, n) / flat tensor with ``counts`` samples per
        recording, or a list of 1-D arrays/tensors.  Host tensors should be pinned."""
        if len(self._pending) >= self.depth:
            raise RuntimeError("pipeline full: call collect() first")
        t_host = time.perf_counter()
        dev = self.device
        cur = torch.cuda.current_stream(dev)
        k = self._n_submitted % len(self.s_computes)
        self._n_submitted += 1
        s_compute = self.s_computes[k]
        self.model.workspace_slot = k
        perms = p = None
        if isinstance(audio, torch.Tensor) and counts is not None:
            if audio.is_cuda:
                s_compute.wait_stream(cur)        # whoever produced it
                dev_audio = audio
            else:
                with torch.cuda.stream(self.s_in):
                    # The frame shuffles go up FIRST, on the copy stream, ahead of this batch's audio: uploaded from the
                    # compute stream they become eligible only after the batch's encoder, by which time the 51 MB audio
                    # copies of the NEXT batches occupy the H2D engine and the LSTM waits ~2 ms per queued batch for a
                    # 128 KB copy (e2e 14.2 -> 11.9 M frames/s once the host had become fast enough to run that far ahead).
                    p = self._plan(self.fe.rows_of(counts))
                    perms = self.model.draw_perms(p["lens"], p["order"], dev)       # the driver's draw order
                    dev_audio = audio.to(dev, non_blocking=True)
                s_compute.wait_stream(self.s_in)
                for t in perms:
                    t.record_stream(s_compute)
        else:
            dev_audio = audio                     # list of recordings: FrontEnd stages them itself
        n_att = self.max_n_speakers - 1
        with torch.cuda.stream(s_compute), L.use_stream(s_compute):
            feats, rows = self.fe.features_flat(dev_audio, ld=self.ld, counts=counts)
            if p is None:
                p = self._plan(rows)
            total = int(p["rec_off"][-1])
            post = torch.empty((total, n_att), dtype=torch.float32, device=dev)
            labels = torch.empty((total, n_att), dtype=torch.uint8, device=dev)
            mask = torch.empty((total,), dtype=torch.int16, device=dev) if self.gather is not None else None
            if perms is None:
                perms = self.model.draw_perms(p["lens"], p["order"], dev)       # the driver's draw order
            probs = []
            for lens, offs, perm in zip(p["lens"], p["offs"], perms):
                _, pr, _ = self.model.infer_device(feats, lens, self.max_n_speakers, self.activity_threshold,
                                                   independent=self.independent, offsets=offs, packed_out=(post, labels, mask),
                                                   perm_dev=perm)
                probs.append(pr)
            probs = probs[0] if len(probs) == 1 else torch.cat(probs)
            if self.gather is not None:
                t_g = time.perf_counter()
                self.gather(post, mask)
                self.stats["gather_s"] += time.perf_counter() - t_g
            if isinstance(dev_audio, torch.Tensor):
                dev_audio.record_stream(s_compute)
        # results belong to the caller: fresh pinned buffers per batch (the pinned caching allocator recycles a
        # block only after the copy that last touched it has completed AND the caller dropped the tensor)
        t_r = time.perf_counter()
        h_post = torch.empty((total, n_att), dtype=torch.float32, pin_memory=True)
        h_labels = torch.empty((total, n_att), dtype=torch.uint8, pin_memory=True)
        h_probs = torch.empty((len(p["chunks"]), self.max_n_speakers), dtype=torch.float32, pin_memory=True)
        self.s_out.wait_stream(s_compute)
        with torch.cuda.stream(self.s_out):
            h_post.copy_(post, non_blocking=True)
            h_probs.copy_(probs, non_blocking=True)
            h_labels.copy_(labels, non_blocking=True)
            for t in (post, probs, labels):
                t.record_stream(self.s_out)
            done = torch.cuda.Event()
            done.record(self.s_out)
        self._pending.append((done, p, h_post, h_probs, h_labels, list(rows)))
        self.stats["results_s"] += time.perf_counter() - t_r
        self.stats["submit_s"] += time.perf_counter() - t_host
        self.stats["submits"] += 1

    def collect(self) -> dict:
        """Block until the OLDEST submitted batch is on the host and return it (see the module docstring)."""
        done, p, h_post, h_probs, h_labels, rows = self._pending.popleft()
        done.synchronize()
        if len(p["buckets"]) > 1:                                  # back to the driver's chunk order
            h_probs = h_probs[torch.from_numpy(p["inv"])]
        below = h_probs.numpy() < self.attractor_threshold          # models.py:340-343, all chunks at once
        first = below.argmax(axis=1)
        n_spk = [int(f) if hit else self.max_n_speakers - 1 for f, hit in zip(first, below.any(axis=1))]
        return dict(posteriors=h_post, probs=h_probs, labels=h_labels, n_speakers=n_spk, lengths=p["lengths"],
                    chunk_rec=p["chunk_rec"], rec_offsets=p["rec_off"], rec_lengths=rows)

    @staticmethod
    def t_hats(result: dict) -> List[torch.Tensor]:
        """Per recording the matrix the driver stores as ``T_hat`` (infer_eda.py:113-124): the chunks' (T, n_spk)
        posteriors stacked on the time axis.  Like ``np.vstack`` there, it refuses chunks that disagree on n_spk."""
        out = []
        off = result["rec_offsets"]
        for r in range(len(off) - 1):
            ns = {n for n, cr in zip(result["n_speakers"], result["chunk_rec"]) if cr == r}
            if len(ns) > 1:
                raise ValueError(f"recording {r}: chunks decided different speaker counts {sorted(ns)}; the reference's "
                                 "np.vstack(out_chunks) fails on this too (infer_eda.py:121)")
            n = ns.pop() if ns else 0
            out.append(result["posteriors"][int(off[r]):int(off[r + 1]), :n])
        return out

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

"""Multi-GPU partitioning of the EEND-EDA inference path.

Chunks and recordings are independent in the reference (infer_eda.py:92-113: one recording at
a time, ``chunk_size`` frames per model call, no carried state), so the path shards by
recording with NO data-path collective; the only exchange is result collection on rank 0
(SURVEY 8e).  One process per GPU, ``torch.distributed`` (NCCL on the GPUs, gloo in the CPU
tests) for that one gather.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch
import torch.distributed as dist


def gen_chunk_indices(data_len: int, chunk_size: int) -> List[Tuple[int, int]]:
    """infer_eda.py:21-27 (``_gen_chunk_indices``): [start, end) of each model call."""
    out, start = [], 0
    step = chunk_size if chunk_size and chunk_size > 0 else max(data_len, 1)
    while start < data_len:
        out.append((start, min(start + step, data_len)))
        start += step
    return out


def shard_recordings(lengths: Sequence[int], world_size: int) -> List[List[int]]:
    """Longest-first greedy assignment of recordings (by frame count) to ranks: the scaling
    limit of this path is load balance, not bandwidth.  Deterministic; returns the recording
    indices of every rank in processing order."""
    order = sorted(range(len(lengths)), key=lambda i: (-int(lengths[i]), i))
    load = [0] * world_size
    shards: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        shards[r].append(i)
        load[r] += int(lengths[i])
    return shards


def gather_posteriors(local: Dict[int, torch.Tensor], n_cols: int, dst: int = 0, device=None):
    """Collect per-recording posterior matrices ((T_i, n_spk_i <= n_cols) float32) on ``dst``.

    One fixed-size metadata all_gather + ONE gather of a padded flat buffer per rank (the
    path's only collective).  Returns {recording index: tensor} on ``dst``, None elsewhere."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return dict(local)
    world, rank = dist.get_world_size(), dist.get_rank()
    device = device or (next(iter(local.values())).device if local else torch.device("cpu"))
    ids = sorted(local)
    meta = torch.tensor([[i, local[i].shape[0], local[i].shape[1]] for i in ids], dtype=torch.int64).reshape(-1, 3)
    count = torch.tensor([len(ids), sum(int(local[i].numel()) for i in ids)], dtype=torch.int64, device=device)
    counts = [torch.zeros(2, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(counts, count)
    max_rec, max_elems = int(max(c[0] for c in counts)), int(max(c[1] for c in counts))
    meta_pad = torch.zeros((max(max_rec, 1), 3), dtype=torch.int64, device=device)
    meta_pad[: len(ids)] = meta.to(device)
    flat = torch.zeros((max(max_elems, 1),), dtype=torch.float32, device=device)
    pos = 0
    for i in ids:
        n = local[i].numel()
        flat[pos:pos + n] = local[i].reshape(-1).to(device=device, dtype=torch.float32)
        pos += n
    metas = [torch.zeros_like(meta_pad) for _ in range(world)] if rank == dst else None
    flats = [torch.zeros_like(flat) for _ in range(world)] if rank == dst else None
    dist.gather(meta_pad, metas, dst=dst)
    dist.gather(flat, flats, dst=dst)
    if rank != dst:
        return None
    out: Dict[int, torch.Tensor] = {}
    for r in range(world):
        pos = 0
        for k in range(int(counts[r][0])):
            i, t, s = (int(v) for v in metas[r][k])
            assert s <= n_cols
            out[i] = flats[r][pos:pos + t * s].reshape(t, s).clone()
            pos += t * s
    return out


# ------------------------------------------------------------ chunk-level sharding of ONE recording ----
# When there are fewer recordings than GPUs (a few long files), the chunks of one recording are spread over the
# ranks.  Chunks are independent in the model (infer_eda.py:102-113), but the front end normalises with the
# recording-wide mean (feature.py:72-73): every rank computes log-mel for its own frames (+ a halo for the +-7-frame
# splice), adds its own frames' column sums, and ONE all-reduce of 23 sums + 1 count gives every rank the mean.
HALO_FRAMES = 10          # frames before an internal slice start: >= 7 (splice context) + 2 (frames whose window would
                          # reach before the slice), and a multiple of every subsampling factor in use (10)


def stft_frames(n_samples: int, frame_shift: int) -> int:
    """feature.py:186-193 (``_count_frames``)."""
    n = 1 + n_samples // frame_shift
    return n - 1 if n_samples % frame_shift == 0 else n


def plan_chunk_shards(n_samples: int, frame_shift: int, n_fft: int, subsampling: int, chunk_size: int, world: int,
                      context: int = 7) -> List[Dict[str, int]]:
    """Contiguous chunk ranges of one recording for ``world`` ranks (as equal as possible; trailing ranks may get
    nothing).  Per rank: ``chunks`` [c0, c1), model ``rows`` [R0, R1), own STFT ``frames`` [F0, F1) (they partition the
    recording: the mean counts each frame once), the audio ``samples`` [a0, a1) to upload, ``lead`` = halo frames in
    front of F0 inside the slice, ``lead_rows`` = halo rows pass 2 produces in front of the first own row, and
    ``sum_range`` = the own frames as indices within the slice."""
    assert HALO_FRAMES % subsampling == 0 and HALO_FRAMES >= context + 2
    half = n_fft // 2
    n_frames = stft_frames(n_samples, frame_shift)
    rows_total = (n_frames + subsampling - 1) // subsampling
    n_chunks = (rows_total + chunk_size - 1) // chunk_size
    base, extra = divmod(n_chunks, world)
    out, c0 = [], 0
    for r in range(world):
        c1 = c0 + base + (1 if r < extra else 0)
        R0, R1 = min(c0 * chunk_size, rows_total), min(c1 * chunk_size, rows_total)
        F0 = min(R0 * subsampling, n_frames)
        F1 = n_frames if R1 == rows_total else R1 * subsampling
        if c1 == c0:                                    # nothing for this rank (more ranks than chunks)
            F0 = F1 = n_frames
        lead = HALO_FRAMES if (F0 > 0 and c1 > c0) else 0
        a0 = (F0 - lead) * frame_shift if c1 > c0 else 0
        last_needed = min(n_frames, F1 + context + 1) - 1                 # last frame whose log-mel this rank reads
        a1 = min(n_samples, (last_needed + 2) * frame_shift + half) if c1 > c0 else a0
        out.append(dict(rank=r, chunks=(c0, c1), rows=(R0, R1), frames=(F0, F1), samples=(a0, max(a0, a1)), lead=lead,
                        lead_rows=lead // subsampling, sum_range=(lead, lead + F1 - F0), n_frames=n_frames,
                        rows_total=rows_total, chunk_lens=[min(chunk_size, rows_total - c * chunk_size) for c in range(c0, c1)],
                        all_chunk_lens=[min(chunk_size, rows_total - c * chunk_size) for c in range(n_chunks)]))
        c0 = c1
    return out


def allreduce_sum_(t: torch.Tensor, group=None) -> torch.Tensor:
    """The chunk-level path's one data collective: 23 column sums + 1 frame count per recording."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, group=group)
    return t


def shard_features(fe, audio, shard: Dict[str, int], allreduce=allreduce_sum_) -> torch.Tensor:
    """Features (rows R0..R1 of the recording, feat_dim) of one rank's slice, identical to those rows of
    ``fe.features([whole recording])``.  ``audio``: the whole recording (any 1-D array/tensor; only the slice is
    uploaded).  ``allreduce``: sums a float64 CUDA vector over the ranks (tests emulate several ranks with it)."""
    a0, a1 = shard["samples"]
    R0, R1 = shard["rows"]
    if R1 <= R0:
        allreduce(torch.zeros(fe.n_mels + 1, dtype=torch.float64, device=fe.device))       # take part in the collective
        return torch.empty((0, fe.feat_dim), dtype=torch.float32, device=fe.device)
    piece = torch.as_tensor(audio[a0:a1])
    lm, colsum, (offs, frames, rows) = fe.logmel([piece], sum_ranges=[shard["sum_range"]])
    part = torch.cat([colsum[0], torch.tensor([float(shard["frames"][1] - shard["frames"][0])], dtype=torch.float64,
                                              device=fe.device)])
    tot = allreduce(part)
    # pass 2 divides the sums it is given by the SLICE's frame count: hand it sums rescaled to the recording's mean
    scaled = (tot[:-1] * (float(frames[0]) / tot[-1])).reshape(1, -1).contiguous()
    feats = fe.splice(lm, scaled, offs, rows)
    k = shard["lead_rows"]
    return feats[k:k + (R1 - R0)]


def diarize_sharded_recording(model, fe, audio, rank: int, world: int, chunk_size: int = 2000, max_n_speakers: int = 15,
                              activity_threshold: float = 0.5, allreduce=allreduce_sum_):
    """One long recording over ``world`` ranks: this rank's chunks through the front end (with the mean all-reduce)
    and the model.  The frame shuffles are drawn by EVERY rank for ALL chunks in the reference's order
    (infer_eda.py:102-113) so that each rank uses exactly the draws the single-process driver would.
    Returns (posteriors (own rows, max_n_speakers-1), probs (own chunks, max_n_speakers), labels uint8, shard)."""
    plan = plan_chunk_shards(len(audio), fe.frame_shift, 1 << (fe.frame_size - 1).bit_length(), fe.subsampling,
                             chunk_size, world, fe.context_size)
    sh = plan[rank]
    feats = shard_features(fe, audio, sh, allreduce)
    perms_all = [torch.randperm(n) for n in sh["all_chunk_lens"]]
    c0, c1 = sh["chunks"]
    n_att = max_n_speakers - 1
    rows = sh["rows"][1] - sh["rows"][0]
    post = torch.empty((rows, n_att), dtype=torch.float32, device=fe.device)
    labels = torch.empty((rows, n_att), dtype=torch.uint8, device=fe.device)
    probs = torch.empty((0, max_n_speakers), dtype=torch.float32, device=fe.device)
    if c1 > c0:
        lens = sh["chunk_lens"]
        offs = [sum(lens[:i]) for i in range(len(lens))]
        _, probs, _ = model.infer_device(feats, lens, max_n_speakers, activity_threshold, perms=perms_all[c0:c1],
                                         independent=True, offsets=offs, packed_out=(post, labels))
    return post, probs, labels, sh


def allreduce_mean_(flat: torch.Tensor, group=None) -> torch.Tensor:
    """Data-parallel training (train.py's DDP wrapper averages gradients over ranks): ONE all-reduce of
    the whole flat gradient buffer, in place, then the mean.  No-op without an initialised group."""
    if not (dist.is_available() and dist.is_initialized()):
        return flat
    world = dist.get_world_size(group)
    if world > 1:
        dist.all_reduce(flat, group=group)
        flat.mul_(1.0 / world)
    return flat


def shard_batch(n_items: int, rank: int, world_size: int) -> List[int]:
    """Indices of a mini-batch's sequences this rank trains on (strided, like DistributedSampler)."""
    return list(range(rank, n_items, world_size))

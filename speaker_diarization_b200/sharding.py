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

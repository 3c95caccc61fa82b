"""Host-side multi-GPU logic on CPU: recording sharding, chunk indices, and the path's one
collective (result gather on rank 0) over a world_size-2 gloo group."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from speaker_diarization_b200.sharding import (allreduce_mean_, gather_posteriors, gen_chunk_indices, shard_batch,
                                               shard_recordings)


def test_chunk_indices_match_reference_generator():
    # infer_eda.py:21-27
    assert gen_chunk_indices(4800, 2000) == [(0, 2000), (2000, 4000), (4000, 4800)]
    assert gen_chunk_indices(2000, 2000) == [(0, 2000)]
    assert gen_chunk_indices(5, 0) == [(0, 5)]
    assert gen_chunk_indices(0, 2000) == []


def test_shard_recordings_balanced_and_complete():
    lens = [18000, 500, 4800, 4800, 12000, 700, 9000, 100, 18000]
    for world in (1, 2, 4, 8):
        shards = shard_recordings(lens, world)
        assert sorted(i for s in shards for i in s) == list(range(len(lens)))
        loads = [sum(lens[i] for i in s) for s in shards]
        assert max(loads) - min(loads) <= max(lens)          # LPT bound
    assert shard_recordings(lens, 2) == shard_recordings(lens, 2)   # deterministic


def _fake_post(i, t, s):
    g = torch.Generator().manual_seed(100 + i)
    return torch.rand((t, s), generator=g)


def _worker(rank, world, port, lens, nspk, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = shard_recordings(lens, world)[rank]
        local = {i: _fake_post(i, lens[i], nspk[i]) for i in mine}
        got = gather_posteriors(local, n_cols=14, dst=0)
        if rank == 0:
            ok = sorted(got) == list(range(len(lens))) and all(
                torch.equal(got[i], _fake_post(i, lens[i], nspk[i])) for i in got)
            q.put(bool(ok))
        else:
            q.put(got is None)
    finally:
        dist.destroy_process_group()


def test_gather_posteriors_world2_gloo():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    lens = [50, 7, 33, 33, 12, 1]
    nspk = [2, 1, 4, 14, 3, 2]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, lens, nspk, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [True, True]


def _grad_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(7)
        per_seq = torch.randn((6, 1000), generator=g)                 # one "gradient" per sequence of the batch
        mine = shard_batch(6, rank, world)
        flat = per_seq[mine].mean(0)                                  # this rank's mini-batch mean
        allreduce_mean_(flat)
        q.put(bool(torch.allclose(flat, per_seq.mean(0), atol=1e-6)) and mine == list(range(rank, 6, world)))
    finally:
        dist.destroy_process_group()


def test_training_gradient_allreduce_world2_gloo():
    """The training step's one collective (NoamAdam.step): mean of the flat gradient over ranks."""
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_grad_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [True, True]
    assert torch.equal(allreduce_mean_(torch.ones(3)), torch.ones(3))     # no process group: untouched

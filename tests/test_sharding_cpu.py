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


# ------------------------------------------------------ chunk-level sharding of one recording ----------
def test_chunk_shard_plan_partitions_the_recording():
    from speaker_diarization_b200.sharding import HALO_FRAMES, plan_chunk_shards, stft_frames
    for n, hop, n_fft, world in [(8000 * 1800, 80, 256, 4), (8000 * 430 + 2400, 80, 256, 2), (16000 * 480 + 911, 160, 512, 8),
                                 (8000 * 20, 80, 256, 3), (79, 80, 256, 2)]:
        plan = plan_chunk_shards(n, hop, n_fft, 10, 2000, world)
        nf = stft_frames(n, hop)
        assert [p["frames"][0] for p in plan][0] == 0 and plan[-1]["frames"][1] == nf
        assert all(a["frames"][1] == b["frames"][0] and a["rows"][1] == b["rows"][0] and a["chunks"][1] == b["chunks"][0]
                   for a, b in zip(plan, plan[1:]))                       # frames / rows / chunks: each exactly once
        assert sum(sum(p["chunk_lens"]) for p in plan) == plan[0]["rows_total"] == (nf + 9) // 10
        for p in plan:
            a0, a1 = p["samples"]
            assert 0 <= a0 <= a1 <= n and p["lead"] in (0, HALO_FRAMES)
            assert (p["lead"] == 0) == (p["frames"][0] == 0 or p["rows"][1] == p["rows"][0])
            if p["rows"][1] > p["rows"][0]:
                # every frame the rank reads (own frames +- 7) has all its samples inside the slice or beyond the TRUE ends
                lo, hi = max(0, p["frames"][0] - 7), min(nf, p["frames"][1] + 7)
                assert max(0, lo * hop - n_fft // 2) >= a0 and min(n, (hi - 1) * hop + n_fft // 2) <= a1


def _slice_worker(rank, world, port, n, q):
    """Each rank runs the chunk-level front end on ITS slice with the oracle's numpy arithmetic standing in for the
    kernels and the real gloo all-reduce for the mean; the rows must equal those of the whole-recording front end."""
    import numpy as np
    from oracle import feature_oracle as fo
    from speaker_diarization_b200.sharding import allreduce_sum_, plan_chunk_shards
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        x, _ = fo.synth_audio(n, 8000, 2, 77, dtype=np.float64)
        whole = fo.frontend(x, 8000, 200, 80, 7, 10)
        sh = plan_chunk_shards(n, 80, 256, 10, 500, world)[rank]
        a0, a1 = sh["samples"]
        lm = fo.logmel23(fo.stft(x[a0:a1], 200, 80), 8000)                       # slice as if it were a recording
        lo, hi = sh["sum_range"]
        part = torch.from_numpy(np.concatenate([lm[lo:hi].sum(0), [float(hi - lo)]]))
        tot = allreduce_sum_(part)                                                # the path's one data collective
        mean = (tot[:-1] / tot[-1]).numpy()
        feats = np.ascontiguousarray(fo.splice((lm - mean).astype(np.float32), 7)[::10])
        k = sh["lead_rows"]
        R0, R1 = sh["rows"]
        mine = feats[k:k + R1 - R0]
        q.put((rank, float(tot[-1]), bool(mine.shape == whole[R0:R1].shape), float(np.abs(mine - whole[R0:R1]).max()) if R1 > R0 else 0.0))
    finally:
        dist.destroy_process_group()


def test_chunk_level_front_end_world2_gloo():
    from speaker_diarization_b200.sharding import stft_frames
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    n = 8000 * 130 + 333                                   # 13005 frames -> 1301 rows -> 3 chunks of 500 (2 + 1)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_slice_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, count, shape_ok, err in res:
        assert count == stft_frames(n, 80) and shape_ok and err < 2e-6, (rank, count, shape_ok, err)


def test_pipeline_chunk_plan_is_the_drivers_chunking():
    """pipeline.chunk_plan over a batch of recordings == infer_eda.py:21-27 applied per recording, in the driver's order."""
    from speaker_diarization_b200.pipeline import chunk_plan
    rows = [4800, 2000, 1, 18000, 4303]
    plan = chunk_plan(rows, 2000)
    want, base = [], 0
    for r, n in enumerate(rows):
        want += [(r, base + s, e - s) for s, e in gen_chunk_indices(n, 2000)]
        base += n
    assert plan == want
    assert [c[2] for c in plan if c[0] == 0] == [2000, 2000, 800] and sum(c[2] for c in plan) == sum(rows)
    assert chunk_plan([500, 300], None) == [(0, 0, 500), (1, 500, 300)]          # no chunking: one call per recording

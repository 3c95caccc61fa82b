#!/usr/bin/env python
"""configs[2] and configs[3] of BASELINE.json as workloads (builder-run lines under profiles/; bench.py --workload c3|c4).

  c3  64 far-field-shaped 4-speaker sessions of ~8 min @ 16 kHz (frame 400 / hop 160 -> ~4800 model frames each ->
      chunks 2000 / 2000 / tail), sharded over the ranks by RECORDING (longest-processing-time greedy)
  c4  16 long-form recordings of 30 min @ 8 kHz (18 000 model frames -> 9 chunks of 2000), speaker count decided by the
      attractor threshold; with fewer recordings than ranks (--recordings 4 at 8 GPUs) the CHUNKS of each recording are
      spread over the ranks and the recording-wide mel mean comes from one 24-value all-reduce

A step = the whole job (all recordings) once; total work is fixed, so scaling is "strong".  Every rank pushes its
recordings through InferencePipeline (chunked, independent chunks = the reference driver's semantics, infer_eda.py:92-124)
in batches of --per-batch recordings.  value = total model frames / (max over ranks of the device time of a step);
`imbalance` = max / mean of the ranks' device times - the scaling risk of this path is load balance, not bandwidth.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import bench as B0                                                                          # noqa: E402

WORKLOADS = {
    "c3": dict(idx=2, rate=16000, frame=400, shift=160, n_rec=64, seconds=480.0, spread=0.12, per_batch=16,
               what="AliMeeting-shaped 4-speaker 8-min sessions @ 16 kHz"),
    "c4": dict(idx=3, rate=8000, frame=200, shift=80, n_rec=16, seconds=1800.0, spread=0.0, per_batch=8,
               what="long-form 30-min recordings @ 8 kHz, 1-6 speakers by attractor threshold"),
}
LAYERS, HEADS, UNITS, FF, IN, MAXSPK, CTX, SUB, CHUNK = 4, 4, 256, 2048, 345, 15, 7, 10, 2000


def rec_seconds(w, i):
    """Deterministic spread of session lengths around the nominal one (real sessions are not all equal)."""
    return w["seconds"] * (1.0 + w["spread"] * np.sin(1.7 * i + 0.3))


def synth_pcm(n: int, rate: int, seed: int) -> torch.Tensor:
    """int16 synthetic mixture of length n: a 30 s block of gated coloured noise, tiled under a slow gain (the content
    does not influence the timing; generating 64 x 7.7 M samples from scratch would take minutes of host time)."""
    g = torch.Generator().manual_seed(0xC3C4 + seed)
    blk = 30 * rate
    noise = torch.randn((1, 1, blk), generator=g)
    k = torch.hann_window(33).view(1, 1, -1)
    col = torch.nn.functional.conv1d(noise, k / k.sum(), padding=16).view(-1)
    gate = (torch.rand(blk // 4000 + 1, generator=g) < 0.6).float().repeat_interleave(4000)[:blk]
    x = (0.6 * gate * col + 1e-3 * torch.randn(blk, generator=g))
    reps = (n + blk - 1) // blk
    gain = 0.4 + 0.3 * torch.rand(reps, generator=g)
    y = (x.view(1, -1) * gain.view(-1, 1)).reshape(-1)[:n]
    return B0.pcm16(y.clamp_(-1, 1))


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--precision", default="bf16x3", choices=["fp32", "bf16x3", "bf16"])
    ap.add_argument("--recordings", type=int, default=0, help="override the number of recordings")
    ap.add_argument("--per-batch", type=int, default=0, help="recordings per pipeline batch")
    ap.add_argument("--skip-cpu", action="store_true")
    args = ap.parse_args(argv)
    w = dict(WORKLOADS[args.workload])
    if args.recordings:
        w["n_rec"] = args.recordings
    if args.per_batch:
        w["per_batch"] = args.per_batch

    import torch.distributed as dist
    from speaker_diarization_b200 import _lib as L, feature, sharding
    from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
    from speaker_diarization_b200.pipeline import InferencePipeline
    from speaker_diarization_b200.synth import EdaConfig, synth_state_dict

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = B0.pin_to_gpu_numa(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    model = TransformerEdaModel(4, IN, HEADS, UNITS, LAYERS, dim_feedforward=FF, precision=args.precision)
    model.load_state_dict(synth_state_dict(EdaConfig(n_layers=LAYERS), 777))
    model = model.to(dev).eval()
    fe = feature.FrontEnd(w["rate"], w["frame"], w["shift"], CTX, SUB, device=dev)

    n_samples = [int(round(rec_seconds(w, i) * w["rate"])) for i in range(w["n_rec"])]
    rows = [(sharding.stft_frames(n, w["shift"]) + SUB - 1) // SUB for n in n_samples]
    total_frames = int(sum(rows))
    chunk_level = w["n_rec"] < world                      # fewer recordings than ranks: split each recording's chunks
    mine = list(range(w["n_rec"])) if chunk_level else sharding.shard_recordings(rows, world)[rank]
    audio = {i: synth_pcm(n_samples[i], w["rate"], i).pin_memory() for i in mine}
    my_frames = 0

    if not chunk_level:
        pipe = InferencePipeline(model, fe, max_n_speakers=MAXSPK, depth=3, compute_streams=3, chunk_size=CHUNK)
        batches = [mine[k:k + w["per_batch"]] for k in range(0, len(mine), w["per_batch"])]
        host_batches = [(torch.cat([audio[i] for i in b]).pin_memory(), [n_samples[i] for i in b]) for b in batches]
        dev_batches = [(a.to(dev), c) for a, c in host_batches]
        my_frames = sum(rows[i] for i in mine)

        def run_steps(bs, n):
            last = None
            for last in pipe.run(bs[k % len(bs)] for k in range(n * len(bs))):
                pass
            return last
        first, last_stream = pipe.s_compute, pipe.s_out
    else:
        s_main = torch.cuda.Stream()
        group_sum = (lambda t: sharding.allreduce_sum_(t))
        dev_audio = {i: audio[i].to(dev) for i in mine}
        host_out = {}

        def run_steps(src, n):
            res = None
            with torch.cuda.stream(s_main):
                for _ in range(n):
                    for i in mine:
                        post, probs, labels, sh = sharding.diarize_sharded_recording(model, fe, src[i], rank, world, CHUNK,
                                                                                     MAXSPK, 0.5, group_sum)
                        if src is audio:                                        # e2e: results land on the host
                            host_out[i] = (post.to("cpu", non_blocking=True), probs.to("cpu", non_blocking=True))
                        res = post
            return res
        host_batches, dev_batches = audio, dev_audio
        plan0 = [sharding.plan_chunk_shards(n_samples[i], w["shift"], 1 << (w["frame"] - 1).bit_length(), SUB, CHUNK, world)[rank]
                 for i in mine]
        my_frames = sum(p["rows"][1] - p["rows"][0] for p in plan0)
        first = last_stream = s_main

    torch.manual_seed(1234)
    if bool(mine) or chunk_level:
        run_steps(dev_batches, max(1, args.warmup))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = B0.ClockSampler(local)
    sampler.start()
    n0 = L.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(first)
    if bool(mine) or chunk_level:
        run_steps(dev_batches, args.steps)
    ev1.record(last_stream)
    torch.cuda.synchronize()
    my_ms = ev0.elapsed_time(ev1) / args.steps
    launches = L.launch_count() - n0
    all_ms = torch.tensor([my_ms], device=dev)
    if world > 1:
        dist.barrier()
        gathered = [torch.zeros_like(all_ms) for _ in range(world)]
        dist.all_gather(gathered, all_ms)
        per_rank = [float(g.item()) for g in gathered]
    else:
        per_rank = [my_ms]
    ms_step = max(per_rank)

    # ---- end to end: pinned host audio in, host results out ----
    passes = []
    for _ in range(3):
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        if bool(mine) or chunk_level:
            run_steps(host_batches, args.steps)
        torch.cuda.synchronize()
        dt = torch.tensor([(time.perf_counter() - t0) / args.steps], device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        passes.append(dt.item())
    passes.sort()
    clocks = sampler.stop()

    out = None
    prof_all = None
    if chunk_level:
        # the roofline leg of the chunk-level path contains the all-reduce: EVERY rank runs it (rank 0 alone would wait
        # for its peers forever), only rank 0 keeps the timings
        if rank == 0:
            L.PROFILE = {}
        run_steps(dev_batches, 1)
        torch.cuda.synchronize()
        if rank == 0:
            prof_all = L.profile_ms()
            L.PROFILE = None
    if rank == 0:
        pk = B0.peaks()
        # roofline leg: ONE batch (or one recording's shard) on one stream, every C-ABI call bracketed by CUDA events
        if not chunk_level:
            serial = InferencePipeline(model, fe, max_n_speakers=MAXSPK, depth=1, compute_streams=1, chunk_size=CHUNK)
            serial.submit(*dev_batches[0]); serial.collect()
            L.PROFILE = {}
            with torch.cuda.stream(serial.s_compute):
                pass
            serial.submit(*dev_batches[0]); serial.collect()
            my_prof_frames = sum(rows[i] for i in batches[0])
            prof = L.profile_ms()
            L.PROFILE = None
        else:
            prof, my_prof_frames = prof_all, my_frames
        tot = sum(m for v in prof.values() for m, _ in v)
        shares = {k: round(sum(m for m, _ in v) / tot, 4) for k, v in prof.items()}
        gemm = [(m, 2.0 * a[16] * a[17] * min(a[18], IN if a[18] == 352 else a[18])) for m, a in prof.get("eda_gemm_split", [])]
        gemm += [(m, 2.0 * a[12] * a[13] * 3 * a[15] * a[15]) for m, a in prof.get("eda_gemm_qkv", [])]
        g_ms, g_fl = sum(m for m, _ in gemm), sum(f for _, f in gemm)
        ach = g_fl / (g_ms * 1e-3) / 1e12
        issued = 3.0 if args.precision == "bf16x3" else 1.0
        fe_ms = sum(m for k in ("eda_frontend_transform", "eda_frontend_logmel_ranges", "eda_frontend_splice_subsample_w")
                    for m, _ in prof.get(k, []))
        fe_bytes = my_prof_frames * (4 * w["shift"] * SUB + 4 * IN)
        roof = dict(bound="tensor", kernel="gemm_split_kernel", achieved=round(ach, 2), peak=pk["tf_sust"], unit="TFLOP/s",
                    frac=round(ach / pk["tf_sust"], 4), traffic=None, issued_mma_factor=issued,
                    tensor_pipe_frac_issued=round(issued * ach / pk["tf_sust"], 4), launches_per_step=len(gemm),
                    avg_launch_ms=round(g_ms / max(1, len(gemm)), 4), share_of_step=shares, peak_source=pk["src"] + " (bf16 sustained)",
                    frontend=dict(bound="hbm", achieved=round(fe_bytes / (fe_ms * 1e-3) / 1e9, 1), peak=pk["hbm"], unit="GB/s",
                                  frac=round(fe_bytes / (fe_ms * 1e-3) / 1e9 / pk["hbm"], 4)))
        cpu = None
        if not args.skip_cpu and world == 1:
            # bounded sample: ONE recording through the reference driver's loop (features, then one chunk per model call)
            if B0.ORIG_AFFINITY:
                os.sched_setaffinity(0, B0.ORIG_AFFINITY)
            cores = os.cpu_count() or 1
            save = (B0.RATE, B0.FRAME, B0.SHIFT)
            B0.RATE, B0.FRAME, B0.SHIFT = w["rate"], w["frame"], w["shift"]
            try:
                arm = B0.CpuArm(cores)
                a = (audio[mine[0]].double() / 32768.0).numpy()
                t0 = time.perf_counter()
                with torch.no_grad():
                    Y = arm.features(a)
                    for s0 in range(0, len(Y), CHUNK):
                        arm.model.infer([Y[s0:s0 + CHUNK]], attractor_threshold=0.5, max_n_speakers=MAXSPK)
                dt = time.perf_counter() - t0
            finally:
                B0.RATE, B0.FRAME, B0.SHIFT = save
            cpu = {"value": round(rows[mine[0]] / dt, 1), "unit": B0.UNIT, "cores": cores, "kind": arm.kind,
                   "sample": f"1 of the {w['n_rec']} recordings ({rows[mine[0]]} frames, {dt:.1f} s): features + one chunk per model "
                             f"call as in infer_eda.py:92-113, all host threads; {arm.what}"}
        out = {
            "metric": B0.METRIC, "value": round(total_frames / (ms_step * 1e-3), 1), "unit": B0.UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 3), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None,
            "dtype": {"fp32": "f32", "bf16x3": "bf16x3(f32 accum)", "bf16": "bf16"}[args.precision], "data": "synthetic",
            "config": {"workload": f"configs[{w['idx']}]: {w['n_rec']} {w['what']} ({total_frames} model frames per step), "
                                   f"{LAYERS}-layer {UNITS}-d encoder, chunks of {CHUNK} frames, whole job per step",
                       "frames_per_step": total_frames, "precision": args.precision,
                       "parallelism": (f"dp{world}: chunks of every recording spread over the ranks, one 24-value all-reduce per "
                                       "recording for the mel mean" if chunk_level else
                                       f"dp{world}: recordings sharded longest-first ({w['per_batch']} per pipeline batch)"),
                       "l2": "audio of a step (>= 460 MB) and the FFN intermediates exceed the 126 MB L2",
                       "roofline_leg": "one pipeline batch on one stream (the timed run overlaps batches on 3 streams)",
                       "host": f"rank thread pinned to {numa}"},
            "rank_ms_per_step": [round(v, 3) for v in per_rank],
            "imbalance": round(max(per_rank) / (sum(per_rank) / len(per_rank)), 4),
            "e2e": {"value": round(total_frames / passes[len(passes) // 2], 1), "unit": B0.UNIT,
                    "h2d_bytes_per_step": int(sum(2 * n_samples[i] for i in mine)), "d2h_bytes_per_step": int(my_frames * (MAXSPK - 1) * 5),
                    "passes": 3, "min": round(total_frames / passes[-1], 1), "max": round(total_frames / passes[0], 1),
                    "note": "bytes are rank 0's"},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "cpu_baseline": cpu,
        }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if out is not None:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()

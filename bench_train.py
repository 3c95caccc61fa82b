#!/usr/bin/env python
"""EEND-EDA TRAINING step on B200 (BASELINE.json:configs[4]): frames/s of
forward + PIT loss + backward + Adam/Noam update, one process per GPU (data parallel, one NCCL
all-reduce of the flat gradient per step).

    python bench_train.py [--steps K] [--warmup W] [--precision bf16x3|bf16] [--impl ours|reference]
    (also reachable as  python bench.py --workload train ...)

Workload C5: batch 32 x 2000-frame chunks of 345-d features (randn), 3 speakers (Bernoulli-run
labels), 4-layer 256-d encoder, dropout 0.1 (the reference's training yaml).  ``value``: features and labels resident in HBM;
``e2e``: host tensors through ``model(src, ts)`` / ``loss.backward()`` / ``opt.step()`` with the
H2D of the step's features + labels and the D2H of its loss / DER counters inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

B, T, S, LAYERS, HEADS, UNITS, FF, IN = 32, 2000, 3, 4, 4, 256, 2048, 345
METRIC, UNIT = "eend_eda_train_frames_per_sec", "frames/s"


def make_batch(seed: int):
    from speaker_diarization_b200.synth import synth_features, synth_labels      # seeded inputs only (not the oracle path)
    return synth_features([T] * B, IN, seed), synth_labels([T] * B, S, seed)


def train_flops_per_frame() -> float:
    """Algorithmic flops of one training step per model frame: 3 x the forward's GEMM + attention
    work (forward, input gradient, weight gradient) + the attention backward's extra recompute."""
    fwd = 2 * IN * UNITS + LAYERS * (2 * UNITS * 3 * UNITS + 2 * UNITS * UNITS + 4 * UNITS * FF) + 2 * UNITS * 4 * UNITS
    att_f = LAYERS * 4 * T * UNITS              # QK^T and PV
    return 3 * fwd + att_f * (1 + 2.5) + 2 * 2 * UNITS * 4 * UNITS


def cpu_reference(n_seq: int, steps: int, warmup: int, threads: int):
    from oracle.eend_eda_oracle import adam_step, forward_backward, noam_lr
    from oracle.weights import EdaConfig, synth_features, synth_labels, synth_state_dict
    torch.set_num_threads(threads)
    params = {k: v.clone() for k, v in synth_state_dict(EdaConfig(n_layers=LAYERS, n_speakers=S), 777).items()}
    src, ts = synth_features([T] * n_seq, IN, 0), synth_labels([T] * n_seq, S, 0)
    state = {}
    def step(i):
        loss, grads, _ = forward_backward(params, src, ts)
        adam_step(params, grads, state, i + 1, noam_lr(i, UNITS, 25000))
    for i in range(warmup):
        step(i)
    t0 = time.perf_counter()
    for i in range(steps):
        step(warmup + i)
    dt = (time.perf_counter() - t0) / steps
    return n_seq * T / dt, dt


def run_ours(args):
    import torch.distributed as dist
    from speaker_diarization_b200.synth import EdaConfig, synth_state_dict
    from speaker_diarization_b200 import _lib as L
    from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
    from speaker_diarization_b200.eend_eda.training import NoamAdam
    from bench import ClockSampler, peaks

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    model = TransformerEdaModel(S, IN, HEADS, UNITS, LAYERS, dim_feedforward=FF, dropout=args.dropout, precision=args.precision)
    model.load_state_dict(synth_state_dict(EdaConfig(n_layers=LAYERS, n_speakers=S), 777))
    model = model.to(dev).train()
    opt = NoamAdam(model, lr=1.0, warmup_steps=25000)
    host = [make_batch(2 * rank + i) for i in range(2)]
    host = [([x.pin_memory() for x in s], [t.pin_memory() for t in l]) for s, l in host]
    resident = [([x.to(dev) for x in s], [t.to(dev) for t in l]) for s, l in host]

    def step(batch):
        src, ts = batch
        loss, stats = model(src, ts)              # stats: the step's D2H (DER counters)
        loss.backward()
        opt.step()
        opt.zero_grad()
        return loss

    torch.manual_seed(77 + rank)
    for i in range(max(3, args.warmup)):
        step(resident[i & 1])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    n0 = L.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(args.steps):
        step(resident[i & 1])
    ev1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    launches = L.launch_count() - n0
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_step = ms.item() / args.steps
    value = world * B * T / (ms_step * 1e-3)

    from speaker_diarization_b200.eend_eda.training import DeviceBatches
    feeder = DeviceBatches(None, dev)
    for batch in feeder.iterate(host[i & 1] for i in range(4)):      # warm-up: allocates the staging sets
        step(batch)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    # the step's inputs go pinned host -> device inside the timed region, staged one batch ahead on a copy stream
    # (DeviceBatches = the reference's DataLoader(pin_memory) + non_blocking .to(device)); the loss is read back every
    # step, one step late so that the host never stalls the GPU
    prev = None
    for batch in feeder.iterate(host[i & 1] for i in range(args.steps)):
        loss_t = step(batch)
        if prev is not None:
            last = prev.item()
        prev = loss_t.detach().clone()
    last = prev.item()
    torch.cuda.synchronize()
    e2e_s = torch.tensor([(time.perf_counter() - t0) / args.steps], device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    clocks = sampler.stop()
    out = None
    # one instrumented step (per-entry-point CUDA events on rank 0); every rank runs it: it contains the all-reduce
    if rank == 0:
        L.PROFILE = {}
    step(resident[0])
    if rank == 0:
        prof = L.profile_ms()
        L.PROFILE = None
        tot = sum(m for v in prof.values() for m, _ in v)
        shares = {k: round(sum(m for m, _ in v) / tot, 4) for k, v in prof.items()}
        times = {k: round(sum(m for m, _ in v), 4) for k, v in prof.items()}
        # GEMM launches (forward, input gradient, weight gradient): algorithmic 2*M*N*K each
        gemm = [(m, 2.0 * a[20] * a[21] * min(a[22], IN if a[22] == 352 else a[22])) for m, a in prof.get("eda_gemm_split_train", [])]
        gemm += [(m, 2.0 * a[16] * a[17] * a[18]) for m, a in prof.get("eda_gemm_split", [])]
        gemm += [(m, 2.0 * a[12] * a[13] * 3 * a[15] * a[15]) for m, a in prof.get("eda_gemm_qkv", [])]
        gemm += [(m, 2.0 * a[16] * a[17] * a[18]) for m, a in prof.get("eda_gemm_dgrad", [])]
        gemm += [(m, 2.0 * a[6] * a[7] * a[8]) for m, a in prof.get("eda_gemm_wgrad", [])]
        g_ms, g_fl = sum(m for m, _ in gemm), sum(f for _, f in gemm)
        pk = peaks()
        issued = 3.0 if args.precision == "bf16x3" else 1.0
        ach = g_fl / (g_ms * 1e-3) / 1e12
        roof = dict(bound="tensor", kernel="gemm_split_kernel (forward, input-gradient and weight-gradient launches)",
                    achieved=round(ach, 2), peak=pk["tf_sust"], unit="TFLOP/s", frac=round(ach / pk["tf_sust"], 4),
                    traffic=None, peak_source=pk["src"] + " (bf16 sustained)", issued_mma_factor=issued,
                    tensor_pipe_frac_issued=round(issued * ach / pk["tf_sust"], 4), launches_per_step=len(gemm),
                    avg_launch_ms=round(g_ms / len(gemm), 4), share_of_step=shares, ms_by_entry_point=times,
                    step_tflops_algorithmic=round(train_flops_per_frame() * B * T / (ms_step * 1e-3) / 1e12, 2))
        ab = prof.get("eda_mha_bwd", [])
        if ab:
            ab_ms = sum(m for m, _ in ab)
            ab_fl = sum(10.0 * a[11] * a[13] * a[12] * a[12] * 64 for _, a in ab)       # 5 products of 2*T*T*64 per (b, h)
            roof["attention_bwd"] = dict(bound="tensor", achieved=round(ab_fl / (ab_ms * 1e-3) / 1e12, 2), peak=pk["tf_sust"],
                                         unit="TFLOP/s", frac=round(ab_fl / (ab_ms * 1e-3) / 1e12 / pk["tf_sust"], 4),
                                         avg_launch_ms=round(ab_ms / len(ab), 4))
        cores = os.cpu_count() or 1
        cpu_fps, cpu_dt = (0.0, 0.0) if args.skip_cpu else cpu_reference(2, 1, 1, cores)
        out = {
            "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": {"bf16x3": "bf16x3(f32 accum)", "bf16": "bf16"}[args.precision], "data": "synthetic",
            "config": {"workload": f"configs[4]: EEND-EDA training step (forward + batched PIT loss + backward + Adam/Noam), "
                                   f"{LAYERS}-layer {UNITS}-d encoder, batch {B} x {T}-frame chunks, {S} speakers, dropout {args.dropout}, per GPU",
                       "global_batch": world * B, "frames_per_step": world * B * T, "precision": args.precision,
                       "parallelism": f"dp{world} (one NCCL all-reduce of the 20 MB flat gradient per step)",
                       "l2": "two alternating batches; > 5 GB of saved activations per step exceed the 126 MB L2"},
            "e2e": {"value": round(world * B * T / e2e_s.item(), 1), "unit": UNIT,
                    "h2d_bytes_per_step": B * T * (IN + S) * 4 + B * T * 4, "d2h_bytes_per_step": 4 + 9 * 8},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "last_loss": last,
            "cpu_baseline": {"value": round(cpu_fps, 1), "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"2 of the {B} chunks per step, 1 timed step after 1 warm-up ({cpu_dt:.2f} s/step), "
                                       "oracle/ torch-fp32 autograd + restated Adam/Noam"},
        }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def run_reference(args):
    if int(os.environ.get("RANK", 0)) != 0:
        return None
    cores = os.cpu_count() or 1
    fps, dt = cpu_reference(2, max(1, args.steps), 1, cores)
    return {"impl": "reference", "metric": METRIC, "value": round(fps, 1), "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt * 1e3, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"configs[4] bounded sample: 2 of {B} chunks x {T} frames per step, CPU"},
            "cpu_baseline": {"value": round(fps, 1), "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "oracle/ torch autograd restatement of model(y,t) + loss.backward() + Adam/Noam"},
            "e2e": {"value": round(fps, 1), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="bf16x3", choices=["bf16x3", "bf16"])
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--dropout", type=float, default=0.1, help="transformer_encoder_dropout (conf/eend_eda/*train.yaml: 0.1)")
    args = ap.parse_args(argv)
    out = run_reference(args) if args.impl == "reference" else run_ours(args)
    if out is not None:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()

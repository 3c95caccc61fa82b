#!/usr/bin/env python
"""EEND-EDA diarized frames/sec on B200 (BASELINE.json metric), one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step = one pass of the hot path over one batch of synthetic mixtures of the shape
BASELINE.json:configs[1] names: 64 recordings x 50 s @ 8 kHz -> feature front end
(STFT/log-mel/splice/subsample) -> 64 x 500-frame chunks -> 4-layer 256-d encoder -> EDA LSTM
-> 14 posteriors per frame + thresholded labels.  Weights are seeded random-init.
See DESIGN.md §Measurement for what each key means.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

B, SECONDS, RATE, FRAME, SHIFT, CTX, SUB = 64, 50, 8000, 200, 80, 7, 10
LAYERS, HEADS, UNITS, FF, IN, MAXSPK = 4, 4, 256, 2048, 345, 15
T = SECONDS * RATE // SHIFT // SUB            # 500 model frames per recording
METRIC, UNIT = "eend_eda_diarized_frames_per_sec", "frames/s"
E2E_PASSES = 5
ORIG_AFFINITY = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else None


# DRAM traffic of the dominant kernel from the committed ncu capture (bytes per launch, bf16x3, configs[1])
NCU_GEMM_DRAM_BYTES_PER_LAUNCH = 170.0e6


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=float(p["hbm_gbs"]), tf_burst=float(p["bf16_tflops"]),
                    tf_sust=float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), src="measured")
    except Exception:
        return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


def pcm16(x: torch.Tensor) -> torch.Tensor:
    """float [-1, 1) -> 16-bit PCM (what a wav file stores)."""
    return (x * 32768.0).round().clamp_(-32768, 32767).to(torch.int16).contiguous()


def synth_audio_batch(n_rec: int, seed: int) -> torch.Tensor:
    """(n_rec, SECONDS*RATE) float32 synthetic 2-speaker mixtures: gated, low-passed noise.
    Cheap vectorised stand-in for oracle.feature_oracle.synth_audio (same spirit: talk spurts
    of coloured noise), generated with torch on the CPU."""
    g = torch.Generator().manual_seed(0xBE7C4 + seed)
    n = SECONDS * RATE
    mix = torch.zeros((n_rec, n))
    for _ in range(2):
        noise = torch.randn((n_rec, n), generator=g)
        k = torch.hann_window(33).view(1, 1, -1)
        col = torch.nn.functional.conv1d(noise.unsqueeze(1), k / k.sum(), padding=16).squeeze(1)
        spurts = (torch.rand((n_rec, n // 4000 + 1), generator=g) < 0.55).float()
        gate = spurts.repeat_interleave(4000, dim=1)[:, :n]
        mix += (0.05 + 0.45 * torch.rand((n_rec, 1), generator=g)) * gate * col * 4.0
    mix += 1e-3 * torch.randn((n_rec, n), generator=g)
    return mix.clamp_(-1, 1).contiguous()


class ClockSampler(threading.Thread):
    """nvidia-smi-equivalent clock / throttle-reason samples (NVML) during the timed region."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._halt = threading.Event()

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                     "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                     "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                     "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
            while not self._halt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.reasons.update(k for k, bit in names.items() if r & bit)
                time.sleep(0.02)
        except Exception as e:  # pragma: no cover
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return dict(sm_mhz=(s[len(s) // 2] if s else None), sm_max_mhz=self.max_mhz, reasons=sorted(self.reasons))


# ------------------------------------------------------------------ CPU reference arm ----
def pin_to_gpu_numa(index: int):
    """Run this rank's host thread (and first-touch its pinned buffers) on the cores NVML reports as local to
    the GPU; with several ranks per NUMA node each rank takes its own slice of them."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(index)
        words = nv.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cores = [64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1]
        cores = sorted(set(cores) & set(os.sched_getaffinity(0)))
        if not cores:
            return None
        n_local = max(1, sum(1 for i in range(nv.nvmlDeviceGetCount())
                             if list(nv.nvmlDeviceGetCpuAffinity(nv.nvmlDeviceGetHandleByIndex(i), len(words))) == list(words)))
        # the whole NUMA-local set, shared with the other ranks of that node: slicing it per rank (4 cores each on a
        # 32-core box with 8 ranks) starved the rank's helper threads and cost 25 % of the 8-GPU device-timed throughput
        del n_local
        os.sched_setaffinity(0, cores)
        return f"{len(cores)} cores local to GPU {index} (shared by the ranks of that NUMA node)"
    except Exception as e:  # pragma: no cover
        return f"unpinned ({type(e).__name__})"


class CpuArm:
    """The reference's CPU path for configs[1] on this box's host cores.

    kind "reference": the reference's OWN files (feature.py, eend_eda/models.py, encoder_decoder_attractor.py) placed
    unmodified in the git-ignored oracle/_ref by oracle/build_ref.py, feature.py running on oracle/librosa_stub
    (librosa is not installable); kind "port": oracle/'s restatement, when oracle/_ref is absent.
    ``step(audio)``: best-effort use of the reference = front end per recording, ONE model.infer over the batch, all
    host threads.  ``as_shipped(audio)``: the reference driver's own loop (infer_eda.py:92-113: one recording at a
    time, one chunk per model call) with torch.set_num_threads(1) as train.py:87-88 sets it."""

    def __init__(self, threads: int):
        from oracle import build_ref
        from speaker_diarization_b200.synth import EdaConfig, synth_state_dict
        self.threads = threads
        torch.set_num_threads(threads)
        sd = synth_state_dict(EdaConfig(n_layers=LAYERS), 777)
        if build_ref.available():
            self.kind = "reference"
            self.feature, Model = build_ref.import_reference()
            self.model = Model(n_speakers=2, in_size=IN, n_units=UNITS, n_heads=HEADS, n_layers=LAYERS, has_pos=False)
            self.model.load_state_dict(sd)
            self.model.eval()
            self.what = ("unmodified reference files in oracle/_ref (feature.py on the librosa stand-in, "
                         "TransformerEdaModel.infer), fp32 torch CPU")
        else:
            from oracle import feature_oracle as fo
            from oracle.eend_eda_oracle import OracleEda
            self.kind = "port"
            self.fo, self.model = fo, OracleEda(sd, n_heads=HEADS, library_lstm=True)
            self.what = "oracle/ restatement of feature.py + models.py infer (oracle/_ref absent), fp32 torch CPU"

    def features(self, a):
        if self.kind == "reference":
            F = self.feature
            Y = F.stft(a, FRAME, SHIFT)
            Y = F.transform(Y, transform_type="logmel23_mn", sample_rate=RATE)
            return torch.from_numpy(F.splice(Y, context_size=CTX)[::SUB].copy())
        return torch.from_numpy(self.fo.frontend(a, RATE, FRAME, SHIFT, CTX, SUB))

    def step(self, audio):
        with torch.no_grad():
            feats = [self.features(a) for a in audio]
            return self.model.infer(feats, attractor_threshold=0.5, max_n_speakers=MAXSPK)

    def as_shipped(self, audio):
        torch.set_num_threads(1)
        try:
            t0 = time.perf_counter()
            with torch.no_grad():
                for a in audio:
                    Y = self.features(a)
                    for s0 in range(0, len(Y), 2000):
                        self.model.infer([Y[s0:s0 + 2000]], attractor_threshold=0.5, max_n_speakers=MAXSPK)
            return len(audio) * T / (time.perf_counter() - t0)
        finally:
            torch.set_num_threads(self.threads)


def cpu_audio(n_rec: int):
    # soundfile hands the reference float64 = int16 PCM / 32768 (kaldi_data.py:82)
    return (pcm16(synth_audio_batch(n_rec, 0)).double() / 32768.0).numpy()


def cpu_reference(n_rec: int, steps: int, warmup: int, threads: int, as_shipped_recs: int = 2):
    """Returns (arm, frames/s, seconds per step, as-shipped frames/s)."""
    arm = CpuArm(threads)
    audio = cpu_audio(n_rec)
    torch.manual_seed(0)
    for _ in range(warmup):
        arm.step(audio)
    t0 = time.perf_counter()
    for _ in range(steps):
        arm.step(audio)
    dt = (time.perf_counter() - t0) / steps
    shipped = arm.as_shipped(audio[:as_shipped_recs]) if as_shipped_recs else None
    return arm, n_rec * T / dt, dt, shipped


# ------------------------------------------------------------------------- B200 arm ----
def flops_linear(args) -> float:
    # eda_linear(A, lda, W, ldw, bias, resid, ldr, g, b, C, ldc, M, N, K, epi, prec, stream)
    M, N, K = args[11], args[12], args[13]
    return 2.0 * M * N * min(K, IN if K == 352 else K)


def run_ours(args):
    import torch.distributed as dist
    from speaker_diarization_b200.synth import EdaConfig, synth_state_dict     # seeded weights (neutral module)
    from speaker_diarization_b200 import _lib as L, feature
    from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
    from speaker_diarization_b200.pipeline import InferencePipeline

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = pin_to_gpu_numa(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    precision = os.environ.get("EDA_B200_PRECISION", args.precision)
    model = TransformerEdaModel(2, IN, HEADS, UNITS, LAYERS, dim_feedforward=FF, precision=precision)
    model.load_state_dict(synth_state_dict(EdaConfig(n_layers=LAYERS), 777))
    model = model.to(dev).eval()
    fe = feature.FrontEnd(RATE, FRAME, SHIFT, CTX, SUB, device=dev)
    pipe = InferencePipeline(model, fe, max_n_speakers=MAXSPK, attractor_threshold=0.5, activity_threshold=0.5, depth=args.depth,
                             compute_streams=args.compute_streams)
    # host audio = 16-bit PCM, the sample format of the wav files the reference reads (kaldi_data.py:59-83)
    host_audio = [pcm16(synth_audio_batch(B, 2 * rank + i)).pin_memory() for i in range(2)]    # 2 x 51 MB
    dev_audio = [a.to(dev) for a in host_audio]
    ns = [SECONDS * RATE] * B
    collect_results = None
    if world > 1:
        # Recordings are independent (infer_eda.py:92-113): the sharded hot path has NO data-path collective, and every rank
        # lands its own recordings' results in its own pinned host memory (the reference writes one file per recording).
        # Result collection on rank 0 - the one NCCL gather north_star names - happens ONCE PER RUN: each step's thresholded
        # decisions are kept as 16-bit masks (2 bytes per frame, 64 KB per step; SURVEY 8e: "posteriors ... or the thresholded
        # bitmask") in a device ring, which is gathered after the last batch; it is timed on its own (`result_collection`),
        # outside the per-step figure.  Measured alternatives on this pool (1.93 ms/step on one GPU), device-timed ms/step at
        # 8 GPUs: a list gather of the fp32 posteriors per step 2.73, an all-gather per step 3.68, the gather of the run's
        # posteriors counted inside a 20-step window 4.18 - every collective call blocked the issuing host thread for
        # 1.3-7 ms, per-step collectives put the eight ranks in lock-step, and a gather issued after the pipeline has drained
        # shows a ~30 ms fixed latency of the send/recv path on these boxes whatever its size.
        ring = torch.empty((max(args.steps, 1), B * T), dtype=torch.int16, device=dev)
        ring_bytes = ring.view(torch.uint8)                     # NCCL (through torch) carries no 16-bit integers
        gather_list = [torch.empty_like(ring_bytes) for _ in range(world)] if rank == 0 else None
        ring_pos = [0]

        def keep(post, mask):
            ring[ring_pos[0] % ring.shape[0]].copy_(mask, non_blocking=True)
            ring_pos[0] += 1
        pipe.gather = keep

        def collect_results():
            """One NCCL gather of the run's decision masks to rank 0; returns its device time in ms."""
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            dist.barrier()
            e0.record()
            dist.gather(ring_bytes, gather_list, dst=0)
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1)

    def run_steps(audio_list, n):
        last = None
        for last in pipe.run((audio_list[i & 1], ns) for i in range(n)):
            pass
        return last

    torch.manual_seed(1234 + rank)
    # the pipeline keeps 3 batches in flight and hands out freshly allocated pinned result buffers: its allocators and the
    # three stream slots reach their steady state only after depth + a few batches, whatever W says (more untimed warm-up
    # than asked for is harmless; with 3 warm-up steps one-time costs of ~50 ms fell into the timed region)
    WARM = max(8, args.warmup)
    run_steps(dev_audio, WARM)
    torch.cuda.synchronize()
    if world > 1:
        collect_results()                      # connections of the send/recv path are made here, not in a timed call
        dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    n0 = L.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    ev0.record(pipe.s_compute)
    run_steps(dev_audio, args.steps)
    ev1.record(pipe.s_out)                 # the result stream trails the compute stream
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    launches = L.launch_count() - n0
    collect_ms = collect_results() if world > 1 else None
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_step = ms.item() / args.steps
    value = world * B * T / (ms_step * 1e-3)

    # ---- end to end through the public API with HOST buffers (pinned), result read back ----
    # K steps per pass, E2E_PASSES passes: the median pass is the value, min / max show the spread (one host hiccup
    # inside a single 40 ms window used to halve the number)
    run_steps(host_audio, 6)
    torch.cuda.synchronize()
    passes = []
    for _ in range(E2E_PASSES):
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        res = run_steps(host_audio, args.steps)       # every step: H2D of its audio, D2H of its posteriors / labels / probs
        torch.cuda.synchronize()
        dt = torch.tensor([(time.perf_counter() - t0) / args.steps], device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        passes.append(dt.item())
    assert res["posteriors"].shape == (B * T, MAXSPK - 1) and len(res["n_speakers"]) == B
    passes.sort()
    e2e_med, e2e_best, e2e_worst = passes[len(passes) // 2], passes[0], passes[-1]
    clocks = sampler.stop()
    h2d = B * SECONDS * RATE * 2 + B * T * 4 + 3 * (B + 1) * 8 + (2 * B + 1) * 8
    d2h = B * T * (MAXSPK - 1) * 5 + B * MAXSPK * 4

    def step(i, audio_list):
        with torch.cuda.stream(pipe.s_compute):
            feats, rows = fe.features_flat(audio_list[i & 1], ld=352, counts=ns)
            return model.infer_device(feats, rows, MAXSPK, 0.5)

    out = None
    if rank == 0:
        # ---- roofline leg: per-entry-point device time of one instrumented step ------------
        L.PROFILE = {}
        with torch.cuda.stream(pipe.s_compute):
            step(0, dev_audio)
        prof = L.profile_ms()
        L.PROFILE = None
        tot = sum(ms_ for v in prof.values() for ms_, _ in v)
        shares = {k: round(sum(m for m, _ in v) / tot, 4) for k, v in prof.items()}
        gemm = [(m, flops_linear(a)) for m, a in prof.get("eda_linear", [])]
        gemm += [(m, 2.0 * a[16] * a[17] * min(a[18], IN if a[18] == 352 else a[18])) for m, a in prof.get("eda_gemm_split", [])]
        gemm += [(m, 2.0 * a[12] * a[13] * 3 * a[15] * a[15]) for m, a in prof.get("eda_gemm_qkv", [])]
        lin_ms, lin_fl = sum(m for m, _ in gemm), sum(f for _, f in gemm)
        pk = peaks()
        issued = 3.0 if precision == "bf16x3" else 1.0
        achieved = lin_fl / (lin_ms * 1e-3) / 1e12            # ALGORITHMIC flops (2*M*N*K) per second
        roof = dict(bound="tensor", kernel="gemm_split_kernel (eda_linear / eda_gemm_split / eda_gemm_qkv)",
                    achieved=round(achieved, 2), peak=pk["tf_sust"],
                    unit="TFLOP/s", frac=round(achieved / pk["tf_sust"], 4),
                    traffic=NCU_GEMM_DRAM_BYTES_PER_LAUNCH if precision == "bf16x3" else None,
                    traffic_source="dram__bytes_read.sum + dram__bytes_write.sum averaged over the step's 18 gemm_split_kernel "
                                   "launches, ncu --set full (profiles/r02_ncu_step_v1.md)",
                    peak_source=pk["src"] + " (bf16 sustained)", issued_mma_factor=issued,
                    issued_tflops=round(issued * achieved, 2), tensor_pipe_frac_issued=round(issued * achieved / pk["tf_sust"], 4),
                    launches_per_step=len(gemm), avg_launch_ms=round(lin_ms / len(gemm), 4),
                    share_of_step=shares)
        att = prof.get("eda_mha_split", [])
        if att:
            att_ms = sum(m for m, _ in att)
            att_fl = sum(4.0 * a[6] * a[8] * a[7] * a[7] * 64 for _, a in att)
            roof["attention"] = dict(bound="tensor", achieved=round(att_fl / (att_ms * 1e-3) / 1e12, 2),
                                     peak=pk["tf_sust"], unit="TFLOP/s", issued_mma_factor=issued,
                                     frac=round(att_fl / (att_ms * 1e-3) / 1e12 / pk["tf_sust"], 4),
                                     tensor_pipe_frac_issued=round(issued * att_fl / (att_ms * 1e-3) / 1e12 / pk["tf_sust"], 4),
                                     avg_launch_ms=round(att_ms / len(att), 4))
        lstm_ms = sum(m for m, _ in prof.get("eda_lstm_attractors", []))
        if lstm_ms:
            lstm_bytes = B * T * (4 * 4 * UNITS + 4)           # xproj row (4 gates x 256 fp32) + perm index
            roof["lstm"] = dict(bound="latency (T sequential steps)", achieved=round(lstm_bytes / (lstm_ms * 1e-3) / 1e9, 1),
                                peak=pk["hbm"], unit="GB/s", frac=round(lstm_bytes / (lstm_ms * 1e-3) / 1e9 / pk["hbm"], 4),
                                us_per_step=round(lstm_ms * 1e3 / (T + MAXSPK), 3))
        fe_ms = sum(m for k in ("eda_frontend_transform", "eda_frontend_splice_subsample_w") for m, _ in prof[k])
        fe_bytes = B * T * (4 * SHIFT * SUB + 4 * IN)            # 4580 B / model frame @8 kHz
        roof["frontend"] = dict(bound="hbm", achieved=round(fe_bytes / (fe_ms * 1e-3) / 1e9, 1), peak=pk["hbm"],
                                unit="GB/s", frac=round(fe_bytes / (fe_ms * 1e-3) / 1e9 / pk["hbm"], 4))
        # ---- CPU baseline on this box's host cores, bounded sample of the same workload ------
        cores = os.cpu_count() or 1
        # the reference's CPU path on this box's host cores, whole configs[1] batch per step; at N > 1 it is left to
        # the N = 1 run (no rank may idle in a barrier while rank 0 runs numpy)
        cpu = None
        if not args.skip_cpu and world == 1:
            if ORIG_AFFINITY:
                os.sched_setaffinity(0, ORIG_AFFINITY)      # the CPU leg may use every host core again
            arm, cpu_fps, cpu_dt, cpu_shipped = cpu_reference(B, 2, 1, cores)
            cpu = {"value": round(cpu_fps, 1), "unit": UNIT, "cores": cores, "kind": arm.kind,
                   "sample": f"all {B} recordings of the workload per step, 2 timed steps after 1 warm-up ({cpu_dt:.2f} s/step): "
                             f"{arm.what}",
                   "as_shipped": {"value": round(cpu_shipped, 1), "unit": UNIT, "cores": 1,
                                  "sample": "2 recordings, the driver's own loop (infer_eda.py:92-113: one recording, one chunk "
                                            "per model call) with torch.set_num_threads(1) (train.py:87-88)"}}
        out = {
            "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": WARM, "ms_per_step": round(ms_step, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": {"fp32": "f32", "bf16x3": "bf16x3(f32 accum)", "bf16": "bf16"}[precision],
            "data": "synthetic",
            "config": {"workload": f"configs[1]: EEND-EDA {LAYERS}-layer {UNITS}-d encoder, batch {B} x {T}-frame chunks "
                                   f"({B} x {SECONDS}s @ {RATE} Hz audio through the front end), 2 speakers, per GPU",
                       "global_batch": world * B, "frames_per_step": world * B * T, "precision": precision,
                       "parallelism": f"dp{world} (recordings sharded, no data-path collective; ONE NCCL gather per run of the thresholded "
                                      "decisions (16-bit masks kept per step in a device ring) to rank 0, timed separately: result_collection; in the e2e passes every rank lands its OWN recordings' results in its own "
                                      "pinned host memory, as the reference writes one output file per recording)",
                       "host": f"rank thread pinned to {numa}",
                       "l2": "two alternating 51 MB int16 audio batches + 524 MB of FFN intermediates per step exceed the 126 MB L2",
                       "pipeline": f"InferencePipeline depth {args.depth}: H2D / {args.compute_streams} alternating compute / D2H streams "
                                   "(the next batch's front end fills the SMs the LSTM recurrence leaves idle), "
                                   "host never waits on the batch just submitted"},
            "e2e": {"value": round(world * B * T / e2e_med, 1), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "passes": E2E_PASSES, "steps_per_pass": args.steps,
                    "min": round(world * B * T / e2e_worst, 1), "max": round(world * B * T / e2e_best, 1),
                    "stat": "median over passes of (max over ranks of host wall time per step)"},
            "result_collection": None if collect_ms is None else {
                "what": f"one NCCL gather to rank 0 of the {args.steps} timed steps' thresholded decisions as 16-bit masks",
                "ms": round(collect_ms, 3), "bytes_per_rank": int(ring_bytes.numel()), "inside_timed_region": False},
            "host_submit_ms": round(1e3 * pipe.stats["submit_s"] / max(1, pipe.stats["submits"]), 3),
            "host_submit_parts_ms": {"gather_hook": round(1e3 * pipe.stats["gather_s"] / max(1, pipe.stats["submits"]), 3),
                                     "result_staging": round(1e3 * pipe.stats["results_s"] / max(1, pipe.stats["submits"]), 3)},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof,
            "cpu_baseline": cpu,
        }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def run_reference(args):
    """The reference's CPU implementation of configs[1] on the box's host cores: all 64 recordings per step."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return None
    cores = os.cpu_count() or 1
    arm, fps, dt, shipped = cpu_reference(B, max(1, args.steps), max(1, min(args.warmup, 2)), cores)
    return {
        "impl": "reference", "metric": METRIC, "value": round(fps, 1), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt * 1e3, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32 (front end f64)", "data": "synthetic",
        "config": {"workload": f"configs[1]: EEND-EDA {LAYERS}-layer {UNITS}-d encoder, batch {B} x {T}-frame chunks "
                               f"({B} x {SECONDS}s @ {RATE} Hz audio through the front end), 2 speakers, on the host CPU",
                   "global_batch": B, "frames_per_step": B * T},
        "cpu_baseline": {"value": round(fps, 1), "unit": UNIT, "cores": cores, "kind": arm.kind,
                         "sample": f"all {B} recordings x {SECONDS}s per step ({dt:.2f} s/step): {arm.what}",
                         "as_shipped": {"value": round(shipped, 1), "unit": UNIT, "cores": 1,
                                        "sample": "2 recordings, the driver's own loop (one recording, one chunk per model call, "
                                                  "torch.set_num_threads(1))"}},
        "e2e": {"value": round(fps, 1), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="bf16x3", choices=["fp32", "bf16x3", "bf16"])
    ap.add_argument("--depth", type=int, default=3, help="batches in flight in the pipeline")
    ap.add_argument("--compute-streams", type=int, default=3, help="compute streams the pipeline alternates batches over")
    ap.add_argument("--skip-cpu", action="store_true", help="omit the cpu_baseline leg (profiling runs)")
    ap.add_argument("--workload", default="infer", choices=["infer", "train", "c3", "c4"],
                    help="infer = configs[1] (the headline metric); train = configs[4] training step (bench_train.py); "
                         "c3 / c4 = configs[2] / configs[3], whole-job strong-scaling runs (bench_long.py)")
    ap.add_argument("--recordings", type=int, default=0, help="c3 / c4: override the number of recordings")
    args = ap.parse_args()
    if args.workload in ("c3", "c4"):
        import bench_long
        argv = ["--workload", args.workload, "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup),
                "--precision", args.precision] + (["--skip-cpu"] if args.skip_cpu else []) + \
               (["--recordings", str(args.recordings)] if args.recordings else [])
        return bench_long.main(argv)
    if args.workload == "train":
        import bench_train
        argv = ["--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup), "--impl", args.impl,
                "--precision", args.precision if args.precision != "fp32" else "bf16x3"] + (["--skip-cpu"] if args.skip_cpu else [])
        return bench_train.main(argv)
    out = run_reference(args) if args.impl == "reference" else run_ours(args)
    if out is not None:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()

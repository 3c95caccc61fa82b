#!/usr/bin/env python
"""Throughput of EendEdaModel(encoder_type='conformer').infer on the configs[1] shape (64 x 500 frames, 4 layers):
frames/s from CUDA events + per-entry-point device time.  Usage: python scripts/conformer_bench.py [--steps K]"""
import argparse, json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.weights import synth_conformer_state_dict, synth_features
from speaker_diarization_b200 import _lib as L
from speaker_diarization_b200.eend_eda.models import EendEdaModel

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--layers", type=int, default=4)
ap.add_argument("--precision", default="bf16x3")
a = ap.parse_args()
B, T = 64, 500
dev = torch.device("cuda:0")
m = EendEdaModel(2, 345, 4, 256, a.layers, encoder_type="conformer", precision=a.precision)
m.load_state_dict(synth_conformer_state_dict(a.layers, 1))
m = m.to(dev).eval()
feats = torch.cat(synth_features([T] * B, 345, 0)).to(dev)
lens = [T] * B
for _ in range(3):
    m.infer_device(feats, lens)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.steps):
    m.infer_device(feats, lens)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.steps
L.PROFILE = {}
m.infer_device(feats, lens)
prof = L.profile_ms()
L.PROFILE = None
times = {k: round(sum(t for t, _ in v), 4) for k, v in prof.items()}
# CPU port on a bounded sample (8 of the 64 chunks), all host threads
from oracle.conformer_oracle import OracleEendConformer
o = OracleEendConformer(synth_conformer_state_dict(a.layers, 1))
src = synth_features([T] * 8, 345, 0)
o.infer(src)
t0 = time.perf_counter(); o.infer(src); cpu = 8 * T / (time.perf_counter() - t0)
print(json.dumps({"metric": "eend_eda_conformer_frames_per_sec", "value": round(B * T / (ms * 1e-3), 1), "ms_per_step": round(ms, 4),
                  "config": f"EendEdaModel conformer {a.layers} layers, {B} x {T} frames, {a.precision}, features resident",
                  "ms_by_entry_point": times, "cpu_port_frames_per_sec": round(cpu, 1), "cpu_threads": torch.get_num_threads()}))

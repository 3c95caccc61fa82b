#!/usr/bin/env python
"""Host-side cost of InferencePipeline.submit()/collect() per batch (configs[1]) against the GPU time."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as Bn
from oracle.weights import EdaConfig, synth_state_dict
from speaker_diarization_b200 import feature
from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
from speaker_diarization_b200.pipeline import InferencePipeline

dev = torch.device("cuda", 0)
model = TransformerEdaModel(2, Bn.IN, Bn.HEADS, Bn.UNITS, Bn.LAYERS, dim_feedforward=Bn.FF, precision="bf16x3")
model.load_state_dict(synth_state_dict(EdaConfig(n_layers=Bn.LAYERS), 777))
model = model.to(dev).eval()
fe = feature.FrontEnd(Bn.RATE, Bn.FRAME, Bn.SHIFT, Bn.CTX, Bn.SUB, device=dev)
pipe = InferencePipeline(model, fe, depth=3, compute_streams=3)
host = [Bn.pcm16(Bn.synth_audio_batch(Bn.B, i)).pin_memory() for i in range(2)]
ns = [Bn.SECONDS * Bn.RATE] * Bn.B
for _ in pipe.run((host[i & 1], ns) for i in range(6)):
    pass
torch.cuda.synchronize()
ts, tc = [], []
n = 30
t_all = time.perf_counter()
for i in range(n):
    if len(pipe._pending) >= pipe.depth:
        t0 = time.perf_counter(); pipe.collect(); tc.append(time.perf_counter() - t0)
    t0 = time.perf_counter(); pipe.submit(host[i & 1], ns); ts.append(time.perf_counter() - t0)
pipe.drain()
torch.cuda.synchronize()
t_all = time.perf_counter() - t_all
print(f"submit {1e3 * sum(ts) / len(ts):.3f} ms/batch (min {1e3 * min(ts):.3f}), collect (incl. waiting) {1e3 * sum(tc) / max(1, len(tc)):.3f} ms, "
      f"wall {1e3 * t_all / n:.3f} ms/batch")
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for i in range(10):
    if len(pipe._pending) >= pipe.depth:
        pipe.collect()
    pipe.submit(host[i & 1], ns)
pr.disable(); pipe.drain()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)

#!/usr/bin/env python
"""Timeline of the multi-stream pipeline: start/end of every C-ABI call (CUDA events on its own
stream) relative to one base event for a few steady-state batches, printed as a text Gantt chart
(one lane per batch) so the overlap between batches can be read off.
Usage: pipeline_timeline.py [depth] [compute_streams]"""
import collections
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as Bn  # noqa: E402
from oracle.weights import EdaConfig, synth_state_dict  # noqa: E402
from speaker_diarization_b200 import _lib as L, feature  # noqa: E402
from speaker_diarization_b200.eend_eda.models import TransformerEdaModel  # noqa: E402
from speaker_diarization_b200.pipeline import InferencePipeline  # noqa: E402

SHORT = {"eda_frontend_logmel": "F", "eda_frontend_splice_subsample": "f", "eda_pad_pack_split": "p",
         "eda_gemm_split": "G", "eda_gemm_qkv": "Q", "eda_mha_split": "A", "eda_lstm_attractors": "L",
         "eda_attractor_head": "h", "eda_posterior": "P"}


def main():
    depth = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    streams = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    dev = torch.device("cuda", 0)
    model = TransformerEdaModel(2, Bn.IN, Bn.HEADS, Bn.UNITS, Bn.LAYERS, dim_feedforward=Bn.FF, precision="bf16x3")
    model.load_state_dict(synth_state_dict(EdaConfig(n_layers=Bn.LAYERS), 777))
    model = model.to(dev).eval()
    fe = feature.FrontEnd(Bn.RATE, Bn.FRAME, Bn.SHIFT, Bn.CTX, Bn.SUB, device=dev)
    pipe = InferencePipeline(model, fe, depth=depth, compute_streams=streams)
    audio = [Bn.pcm16(Bn.synth_audio_batch(Bn.B, i)).to(dev) for i in range(2)]
    ns = [Bn.SECONDS * Bn.RATE] * Bn.B
    for _ in pipe.run((audio[i & 1], ns) for i in range(6)):
        pass
    torch.cuda.synchronize()
    base = torch.cuda.Event(enable_timing=True)
    base.record()
    n = 9
    L.PROFILE = {}
    for _ in pipe.run((audio[i & 1], ns) for i in range(n)):
        pass
    torch.cuda.synchronize()
    prof, L.PROFILE = L.PROFILE, None
    lanes = collections.defaultdict(list)            # batch -> [(start, end, letter)]
    for name, v in prof.items():
        k = len(v) // n                              # calls of this entry point per batch, in issue order
        for i, (e0, e1, _) in enumerate(v):
            lanes[i // k].append((base.elapsed_time(e0), base.elapsed_time(e1), SHORT.get(name, "?")))
    t0 = min(s for lane in lanes.values() for s, _, _ in lane)
    t1 = max(e for lane in lanes.values() for _, e, _ in lane)
    width = 220
    print(f"depth {depth}, {streams} compute streams: {n} batches in {t1 - t0:.2f} ms "
          f"({(t1 - t0) / n:.3f} ms/batch incl. fill and drain); one column = {(t1 - t0) / width * 1e3:.0f} us")
    for b in range(n):
        row = [" "] * width
        for s, e, c in sorted(lanes[b]):
            a, z = int((s - t0) / (t1 - t0) * (width - 1)), int((e - t0) / (t1 - t0) * (width - 1))
            for x in range(a, max(a, z) + 1):
                row[x] = c
        dur = {c: 0.0 for c in "FGQAL"}
        for s, e, c in lanes[b]:
            if c in dur:
                dur[c] += e - s
        first, last = min(s for s, _, _ in lanes[b]), max(e for _, e, _ in lanes[b])
        print(f"b{b} |{''.join(row)}| {last - first:.2f} ms  " + " ".join(f"{c}={dur[c]:.2f}" for c in "FGQAL"))


if __name__ == "__main__":
    main()

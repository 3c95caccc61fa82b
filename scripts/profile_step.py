#!/usr/bin/env python
"""ONE bench step (configs[1]) between cudaProfilerStart/Stop, for
  ncu --set full --clock-control none --import-source on --profile-from-start off \
      -o gpurun_out/step_full python scripts/profile_step.py
Every kernel of the step lands in one report.  Without ncu it just runs the step."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as Bn  # noqa: E402
from speaker_diarization_b200.synth import EdaConfig, synth_state_dict  # noqa: E402  (seeded weights only)
from speaker_diarization_b200 import feature  # noqa: E402
from speaker_diarization_b200.eend_eda.models import TransformerEdaModel  # noqa: E402


def main():
    precision = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
    dev = torch.device("cuda", 0)
    model = TransformerEdaModel(2, Bn.IN, Bn.HEADS, Bn.UNITS, Bn.LAYERS, dim_feedforward=Bn.FF, precision=precision)
    model.load_state_dict(synth_state_dict(EdaConfig(n_layers=Bn.LAYERS), 777))
    model = model.to(dev).eval()
    fe = feature.FrontEnd(Bn.RATE, Bn.FRAME, Bn.SHIFT, Bn.CTX, Bn.SUB, device=dev)
    audio = Bn.pcm16(Bn.synth_audio_batch(Bn.B, 0)).to(dev)
    ns = [Bn.SECONDS * Bn.RATE] * Bn.B
    torch.manual_seed(1234)

    def step():
        feats, rows = fe.features_flat(audio, ld=352, counts=ns)
        return model.infer_device(feats, rows, Bn.MAXSPK, 0.5)

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    post, probs, labels = step()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    print("ok", tuple(post.shape), float(post.float().mean()))


if __name__ == "__main__":
    main()

"""Time the front end alone (bench shape: 64 x 50 s @ 8 kHz)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from speaker_diarization_b200 import feature
dev = torch.device("cuda:0")
fe = feature.FrontEnd(8000, 200, 80, 7, 10, device=dev)
B, n = 64, 400000
g = torch.Generator().manual_seed(0)
audio = [(torch.randn((B, n), generator=g) * 0.1).to(dev) for _ in range(2)]
ns = [n] * B
for i in range(3):
    fe.features(audio[i & 1], ld=352, counts=ns)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(10):
    fe.features(audio[i & 1], ld=352, counts=ns)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"front end: {ms:.3f} ms / batch of {B} x 50 s  ({B * 500 * 4580 / ms / 1e6:.1f} GB/s algorithmic)")

"""Time eda_lstm_attractors alone for several batch sizes (device events)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from speaker_diarization_b200 import _lib as L, ops

dev = torch.device("cuda:0")
print("max active clusters:", L.load().eda_lstm_max_active_clusters())
T = int(sys.argv[1]) if len(sys.argv) > 1 else 500
g = torch.Generator().manual_seed(0)
for B in (4, 8, 16, 32, 48, 56, 64, 72, 128):
    xproj = (torch.randn((B * T, 1024), generator=g) * 0.5).to(dev)
    perm = torch.stack([torch.randperm(T, generator=g) for _ in range(B)]).to(torch.int32).to(dev)
    lens = torch.full((B,), T, dtype=torch.int32, device=dev)
    we = (torch.randn((1024, 256), generator=g) / 16).to(dev)
    wd = (torch.randn((1024, 256), generator=g) / 16).to(dev)
    bd = torch.zeros(1024, device=dev)
    for _ in range(3):
        ops.lstm_attractors(xproj, perm, lens, None, we, wd, bd, B, T, 15)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        ops.lstm_attractors(xproj, perm, lens, None, we, wd, bd, B, T, 15)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"B={B:4d} T={T}: {ms:.3f} ms  {ms * 1e3 / (T + 15):.3f} us/step")

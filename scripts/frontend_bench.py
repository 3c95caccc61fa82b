"""Front-end kernels alone on the bench shape (64 recordings x 50 s @ 8 kHz int16): time per pass, GB/s against the
algorithmic bytes of SURVEY 8(d) (4580 B per model frame), old one-warp-per-frame kernel beside the new one.
    python scripts/frontend_bench.py [rate]"""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speaker_diarization_b200 import feature, _lib as L

rate = int(sys.argv[1]) if len(sys.argv) > 1 else 8000
fs, hop = (200, 80) if rate == 8000 else (400, 160)
B, seconds = 64, 50
n = rate * seconds
g = torch.Generator().manual_seed(0)
audio = (torch.randn((B, n), generator=g) * 3000).clamp_(-32768, 32767).to(torch.int16).cuda()
fe = feature.FrontEnd(rate, fs, hop, 7, 10)
counts = [n] * B

def run(label):
    for _ in range(3):
        fe.features_flat(audio, ld=352, counts=counts)
    L.PROFILE = {}
    for _ in range(10):
        fe.features_flat(audio, ld=352, counts=counts)
    prof = L.profile_ms(); L.PROFILE = None
    p1 = sorted(m for m, _ in prof["eda_frontend_transform"])[5]
    p2 = sorted(m for m, _ in prof["eda_frontend_splice_subsample_w"])[5]
    rows = B * (seconds * rate // hop // 10)
    byt = rows * (4 * hop * 10 + 4 * 345)
    print(json.dumps(dict(kernel=label, rate=rate, pass1_us=round(p1 * 1e3, 1), pass2_us=round(p2 * 1e3, 1),
                          algorithmic_GBps=round(byt / ((p1 + p2) * 1e-3) / 1e9, 1))))

run("new")
os.environ["EDA_B200_FRONTEND"] = "warp"
run("warp-per-frame (previous)")

"""Where does the HOST time of InferencePipeline.submit() go?  (cProfile of 40 submits of the bench batch)
    python scripts/host_profile.py"""
import cProfile, pstats, os, sys, io
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench as B0
from speaker_diarization_b200 import feature
from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
from speaker_diarization_b200.pipeline import InferencePipeline
from speaker_diarization_b200.synth import EdaConfig, synth_state_dict
dev = torch.device("cuda", 0)
model = TransformerEdaModel(2, 345, 4, 256, 4, precision="bf16x3")
model.load_state_dict(synth_state_dict(EdaConfig(n_layers=4), 777))
model = model.to(dev).eval()
fe = feature.FrontEnd(8000, 200, 80, 7, 10, device=dev)
pipe = InferencePipeline(model, fe)
audio = B0.pcm16(B0.synth_audio_batch(64, 0)).to(dev)
ns = [400000] * 64
def run(n):
    for _ in pipe.run((audio, ns) for _ in range(n)):
        pass
run(6)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
run(40)
pr.disable()
torch.cuda.synchronize()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45)
print(s.getvalue())
print("host_submit_ms", 1e3 * pipe.stats["submit_s"] / pipe.stats["submits"])

"""Where do thresholded labels differ from the reference's on the full-size goldens, and by how much?
(diagnostic; run on the GPU box: python scripts/flip_probe.py)"""
import json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from full_cases import model_full_case, unpack_bits
from speaker_diarization_b200.eend_eda.models import TransformerEdaModel
manifest = json.load(open(os.path.join(ROOT, "tests/golden/manifest.json")))
for name in ("full_c2", "full_c5"):
    case, g, cfg, sd, src, ts = model_full_case(name, manifest)
    for prec in ("bf16x3", "fp32"):
        m = TransformerEdaModel(case["n_speakers"], 345, 4, 256, case["n_layers"], precision=prec)
        m.load_state_dict(sd); m = m.cuda().eval()
        torch.manual_seed(1000 + case["data_seed"])
        m.infer([s.cuda() for s in src], attractor_threshold=-1.0, activity_threshold=0.5)
        post = m.last["posteriors"].cpu().numpy()
        got = post if case["keep"] is None else post[case["keep"]]
        err = np.abs(got - g["post_all"])
        ref_lab = unpack_bits(g["labels_packed"], post.shape)
        fl = m.last["labels"].cpu().numpy().astype(bool) != ref_lab
        print(name, prec, "max err %.2e mean err %.2e p99.9 %.2e" % (err.max(), err.mean(), np.quantile(err, 0.999)),
              "flips", int(fl.sum()), "of", fl.size, "|ours-0.5| at flips", np.abs(post[fl] - 0.5))
        near = np.abs(g["post_all"] - 0.5)
        print("   reference posteriors within 1e-6/1e-5/1e-4 of 0.5:", int((near < 1e-6).sum()), int((near < 1e-5).sum()), int((near < 1e-4).sum()), "of", near.size)

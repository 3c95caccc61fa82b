"""Pin the RTTM post-processing to the UNMODIFIED reference script and close the DER loop.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Build container only (``python -m oracle.make_golden_rttm``).

``speaker_diarization/bin/make_rttm.py`` is executed as it is (``runpy``, its own argparse, its own file
handling) with ONE stand-in: ``h5py`` is not installed, so a module offering ``h5py.File(path)['T_hat'][:]``
from in-memory arrays is registered for the duration of the run.  The text the script writes is frozen in
tests/golden/rttm.json; inputs are regenerated from seeds at test time.  For the long-recording case
(the reference driver's own T_hat of tests/golden/full_c4_counts.npz) the RTTM is also scored against the
synthetic mixtures' true segments with the reference's vendored NIST scorer
(``SCTK-2.4.12/src/md-eval/md-eval.pl -c 0.25``, perl), and the DER is frozen beside it."""
from __future__ import annotations

import json
import os
import re
import runpy
import subprocess
import sys
import tempfile
import types

import numpy as np

REF_ROOT = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")
MD_EVAL = os.path.join(REF_ROOT, "SCTK-2.4.12", "src", "md-eval", "md-eval.pl")

CASES = {
    # name: (lens, n_spk, seed, script arguments)
    "smooth_default": ([1500, 700, 2300], 3, 1, []),
    "smooth_med11_8k": ([1500, 700, 2300], 3, 1, ["--median", "11", "--frame_shift", "80", "--subsampling", "10",
                                                   "--sampling_rate", "8000"]),
    "smooth_thr_med5": ([64, 4097, 33], 2, 2, ["--threshold", "0.6", "--median", "5"]),
}
C4_ARGS = ["--median", "11", "--frame_shift", "80", "--subsampling", "10", "--sampling_rate", "8000"]


def run_make_rttm(arrays: dict, extra_args) -> str:
    """bin/make_rttm.py:1-42 unmodified on {session: T_hat}."""
    class _File(dict):
        def __init__(self, path, mode="r"):
            super().__init__(T_hat=arrays[os.path.splitext(os.path.basename(path))[0]])
    stub = types.ModuleType("h5py")
    stub.File = _File
    old_mod, old_argv = sys.modules.get("h5py"), sys.argv
    sys.modules["h5py"] = stub
    try:
        with tempfile.TemporaryDirectory() as tmp:
            lst, out = os.path.join(tmp, "files.txt"), os.path.join(tmp, "out.rttm")
            with open(lst, "w") as f:
                f.write("".join(f"/some/dir/{k}.h5\n" for k in arrays))
            sys.argv = ["make_rttm.py", lst, out] + list(extra_args)
            runpy.run_path(os.path.join(REF_ROOT, "speaker_diarization", "bin", "make_rttm.py"), run_name="__main__")
            return open(out).read()
    finally:
        sys.argv = old_argv
        if old_mod is None:
            del sys.modules["h5py"]
        else:
            sys.modules["h5py"] = old_mod


def md_eval(ref_rttm: str, sys_rttm: str, collar: float = 0.25) -> float:
    """DER (percent) by the reference's vendored NIST md-eval.pl."""
    with tempfile.TemporaryDirectory() as tmp:
        r, s = os.path.join(tmp, "ref.rttm"), os.path.join(tmp, "sys.rttm")
        open(r, "w").write(ref_rttm)
        open(s, "w").write(sys_rttm)
        out = subprocess.run(["perl", MD_EVAL, "-c", str(collar), "-r", r, "-s", s], capture_output=True, text=True).stdout
    # the reference's copy of md-eval.pl prints "DER/MS/FA/SC" on one line; stock NIST copies print the long report
    m = re.search(r"OVERALL SPEAKER DIARIZATION ERROR = ([0-9.]+) percent", out) or re.search(r"^([0-9.]+)/[0-9.]+/[0-9.]+/[0-9.]+\s*$", out, re.M)
    assert m, out[-2000:]
    return float(m.group(1))


def c4_truth(manifest) -> str:
    from oracle import feature_oracle as fo
    from oracle.rttm_oracle import truth_rttm
    case = manifest["cases"]["full_c4_counts"]
    txt = ""
    for rec in sorted(case["recs"]):
        seconds, seed, n = case["recs"][rec]
        _, segs = fo.synth_audio(int(round(seconds * case["rate"])), case["rate"], max(int(n), 1), seed, dtype=np.float64)
        txt += truth_rttm(rec, segs)
    return txt


def main():
    sys.path.insert(0, os.path.dirname(HERE))
    from oracle.rttm_oracle import smooth_posteriors
    manifest = json.load(open(os.path.join(GOLD, "manifest.json")))
    out = {}
    for name, (lens, n_spk, seed, args) in CASES.items():
        arrs = smooth_posteriors(lens, n_spk, seed)
        sessions = {f"sess{len(lens) - i}": a for i, a in enumerate(arrs)}          # written in SORTED order (:23)
        out[name] = dict(lens=lens, n_spk=n_spk, seed=seed, args=args, sessions=list(sessions),
                         text=run_make_rttm(sessions, args))
        print(name, len(out[name]["text"].splitlines()), "lines")
    g = np.load(os.path.join(GOLD, "full_c4_counts.npz"))
    t_hats = {k[len("T_hat_"):]: g[k] for k in g.files if k.startswith("T_hat_")}
    text = run_make_rttm(t_hats, C4_ARGS)
    truth = c4_truth(manifest)
    der = md_eval(truth, text)
    out["c4_driver"] = dict(args=C4_ARGS, sessions=list(t_hats), text=text, md_eval_der_percent=der, collar=0.25)
    print("c4_driver", len(text.splitlines()), "lines, md-eval DER", der)
    with open(os.path.join(GOLD, "rttm.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()

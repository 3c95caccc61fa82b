#!/usr/bin/env python
"""Per-CTA clock64 timeline of the attention kernel (probe build with -DEDA_MHA_TRACE).

  python scripts/mha_trace.py build     # here (no GPU): scripts/_mha_trace.so
  python scripts/mha_trace.py           # on the GPU box: prints where a CTA's cycles go
"""
import ctypes as C
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SO = os.path.join(ROOT, "scripts", "_mha_trace.so")


def build():
    from speaker_diarization_b200 import build as B
    B.build()
    obj = os.path.join(ROOT, "scripts", "_mha_trace.o")
    subprocess.check_call([B.nvcc(), *B.NVCC_FLAGS, "-DEDA_MHA_TRACE", "-c", os.path.join(B.CSRC, "mha_tc.cu"), "-o", obj])
    others = [os.path.join(B.OBJ, os.path.basename(c)[:-3] + ".o") for c in glob.glob(os.path.join(B.CSRC, "*.cu"))
              if not c.endswith("mha_tc.cu")]
    subprocess.check_call([B.nvcc(), "-shared", "-o", SO, obj, *others, "-gencode", "arch=compute_100a,code=sm_100a",
                           "-Xcompiler", "-fPIC", "-lcudart"])
    os.remove(obj)
    print(SO)


def main():
    import numpy as np
    import torch
    lib = C.CDLL(SO)
    B, T, H = 64, 500, 4
    D, Tp = H * 64, 504
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(1)
    qk = torch.randn(B * T, 2 * D, device=dev, generator=g)
    vt = torch.randn(B * D, Tp, device=dev, generator=g)
    qk_hi = qk.bfloat16(); qk_lo = (qk - qk_hi.float()).bfloat16()
    vt_hi = vt.bfloat16(); vt_lo = (vt - vt_hi.float()).bfloat16()
    out_hi = torch.empty(B * T, D, device=dev, dtype=torch.bfloat16); out_lo = torch.empty_like(out_hi)
    p = C.c_void_p
    lib.eda_mha_split.argtypes = [p, p, C.c_int, p, p, C.c_int, C.c_int, C.c_int, C.c_int, p, C.c_int, p, p, C.c_int, C.c_int, p]
    for _ in range(3):
        rc = lib.eda_mha_split(qk_hi.data_ptr(), qk_lo.data_ptr(), 2 * D, vt_hi.data_ptr(), vt_lo.data_ptr(), Tp, B, T, H,
                               None, 0, out_hi.data_ptr(), out_lo.data_ptr(), D, 1, None)
        assert rc == 0
    torch.cuda.synchronize()
    n_cta = min(4 * H * B, 2 * torch.cuda.get_device_properties(0).multi_processor_count)
    buf = np.zeros((n_cta, 96), dtype=np.int64)
    lib.eda_mha_trace_read.argtypes = [p, C.c_int]
    assert lib.eda_mha_trace_read(buf.ctypes.data, n_cta) == 0
    life = buf[:, 2] - buf[:, 1]
    print("CTA lifetime cycles: mean %.0f  min %d  max %d ; items per CTA %d..%d ; cycles per item %.0f" % (
        life.mean(), life.min(), life.max(), buf[:, 3].min(), buf[:, 3].max(), (life / buf[:, 3]).mean()))
    print("(per-tile stamps below are each CTA's FIRST item)")
    nk = 8
    mma = buf[:, 8:8 + 4 * nk].reshape(n_cta, nk, 4)
    sm = buf[:, 48:48 + 2 * nk].reshape(n_cta, nk, 2)
    t0 = buf[:, 1:2]
    print("tile   MMA: S(j+1) issued | P(j) ready | V ready | PV issued    softmax: S ready | P arrived   (mean cycles since CTA start)")
    for j in range(nk):
        m = (mma[:, j, :] - t0).mean(0)
        s = (sm[:, j, :] - t0).mean(0)
        print("%3d   %9.0f %9.0f %9.0f %9.0f      %9.0f %9.0f" % (j, *m, *s))
    print("O ready %.0f   end %.0f" % ((buf[:, 80] - buf[:, 1]).mean(), life.mean()))
    print("mean softmax busy per tile (S ready -> P arrived): %.0f" % (sm[:, :, 1] - sm[:, :, 0]).mean())
    print("mean wait for P in the MMA thread: %.0f ; for V: %.0f" % ((mma[:, :, 1] - mma[:, :, 0]).mean(), (mma[:, :, 2] - mma[:, :, 1]).mean()))
    print("mean softmax wait for S (P arrived(j-1) -> S ready(j)): %.0f" % (sm[:, 1:, 0] - sm[:, :-1, 1]).mean())
    # one SM's schedule
    smid = buf[:, 0]
    k = int(smid[0])
    sel = np.where(smid == k)[0]
    base = buf[sel, 1].min()
    print("SM %d ran %d CTAs:" % (k, len(sel)))
    for i in sel[np.argsort(buf[sel, 1])]:
        print("   cta %4d  start %7d  tmem/first S issue %7d  end %7d" % (i, buf[i, 1] - base, buf[i, 8] - base, buf[i, 2] - base))
    span = np.array([buf[smid == s, 2].max() - buf[smid == s, 1].min() for s in np.unique(smid)])
    print("per-SM span cycles: mean %.0f max %d ; CTAs per SM min %d max %d" % (
        span.mean(), span.max(), np.bincount(smid).min(), np.bincount(smid).max()))


if __name__ == "__main__":
    build() if sys.argv[1:] == ["build"] else main()

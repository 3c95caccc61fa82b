#!/usr/bin/env python
"""Per-step clock64 timeline of the LSTM cluster kernels (probe build with -DEDA_LSTM_TRACE).

  python scripts/lstm_trace.py build            # here (no GPU): scripts/_lstm_trace.so
  python scripts/lstm_trace.py [B] [T] [prec]   # on the GPU box: where one recurrence step's cycles go
                                                # prec 0 = fp32 CUDA-core kernel, 1 = tensor-core kernel
"""
import ctypes as C
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SO = os.path.join(ROOT, "scripts", "_lstm_trace.so")


def build():
    from speaker_diarization_b200 import build as B
    B.build()
    obj = os.path.join(ROOT, "scripts", "_lstm_trace.o")
    subprocess.check_call([B.nvcc(), *B.NVCC_FLAGS, "-DEDA_LSTM_TRACE", "-c", os.path.join(B.CSRC, "lstm_cluster.cu"), "-o", obj])
    others = [os.path.join(B.OBJ, os.path.basename(c)[:-3] + ".o") for c in glob.glob(os.path.join(B.CSRC, "*.cu"))
              if not c.endswith("lstm_cluster.cu")]
    subprocess.check_call([B.nvcc(), "-shared", "-o", SO, obj, *others, "-gencode", "arch=compute_100a,code=sm_100a",
                           "-Xcompiler", "-fPIC", "-lcudart"])
    os.remove(obj)
    print(SO)


def main():
    import numpy as np
    import torch
    lib = C.CDLL(SO)
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 500
    prec = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    D, n_dec = 256, 3
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(1)
    xproj = torch.randn(B * T, 4 * D, device=dev, generator=g)
    perm = torch.stack([torch.randperm(T) for _ in range(B)]).to(torch.int32).to(dev)
    lens = torch.full((B,), T, dtype=torch.int32, device=dev)
    w_enc = torch.randn(4 * D, D, device=dev, generator=g) / 16
    w_dec = torch.randn(4 * D, D, device=dev, generator=g) / 16
    bias = torch.randn(4 * D, device=dev, generator=g)
    att = torch.empty(B, n_dec, D, device=dev)
    p = C.c_void_p
    lib.eda_lstm_attractors.argtypes = [p] * 7 + [C.c_int] * 4 + [p, p, C.c_int, p]
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    for it in range(4):
        if it == 3:
            ev[0].record()
        rc = lib.eda_lstm_attractors(xproj.data_ptr(), perm.data_ptr(), lens.data_ptr(), None, w_enc.data_ptr(),
                                     w_dec.data_ptr(), bias.data_ptr(), B, T, D, n_dec, att.data_ptr(), None, prec, None)
        assert rc == 0
    ev[1].record()
    torch.cuda.synchronize()
    ms = ev[0].elapsed_time(ev[1])
    print("B %d T %d prec %d: %.3f ms, %.3f us/step" % (B, T, prec, ms, 1e3 * ms / (T + n_dec)))
    lib.eda_lstm_max_active_clusters.restype = C.c_int
    mc = lib.eda_lstm_max_active_clusters()
    kG = 4
    while (B + kG - 1) // kG > mc and kG < 8:
        kG += 1
    if prec != 0:
        kG = min(8, max(4, B))      # the tensor-core inference launch packs 8 sequences per cluster
    n_cta = 8 * ((B + kG - 1) // kG)
    buf = np.zeros((n_cta, 8, 8), dtype=np.int64)
    lib.eda_lstm_trace_read.argtypes = [p, C.c_int]
    assert lib.eda_lstm_trace_read(buf.ctypes.data, n_cta) == 0
    names = ["wake -> mat-vec done", "-> reduced", "-> cell done", "-> staged + synced", "-> pushes issued"]
    d = np.diff(buf[:, :, :6], axis=2)                 # (cta, step, 5)
    print("kG %d, %d CTAs; mean cycles per phase (thread 0 of each CTA, steps 64..71):" % (kG, n_cta))
    for i, n in enumerate(names):
        print("  %-24s %7.0f  (min %d max %d)" % (n, d[:, :, i].mean(), d[:, :, i].min(), d[:, :, i].max()))
    gap = buf[:, 1:, 0] - buf[:, :-1, 5]               # pushes issued -> next wake
    print("  %-24s %7.0f  (min %d max %d)" % ("-> next wake (exchange)", gap.mean(), gap.min(), gap.max()))
    step = buf[:, 1:, 0] - buf[:, :-1, 0]
    print("  step total %.0f cycles" % step.mean())


if __name__ == "__main__":
    build() if len(sys.argv) > 1 and sys.argv[1] == "build" else main()

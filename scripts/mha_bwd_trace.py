#!/usr/bin/env python
"""Per-tile clock64 timeline of the attention backward kernel (probe build with -DEDA_MHABWD_TRACE).

  python scripts/mha_bwd_trace.py build        # here (no GPU): scripts/_mhabwd_trace.so
  python scripts/mha_bwd_trace.py [B] [T] [p]  # on the GPU box: where one query tile's cycles go (p = dropout, 0 = off)
"""
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SO = os.path.join(ROOT, "scripts", "_mhabwd_trace.so")


def build():
    from speaker_diarization_b200 import build as B
    B.build()
    obj = os.path.join(ROOT, "scripts", "_mhabwd_trace.o")
    subprocess.check_call([B.nvcc(), *B.NVCC_FLAGS, "-DEDA_MHABWD_TRACE", "-c", os.path.join(B.CSRC, "mha_bwd_tc.cu"), "-o", obj])
    others = [os.path.join(B.OBJ, os.path.basename(c)[:-3] + ".o") for c in glob.glob(os.path.join(B.CSRC, "*.cu"))
              if not c.endswith("mha_bwd_tc.cu")]
    subprocess.check_call([B.nvcc(), "-shared", "-o", SO, obj, *others, "-gencode", "arch=compute_100a,code=sm_100a",
                           "-Xcompiler", "-fPIC", "-lcudart"])
    os.remove(obj)
    print(SO)


def main():
    import ctypes as C
    import numpy as np
    import torch
    from speaker_diarization_b200 import _lib as L
    L.LIB_PATH = SO                       # the bindings load the probe build
    from speaker_diarization_b200 import ops, ops_train as OT
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
    p = float(sys.argv[3]) if len(sys.argv) > 3 else 0.1
    H, D = 4, 256
    M, Tp = B * T, (T + 7) // 8 * 8
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(1)
    sp = lambda x: ops.Split(x.bfloat16(), (x - x.bfloat16().float()).bfloat16())
    x = torch.randn(M, D, device=dev, generator=g)
    w = (torch.rand(3 * D, D, device=dev, generator=g) * 2 - 1) * 0.15
    b = torch.zeros(3 * D, device=dev)
    qk, vt = ops.Split.empty(M, 2 * D, dev), ops.Split.empty(B * H * 64, Tp, dev)
    vt.hi.zero_(); vt.lo.zero_()
    P = L.PREC_BF16X3
    ops.gemm_qkv(sp(x), sp(w), b, qk, vt, B, T, Tp, D, H, P)
    ctx, lse = ops.Split.empty(M, D, dev), torch.empty((B * H, T), device=dev)
    drop = (p, 1234, 3) if p > 0 else OT.NO_DROP
    bits = None
    if p > 0:
        bits = torch.empty((B * H * T, (T + 127) // 128 * 4), dtype=torch.int32, device=dev)
        OT.dropout_bits(p, 1234, 3, B * H * T, T, bits)
    OT.mha_split_train(qk, vt, B, T, Tp, H, P, ctx, lse, drop=drop, drop_bits=bits)
    do = sp(torch.randn(M, D, device=dev, generator=g))
    delta = torch.empty((M, H), device=dev)
    OT.mha_delta(do, ctx, M, D, delta)
    dq_acc, dqkv = torch.empty((M, D), device=dev), ops.Split.empty(M, 3 * D, dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    for it in range(3):
        if it == 2:
            ev[0].record()
        OT.mha_bwd(qk, vt, Tp, do, lse, delta, B, T, H, dq_acc, dqkv, P, drop=drop, drop_bits=bits)
    ev[1].record()
    torch.cuda.synchronize()
    n_t = (T + 127) // 128
    n_cta = n_t * H * B
    print("B %d T %d p %.2f: %.3f ms, %d CTAs x %d query tiles" % (B, T, p, ev[0].elapsed_time(ev[1]), n_cta, n_t))
    lib = L.load()
    buf = np.zeros((min(n_cta, 4096), 16), dtype=np.int64)
    lib.eda_mha_bwd_trace_read.argtypes = [C.c_void_p, C.c_int]
    assert lib.eda_mha_bwd_trace_read(buf.ctypes.data, buf.shape[0]) == 0
    t0 = buf[:, 0:1]
    names = {1: "Q/dO landed (MMA thread)", 2: "S, dP issued", 8: "S, dP committed (softmax warp)", 9: "first 32 keys done",
             10: "P/dS stored + arrived", 3: "MMA thread saw P/dS", 4: "gradient MMAs issued", 11: "dQ committed",
             12: "dQ read out of TMEM", 13: "reds issued"}
    print("mean cycles since the TMA producer issued tile 3's loads (i.e. since the gradient MMAs of tile 2 completed):")
    for k in (1, 2, 8, 9, 10, 3, 4, 11, 12, 13):
        d = (buf[:, k:k + 1] - t0)
        print("  %-36s %7.0f  (min %d max %d)" % (names[k], d.mean(), d.min(), d.max()))


if __name__ == "__main__":
    build() if len(sys.argv) > 1 and sys.argv[1] == "build" else main()

"""Time the split-bf16 GEMM shapes of the C2 workload in isolation (device events)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from speaker_diarization_b200 import _lib as L, ops
dev = torch.device("cuda:0")
M = int(sys.argv[1]) if len(sys.argv) > 1 else 32000
g = torch.Generator().manual_seed(0)
def mk(r, c):
    return ops.split_bf16((torch.randn((r, c), generator=g) * 0.5).to(dev))
shapes = [("qkv", 768, 256, "qkv"), ("out_proj", 256, 256, L.EPI_BIAS_RES_LN), ("ffn1", 2048, 256, L.EPI_BIAS_RELU),
          ("ffn2", 256, 2048, L.EPI_BIAS_RES_LN), ("xproj", 1024, 256, L.EPI_BIAS), ("in_lin", 256, 352, L.EPI_BIAS_LN)]
B, T = M // 500, 500
PREC = {"bf16x3": L.PREC_BF16X3, "bf16": L.PREC_BF16}[os.environ.get("PREC", "bf16x3")]
PASSES = 3 if PREC == L.PREC_BF16X3 else 1
for name, N, K, epi in shapes:
    a, w = mk(M, K), mk(N, K)
    bias = torch.zeros(N, device=dev); gm = torch.ones(N, device=dev); bt = torch.zeros(N, device=dev)
    resid = torch.randn((M, 256), generator=g).to(dev)
    outf = torch.empty((M, N), device=dev); outs = ops.Split.empty(M, N, dev)
    qk = ops.Split.empty(M, 512, dev); vt = ops.Split.empty(B * 256, 504, dev)
    def run():
        if epi == "qkv":
            ops.gemm_qkv(a, w, bias, qk, vt, B, T, 504, 256, 4, PREC)
        elif epi == L.EPI_BIAS_RELU:
            ops.gemm_split(a, w, bias, M, N, K, epi, PREC, out_split=outs)
        elif epi == L.EPI_BIAS:
            ops.gemm_split(a, w, bias, M, N, K, epi, PREC, out=outf)
        else:
            ops.gemm_split(a, w, bias, M, N, K, epi, PREC, resid=resid, gamma=gm, beta=bt, out=outf, out_split=outs)
    for _ in range(3):
        run()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    fl = 2.0 * M * N * min(K, 345 if K == 352 else K)
    print(f"{name:9s} M={M} N={N} K={K}: {ms * 1e3:7.1f} us  {fl / ms / 1e9:7.1f} TF algorithmic, {PASSES * fl / ms / 1e9:7.1f} TF issued")

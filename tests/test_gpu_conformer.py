"""Conformer variant: per-kernel parity of the new pieces and the whole
EendEdaModel(encoder_type='conformer') against the oracle and the reference's frozen outputs."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from conftest import cases_of, load_golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _split(x, dev):
    from speaker_diarization_b200 import ops
    return ops.split_bf16(x.to(dev).contiguous())


@pytest.mark.parametrize("M", [1, 130, 3000])
def test_layernorm_split(dev, M):
    from speaker_diarization_b200 import ops
    g = torch.Generator().manual_seed(M)
    x = torch.randn((M, 256), generator=g) * 3 + 1
    gm, bt = 1 + 0.1 * torch.randn((256,), generator=g), 0.1 * torch.randn((256,), generator=g)
    ref = F.layer_norm(x, (256,), gm, bt, 1e-5)
    out, sp = torch.empty((M, 256), device=dev), ops.Split.empty(M, 256, dev)
    ops.layernorm_split(x.to(dev), gm.to(dev), bt.to(dev), out_split=sp, out=out)
    assert (out.cpu() - ref).abs().max().item() < 5e-6
    assert (sp.float().cpu() - ref).abs().max().item() < 6e-5     # a split pair carries 16 mantissa bits: 2^-17 x |x| <= 4


@pytest.mark.parametrize("B,T", [(2, 100), (3, 64), (1, 31), (2, 7)])
def test_dwconv_bn_silu(dev, B, T):
    from speaker_diarization_b200 import ops
    g = torch.Generator().manual_seed(B * 100 + T)
    x = torch.randn((B, T, 256), generator=g)
    w = torch.randn((256, 1, 31), generator=g) / 31 ** 0.5
    bias, bw, bb = (torch.randn((256,), generator=g) * 0.1 for _ in range(3))
    mean, var = torch.randn((256,), generator=g) * 0.1, 0.5 + torch.rand((256,), generator=g)
    y = F.conv1d(x.transpose(1, 2), w, bias, padding=15, groups=256)
    ref = F.silu(F.batch_norm(y, mean, var, 1 + bw, bb, False, 0.0, 1e-5)).transpose(1, 2)
    scale = (1 + bw) / torch.sqrt(var + 1e-5)
    shift = (bias - mean) * scale + bb
    sp = ops.Split.empty(B * T, 256, dev)
    ops.dwconv_bn_silu(x.reshape(B * T, 256).to(dev), w.view(256, 31).contiguous().to(dev), scale.to(dev), shift.to(dev), B, T, sp)
    assert (sp.float().cpu().view(B, T, 256) - ref).abs().max().item() < 6e-5


@pytest.mark.parametrize("prec", ["bf16x3", "bf16"])
@pytest.mark.parametrize("M", [300, 1000])
def test_gemm_silu_glu_scaled_residual(dev, prec, M):
    from speaker_diarization_b200 import _lib as L, ops
    P = L.PRECISIONS[prec]
    tol = {"bf16x3": 1e-4, "bf16": 6e-2}[prec]
    g = torch.Generator().manual_seed(M)
    a = torch.randn((M, 256), generator=g)
    w1, b1 = (torch.rand((2048, 256), generator=g) * 2 - 1) / 16, torch.randn((2048,), generator=g) * 0.1
    sp = ops.Split.empty(M, 2048, dev)
    ops.gemm_split_ex(_split(a, dev), _split(w1, dev), b1.to(dev), M, 2048, 256, L.EPI_BIAS_SILU, P, out_split=sp)
    assert (sp.float().cpu() - F.silu(F.linear(a, w1, b1))).abs().max().item() < tol
    # GLU: rows interleaved in blocks of 32 exactly like EendEdaModel._prepared does
    wp, bp = (torch.rand((512, 256), generator=g) * 2 - 1) / 16, torch.randn((512,), generator=g) * 0.1
    idx = torch.arange(256).view(8, 32)
    order = torch.stack([idx, idx + 256], dim=1).reshape(-1)
    out = torch.empty((M, 256), device=dev)
    ops.gemm_split_ex(_split(a, dev), _split(wp[order].contiguous(), dev), bp[order].contiguous().to(dev), M, 512, 256,
                      L.EPI_BIAS_GLU, P, out=out)
    assert (out.cpu() - F.glu(F.linear(a, wp, bp), dim=-1)).abs().max().item() < tol
    # half-step residual, with and without the final LayerNorm
    h = torch.randn((M, 2048), generator=g)
    w2, b2 = (torch.rand((256, 2048), generator=g) * 2 - 1) / 45, torch.randn((256,), generator=g) * 0.1
    r = torch.randn((M, 256), generator=g)
    gm, bt = 1 + 0.1 * torch.randn((256,), generator=g), 0.1 * torch.randn((256,), generator=g)
    ref = 0.5 * F.linear(h, w2, b2) + r
    o1, o2 = torch.empty((M, 256), device=dev), torch.empty((M, 256), device=dev)
    ops.gemm_split_ex(_split(h, dev), _split(w2, dev), b2.to(dev), M, 256, 2048, L.EPI_BIAS_RES, P, resid=r.to(dev), out=o1, out_scale=0.5)
    ops.gemm_split_ex(_split(h, dev), _split(w2, dev), b2.to(dev), M, 256, 2048, L.EPI_BIAS_RES_LN, P, resid=r.to(dev),
                      gamma=gm.to(dev), beta=bt.to(dev), out=o2, out_scale=0.5)
    assert (o1.cpu() - ref).abs().max().item() < tol
    assert (o2.cpu() - F.layer_norm(ref, (256,), gm, bt, 1e-5)).abs().max().item() < tol


@pytest.mark.parametrize("lens", [[200, 130, 64], [500, 1, 377], [64, 65]])
def test_masked_attention(dev, lens):
    from speaker_diarization_b200 import _lib as L, ops
    B, T, H, D = len(lens), max(lens), 4, 256
    M, Tp = B * T, (T + 7) // 8 * 8
    g = torch.Generator().manual_seed(sum(lens))
    x = torch.randn((M, D), generator=g)
    w = (torch.rand((3 * D, D), generator=g) * 2 - 1) / 8
    b = torch.randn((3 * D,), generator=g) * 0.1
    qkv = F.linear(x.view(B, T, D), w, b).view(B, T, 3, H, 64)
    q, k, v = (qkv[:, :, i].transpose(1, 2) for i in range(3))
    s = q @ k.transpose(-1, -2) / 8.0
    pad = torch.arange(T).unsqueeze(0) >= torch.tensor(lens).unsqueeze(1)
    ref = (torch.softmax(s.masked_fill(pad[:, None, None, :], float("-inf")), -1) @ v).transpose(1, 2).reshape(M, D)
    qk, vt = ops.Split.empty(M, 2 * D, dev), ops.Split.empty(B * H * 64, Tp, dev)
    ops.gemm_qkv(_split(x, dev), _split(w, dev), b.to(dev), qk, vt, B, T, Tp, D, H, L.PREC_BF16X3)
    out = ops.Split.empty(M, D, dev)
    ops.mha_split_masked(qk, vt, B, T, Tp, H, torch.tensor(lens, dtype=torch.int32, device=dev), L.PREC_BF16X3, out)
    assert (out.float().cpu() - ref).abs().max().item() < 6e-5


@pytest.mark.parametrize("prec", ["bf16x3"])
@pytest.mark.parametrize("name", cases_of("conformer"))
def test_conformer_model_matches_oracle_and_reference(dev, manifest, name, prec):
    from oracle.conformer_oracle import OracleEendConformer
    from oracle.weights import synth_conformer_state_dict, synth_features, synth_labels
    from speaker_diarization_b200.eend_eda.models import EendEdaModel
    case = manifest["cases"][name]
    sd = synth_conformer_state_dict(case["n_layers"], case["weight_seed"])
    src = synth_features(case["lengths"], 345, case["data_seed"])
    ts = synth_labels(case["lengths"], case["n_speakers"], case["data_seed"])
    g = load_golden(name)
    m = EendEdaModel(case["n_speakers"], 345, 4, 256, case["n_layers"], encoder_type="conformer", precision=prec)
    m.load_state_dict(sd, strict=True)
    m = m.to(dev).eval()
    o = OracleEendConformer(sd)
    torch.manual_seed(1000 + case["data_seed"])
    ys = m.infer(src, attractor_threshold=-1.0)
    torch.manual_seed(1000 + case["data_seed"])
    yo = o.infer(src, attractor_threshold=-1.0)
    emb_err = (m.last["emb"].cpu() - o.trace["emb"]).abs().max().item()
    assert emb_err < 2e-3, emb_err
    for b, (a, r) in enumerate(zip(ys, yo)):
        assert (a.cpu() - r).abs().max().item() < 1e-3                      # north_star's posterior bar
        assert np.abs(a.cpu().numpy() - g[f"post_all_{b}"]).max() < 1e-3    # and against the reference itself
    torch.manual_seed(2000 + case["data_seed"])
    loss, stats = m(src, ts)
    assert abs(loss.item() - float(g["loss"])) < 1e-3 and abs(stats["der"] - float(g["der"])) < 1e-3
    torch.manual_seed(2000 + case["data_seed"])
    o.forward(src, ts)
    assert m.last["min_idx"].cpu().tolist() == o.trace["min_idx"].tolist()
    assert torch.equal(m.last["label_permute"].cpu(), o.trace["label_permute"])
    m.train()
    with pytest.raises(NotImplementedError):
        m(src, ts)

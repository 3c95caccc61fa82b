"""Per-kernel parity of the training-step entry points (backward passes, Adam) against torch
autograd / closed forms evaluated on the CPU in fp32/fp64 on the same seeded inputs.

Tolerances: bf16x3 products carry ~2^-18 relative error, single bf16 ~2^-9 per operand; the
attention backward rounds P and dS to bf16 once in either mode (documented in mha_bwd_tc.cu),
so its tolerance is relative to the gradient's scale.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _prec(name):
    from speaker_diarization_b200 import _lib as L
    return L.PRECISIONS[name]


def _split(x, dev, ld=0):
    from speaker_diarization_b200 import ops
    return ops.split_bf16(x.to(dev).contiguous(), ld)


def rel_err(got, ref):
    return ((got - ref).abs().max() / ref.abs().max().clamp_min(1e-20)).item()


@pytest.mark.parametrize("prec", ["bf16x3", "bf16"])
@pytest.mark.parametrize("M,N_out,K_in", [(1000, 256, 256), (777, 2048, 256), (4100, 256, 2048), (640, 768, 256),
                                          (130, 256, 352), (64, 1024, 256), (9, 256, 256)])
def test_gemm_wgrad(dev, prec, M, N_out, K_in):
    """dW = dY^T X through the MN-major descriptors + split-K."""
    from speaker_diarization_b200 import ops_train as OT
    g = torch.Generator().manual_seed(M + N_out)
    dy = torch.randn((M, N_out), generator=g)
    x = torch.randn((M, K_in), generator=g)
    ref = (dy.double().t() @ x.double()).float()
    work = torch.empty(24 << 20, dtype=torch.uint8, device=dev)
    kc = 345 if K_in == 352 else K_in                       # the in-projection's ragged width
    dW = torch.full((N_out, kc), 7.0, device=dev)
    OT.gemm_wgrad(_split(dy, dev), _split(x, dev), M, N_out, kc, dW, work, _prec(prec))
    tol = {"bf16x3": 3e-5, "bf16": 2e-2}[prec]
    assert rel_err(dW.cpu(), ref[:, :kc]) < tol
    OT.gemm_wgrad(_split(dy, dev), _split(x, dev), M, N_out, kc, dW, work, _prec(prec), accumulate=True)
    assert rel_err(dW.cpu(), 2 * ref[:, :kc]) < tol


@pytest.mark.parametrize("prec", ["bf16x3", "bf16"])
@pytest.mark.parametrize("mode", ["plain", "resid", "mask"])
@pytest.mark.parametrize("M,K_in,N_out", [(515, 256, 2048), (300, 256, 768), (1000, 256, 256)])
def test_gemm_dgrad(dev, prec, mode, M, K_in, N_out):
    from speaker_diarization_b200 import ops, ops_train as OT
    if mode == "mask":
        K_in, N_out = N_out, K_in                           # linear1's ReLU: dh (M, 2048) = g (M, 256) . W2
    g = torch.Generator().manual_seed(M + K_in)
    dy = torch.randn((M, N_out), generator=g)
    w = (torch.rand((N_out, K_in), generator=g) * 2 - 1) / N_out ** 0.5
    res = torch.randn((M, K_in), generator=g)
    act = torch.randn((M, K_in), generator=g).relu()
    ref = (dy.double() @ w.double()).float()
    if mode == "resid":
        ref = ref + res
    if mode == "mask":
        ref = ref * (act > 0)
    out = torch.empty((M, K_in), device=dev)
    osp = ops.Split.empty(M, K_in, dev)
    OT.gemm_dgrad(_split(dy, dev), _split(w.t().contiguous(), dev), M, K_in, N_out, _prec(prec),
                  resid=res.to(dev) if mode == "resid" else None,
                  relu_mask=_split(act, dev) if mode == "mask" else None, out=out, out_split=osp)
    tol = {"bf16x3": 3e-5, "bf16": 2e-2}[prec]
    assert rel_err(out.cpu(), ref) < tol
    assert rel_err(osp.float().cpu(), ref) < tol + 1e-5


@pytest.mark.parametrize("M", [1, 77, 4000])
def test_ln_bwd_and_forward_rstd(dev, M):
    from speaker_diarization_b200 import _lib as L, ops, ops_train as OT
    g = torch.Generator().manual_seed(M)
    K = 256
    a = torch.randn((M, K), generator=g)
    w = (torch.rand((256, K), generator=g) * 2 - 1) / 16
    b = torch.randn((256,), generator=g) * 0.1
    r = torch.randn((M, 256), generator=g)
    gm = (1 + 0.1 * torch.randn((256,), generator=g)).requires_grad_(True)
    bt = (0.1 * torch.randn((256,), generator=g)).requires_grad_(True)
    z = (F.linear(a, w, b) + r).requires_grad_(True)
    y_ref = F.layer_norm(z, (256,), gm, bt, 1e-5)
    dy = torch.randn((M, 256), generator=g)
    y_ref.backward(dy)
    rstd = torch.empty((M,), device=dev)
    y = torch.empty((M, 256), device=dev)
    OT.gemm_split_train(_split(a, dev), _split(w, dev), b.to(dev), M, 256, K, L.EPI_BIAS_RES_LN, L.PREC_BF16X3, rstd,
                        resid=r.to(dev), gamma=gm.detach().to(dev), beta=bt.detach().to(dev), out=y)
    assert (y.cpu() - y_ref.detach()).abs().max().item() < 1e-4
    rstd_ref = 1.0 / torch.sqrt(z.detach().var(dim=1, unbiased=False) + 1e-5)
    assert rel_err(rstd.cpu(), rstd_ref) < 1e-4
    work = torch.empty(4 << 20, dtype=torch.uint8, device=dev)
    dx = torch.empty((M, 256), device=dev)
    dxs = ops.Split.empty(M, 256, dev)
    dg, db, cs = (torch.empty((256,), device=dev) for _ in range(3))
    OT.ln_bwd(dy.to(dev), y, gm.detach().to(dev), bt.detach().to(dev), rstd, work, dg, db, cs, dx=dx, dx_split=dxs)
    assert rel_err(dx.cpu(), z.grad) < 2e-4
    assert rel_err(dxs.float().cpu(), z.grad) < 2e-4
    assert rel_err(dg.cpu(), gm.grad) < 2e-4 and rel_err(db.cpu(), bt.grad) < 1e-5
    assert rel_err(cs.cpu(), z.grad.sum(0)) < 5e-4 or z.grad.sum(0).abs().max() < 1e-4


def test_colsum_and_scale_split(dev):
    from speaker_diarization_b200 import ops, ops_train as OT
    g = torch.Generator().manual_seed(5)
    x = torch.randn((3001, 2048), generator=g)
    out = torch.empty((2048,), device=dev)
    work = torch.empty(8 << 20, dtype=torch.uint8, device=dev)
    OT.colsum_split(_split(x, dev), 3001, 2048, out, work)
    assert rel_err(out.cpu(), x.double().sum(0).float()) < 1e-5
    out2 = torch.empty((768,), device=dev)
    OT.colsum_split(_split(x[:, :768].contiguous(), dev), 3001, 768, out2, work)
    assert rel_err(out2.cpu(), x[:, :768].double().sum(0).float()) < 1e-5
    dst = ops.Split.empty(3001, 768, dev)
    OT.scale_split(x[:, :256].contiguous().to(dev), 0.125, dst, cols=256)
    assert rel_err(dst.float()[:, :256].cpu(), x[:, :256] * 0.125) < 1e-5


@pytest.mark.parametrize("B,T,S", [(3, 130, 2), (2, 500, 4), (5, 64, 8)])
def test_posterior_and_head_bwd(dev, B, T, S):
    from speaker_diarization_b200 import ops_train as OT
    g = torch.Generator().manual_seed(B * T)
    emb = torch.randn((B, T, 256), generator=g, requires_grad=True)
    att = torch.randn((B, S + 1, 256), generator=g, requires_grad=True)
    w = (torch.randn((256,), generator=g) / 16).requires_grad_(True)
    bias = torch.randn((1,), generator=g).requires_grad_(True)
    dpred = torch.randn((B, T, S), generator=g) * 0.01
    pred = emb @ att[:, :-1].transpose(1, 2)
    probs = torch.sigmoid(att @ w + bias)
    tgt = torch.zeros((B, S + 1)); tgt[:, :S] = 1
    l_att = F.binary_cross_entropy_with_logits(probs, tgt, reduction="none").mean(1).mean() * 0.7
    (l_att + (pred * dpred).sum()).backward()
    demb = torch.empty((B * T, 256), device=dev)
    datt = torch.zeros((B, S + 1, 256), device=dev)
    OT.posterior_bwd(dpred.to(dev), emb.detach().reshape(B * T, 256).to(dev), att.detach().to(dev), B, T, S, demb, datt)
    dw, dbias = torch.empty((256,), device=dev), torch.empty((1,), device=dev)
    OT.attractor_head_bwd(att.detach().to(dev), probs.detach().to(dev), w.detach().to(dev), S, 0.7, datt, dw, dbias)
    assert rel_err(demb.cpu().view(B, T, 256), emb.grad) < 1e-5
    assert rel_err(datt.cpu(), att.grad) < 2e-5
    assert rel_err(dw.cpu(), w.grad) < 2e-5 and rel_err(dbias.cpu(), bias.grad) < 2e-5


def _attention_ref(x, w, b, B, T, H):
    qkv = F.linear(x, w, b).view(B, T, 3, H, 64)
    q, k, v = (qkv[:, :, i].transpose(1, 2) for i in range(3))
    a = torch.softmax(q @ k.transpose(-1, -2) / 8.0, dim=-1) @ v
    return a.transpose(1, 2).reshape(B * T, H * 64)


@pytest.mark.parametrize("prec", ["bf16x3", "bf16"])
@pytest.mark.parametrize("B,T", [(2, 128), (3, 200), (1, 500), (2, 37)])
def test_mha_bwd(dev, prec, B, T):
    """QKV GEMM -> attention (with lse) -> delta -> attention backward, against autograd."""
    from speaker_diarization_b200 import ops, ops_train as OT
    H, D = 4, 256
    M, Tp = B * T, (T + 7) // 8 * 8
    g = torch.Generator().manual_seed(B * 1000 + T)
    x = torch.randn((M, D), generator=g)
    w = (torch.rand((3 * D, D), generator=g) * 2 - 1) * (6.0 / (4 * D)) ** 0.5 * 2
    b = torch.randn((3 * D,), generator=g) * 0.1
    dO = torch.randn((M, D), generator=g)
    qkv_leaf = F.linear(x.view(B, T, D), w, b).detach().requires_grad_(True)
    q, k, v = (qkv_leaf.view(B, T, 3, H, 64)[:, :, i].transpose(1, 2) for i in range(3))
    a = (torch.softmax(q @ k.transpose(-1, -2) / 8.0, dim=-1) @ v).transpose(1, 2).reshape(M, D)
    a.backward(dO)
    dqkv_ref = qkv_leaf.grad.view(M, 3 * D)
    P = _prec(prec)
    qk, vt = ops.Split.empty(M, 2 * D, dev), ops.Split.empty(B * H * 64, Tp, dev)
    vt.hi.zero_(); vt.lo.zero_()
    ops.gemm_qkv(_split(x, dev), _split(w, dev), b.to(dev), qk, vt, B, T, Tp, D, H, P)
    ctx = ops.Split.empty(M, D, dev)
    lse = torch.empty((B * H, T), device=dev)
    OT.mha_split_train(qk, vt, B, T, Tp, H, P, ctx, lse)
    tol_f = {"bf16x3": 2e-5, "bf16": 3e-2}[prec]
    assert rel_err(ctx.float().cpu(), a.detach()) < tol_f
    s_ref = (q @ k.transpose(-1, -2) / 8.0).detach()
    lse_ref = torch.logsumexp(s_ref, dim=-1).reshape(B * H, T) / np.log(2.0)
    assert (lse.cpu() - lse_ref).abs().max().item() < (1e-3 if prec == "bf16x3" else 5e-2)
    do = _split(dO, dev)
    delta = torch.empty((M, H), device=dev)
    OT.mha_delta(do, ctx, M, D, delta)
    delta_ref = (dO * a.detach()).view(M, H, 64).sum(-1)
    assert rel_err(delta.cpu(), delta_ref) < tol_f + 1e-5
    dq_acc = torch.empty((M, D), device=dev)
    dqkv = ops.Split.empty(M, 3 * D, dev)
    OT.mha_bwd(qk, vt, Tp, do, lse, delta, B, T, H, dq_acc, dqkv, P)
    OT.scale_split(dq_acc, 0.125, dqkv, cols=D)
    got = dqkv.float().cpu()
    tol = {"bf16x3": 6e-3, "bf16": 4e-2}[prec]               # P, dS are rounded to bf16 once in both modes
    for name, sl in (("dq", slice(0, D)), ("dk", slice(D, 2 * D)), ("dv", slice(2 * D, 3 * D))):
        e = ((got[:, sl] - dqkv_ref[:, sl]).norm() / dqkv_ref[:, sl].norm()).item()
        assert e < tol, (name, e)


@pytest.mark.parametrize("B,T", [(2, 200), (1, 333)])
def test_mha_dropout_bits_equal_inline_masks_and_autograd(dev, B, T):
    """Attention with dropout on the weights: the kernels fed precomputed keep bits (eda_dropout_bits) and the
    kernels evaluating the Philox function themselves give the same forward output and the same dk, dv bit for
    bit (dq is summed with atomics: compared to rounding); both match autograd with the exported mask."""
    from speaker_diarization_b200 import _lib as L, ops, ops_train as OT
    H, D, p, seed, site = 4, 256, 0.1, 4242, 9
    M, Tp = B * T, (T + 7) // 8 * 8
    g = torch.Generator().manual_seed(B * 77 + T)
    x = torch.randn((M, D), generator=g)
    w = (torch.rand((3 * D, D), generator=g) * 2 - 1) * (6.0 / (4 * D)) ** 0.5 * 2
    b = torch.randn((3 * D,), generator=g) * 0.1
    dO = torch.randn((M, D), generator=g)
    P = L.PREC_BF16X3
    qk, vt = ops.Split.empty(M, 2 * D, dev), ops.Split.empty(B * H * 64, Tp, dev)
    vt.hi.zero_(); vt.lo.zero_()
    ops.gemm_qkv(_split(x, dev), _split(w, dev), b.to(dev), qk, vt, B, T, Tp, D, H, P)
    bits = torch.empty((B * H * T, (T + 31) // 32 + 1), dtype=torch.int32, device=dev)
    OT.dropout_bits(p, seed, site, B * H * T, T, bits)
    do = _split(dO, dev)
    outs = []
    for db in (None, bits):
        ctx, lse = ops.Split.empty(M, D, dev), torch.empty((B * H, T), device=dev)
        OT.mha_split_train(qk, vt, B, T, Tp, H, P, ctx, lse, drop=(p, seed, site), drop_bits=db)
        delta = torch.empty((M, H), device=dev)
        OT.mha_delta(do, ctx, M, D, delta)
        dq_acc, dqkv = torch.empty((M, D), device=dev), ops.Split.empty(M, 3 * D, dev)
        OT.mha_bwd(qk, vt, Tp, do, lse, delta, B, T, H, dq_acc, dqkv, P, drop=(p, seed, site), drop_bits=db)
        OT.scale_split(dq_acc, 0.125, dqkv, cols=D)
        outs.append((ctx.float().cpu(), dqkv.float().cpu()))
    assert torch.equal(outs[0][0], outs[1][0])
    assert torch.equal(outs[0][1][:, D:], outs[1][1][:, D:])
    assert rel_err(outs[0][1][:, :D], outs[1][1][:, :D]) < 1e-5
    # autograd with the exported mask
    keep = OT.dropout_mask(p, seed, site, B * H * T, (T + 7) // 8 * 8, dev)[:, :T].view(B, H, T, T).cpu().float()
    qkv_leaf = F.linear(x.view(B, T, D), w, b).detach().requires_grad_(True)
    q, k, v = (qkv_leaf.view(B, T, 3, H, 64)[:, :, i].transpose(1, 2) for i in range(3))
    a = ((torch.softmax(q @ k.transpose(-1, -2) / 8.0, dim=-1) * keep / (1 - p)) @ v).transpose(1, 2).reshape(M, D)
    a.backward(dO)
    e = rel_err(outs[1][0], a.detach())
    assert e < 4e-5, e            # 2.2e-5 measured: fewer, 1/(1-p) larger terms per row than the dropout-free case
    ref = qkv_leaf.grad.view(M, 3 * D)
    for sl in (slice(0, D), slice(D, 2 * D), slice(2 * D, 3 * D)):
        assert ((outs[1][1][:, sl] - ref[:, sl]).norm() / ref[:, sl].norm()).item() < 6e-3


def _lstm_ref(emb, lens, perms, n_dec, We, Ue, be, Ud, bd):
    B = emb.shape[0]
    hs, cs = [], []
    for bi in range(B):
        h = emb.new_zeros((1, 256)); c = emb.new_zeros((1, 256))
        xg = F.linear(emb[bi, perms[bi]], We, be)
        for t in range(lens[bi]):
            gg = xg[t:t + 1] + h @ Ue.t()
            i, f, g_, o = gg.chunk(4, -1)
            c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(g_)
            h = torch.sigmoid(o) * torch.tanh(c)
        hs.append(h); cs.append(c)
    h, c = torch.cat(hs), torch.cat(cs)
    atts = []
    for _ in range(n_dec):
        gg = bd.unsqueeze(0) + h @ Ud.t()
        i, f, g_, o = gg.chunk(4, -1)
        c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(g_)
        h = torch.sigmoid(o) * torch.tanh(c)
        atts.append(h)
    return torch.stack(atts, 1)


@pytest.mark.parametrize("lens,n_dec", [([40, 17, 40, 1, 33], 3), ([64] * 9, 4), ([5], 1), ([30, 0, 12], 2)])
@pytest.mark.parametrize("lstm_prec", ["fp32", "bf16x3"])     # forward recurrence on CUDA cores / tensor cores
def test_lstm_bwd(dev, lens, n_dec, lstm_prec):
    """Forward with saves + BPTT kernel + the GEMMs that turn gate gradients into parameter / input
    gradients, against autograd through a python-loop LSTM."""
    from speaker_diarization_b200 import _lib as L, ops, ops_train as OT
    B, Tmax, D = len(lens), max(max(lens), 1), 256
    M = B * Tmax
    g = torch.Generator().manual_seed(sum(lens) + n_dec)
    k = 1 / 16
    mk = lambda *s: ((torch.rand(s, generator=g) * 2 - 1) * k).requires_grad_(True)
    We, Ue, be, Ud, bd = mk(4 * D, D), mk(4 * D, D), mk(4 * D), mk(4 * D, D), mk(4 * D)
    emb = torch.randn((B, Tmax, D), generator=g).requires_grad_(True)
    perms = [torch.randperm(n, generator=g) for n in lens]
    datt = torch.randn((B, n_dec, D), generator=g)
    att_ref = _lstm_ref(emb, lens, perms, n_dec, We, Ue, be, Ud, bd)
    att_ref.backward(datt)
    perm = torch.arange(Tmax, dtype=torch.int32).repeat(B, 1)
    for i, p in enumerate(perms):
        perm[i, :lens[i]] = p.to(torch.int32)
    perm_d, lens_d = perm.to(dev), torch.tensor(lens, dtype=torch.int32, device=dev)
    P = L.PREC_BF16X3
    es = _split(emb.detach().view(M, D), dev)
    xproj = torch.empty((M, 4 * D), device=dev)
    ops.gemm_split(es, _split(We.detach(), dev), be.detach().to(dev), M, 4 * D, D, L.EPI_BIAS, P, out=xproj)
    sv = dict(gates=torch.zeros((M, 4 * D), device=dev), c=torch.zeros((M, D), device=dev),
              hprev=ops.Split(torch.zeros((M, D), dtype=torch.bfloat16, device=dev), torch.zeros((M, D), dtype=torch.bfloat16, device=dev)),
              dec_gates=torch.zeros((B * n_dec, 4 * D), device=dev), dec_c=torch.zeros((B * n_dec, D), device=dev),
              dec_hprev=ops.Split.empty(B * n_dec, D, dev), hc=torch.zeros((B, 2, D), device=dev))
    att = OT.lstm_attractors_train(xproj, perm_d, lens_d, None, Ue.detach().to(dev), Ud.detach().to(dev),
                                   bd.detach().to(dev), B, Tmax, n_dec, sv, L.PRECISIONS[lstm_prec])
    assert (att.cpu() - att_ref.detach()).abs().max().item() < (2e-5 if lstm_prec == "fp32" else 1e-4)
    dxp, ddec = ops.Split.empty(M, 4 * D, dev), ops.Split.empty(B * n_dec, 4 * D, dev)
    OT.lstm_attractors_bwd(perm_d, lens_d, None, Ue.detach().to(dev), Ud.detach().to(dev), sv, datt.to(dev), B, Tmax,
                           n_dec, dxp, ddec)
    work = torch.empty(24 << 20, dtype=torch.uint8, device=dev)
    dUe, dWe, dUd = (torch.empty((4 * D, D), device=dev) for _ in range(3))
    OT.gemm_wgrad(dxp, sv["hprev"], M, 4 * D, D, dUe, work, P)
    OT.gemm_wgrad(dxp, es, M, 4 * D, D, dWe, work, P)
    OT.gemm_wgrad(ddec, sv["dec_hprev"], B * n_dec, 4 * D, D, dUd, work, P)
    dbe, dbd = torch.empty((4 * D,), device=dev), torch.empty((4 * D,), device=dev)
    OT.colsum_split(dxp, M, 4 * D, dbe, work)
    OT.colsum_split(ddec, B * n_dec, 4 * D, dbd, work)
    demb = torch.empty((M, D), device=dev)
    OT.gemm_dgrad(dxp, _split(We.detach().t().contiguous(), dev), M, D, 4 * D, P, out=demb)
    tol = 3e-4
    assert rel_err(dbd.cpu(), bd.grad) < tol
    assert rel_err(dUd.cpu(), Ud.grad) < tol
    if sum(lens) > 0:
        assert rel_err(dbe.cpu(), be.grad) < tol
        assert rel_err(dUe.cpu(), Ue.grad) < tol
        assert rel_err(dWe.cpu(), We.grad) < tol
        assert rel_err(demb.cpu().view(B, Tmax, D), emb.grad) < tol


def test_adam_and_scale(dev):
    from speaker_diarization_b200 import ops_train as OT
    g = torch.Generator().manual_seed(3)
    p0 = torch.randn((100003,), generator=g)
    p_ref = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([p_ref], lr=1e-3, betas=(0.9, 0.98), eps=1e-9)
    p, m, v = p0.to(dev), torch.zeros_like(p0, device=dev), torch.zeros_like(p0, device=dev)
    for step in range(1, 4):
        grad = torch.randn((100003,), generator=g)
        p_ref.grad = grad.clone()
        opt.step()
        gd = grad.to(dev) * 4.0
        OT.scale_dev(torch.tensor([0.25], device=dev), gd)
        OT.adam_step(p, gd, m, v, 1e-3, 0.9, 0.98, 1e-9, step)
    assert (p.cpu() - p_ref.detach()).abs().max().item() < 2e-6

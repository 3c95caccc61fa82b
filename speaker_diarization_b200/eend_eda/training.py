"""Training step of ``TransformerEdaModel`` on the hand-written kernels.

What the reference does per step (train_eda.py:182-192)::

    loss, stats = model(y, t); loss.backward(); optimizer.step(); optimizer.zero_grad(); scheduler.step()

with ``optimizer = Adam(betas=(0.9, 0.98), eps=1e-9)`` and ``NoamScheduler`` (lr_scheduler.py:6-33)
for ``--optimizer noam``.  Here:

* ``forward_train`` is the tensor-core forward with the few extra quantities the backward needs
  (per-layer activations as split-bf16 pairs, LayerNorm rstd, attention log-sum-exp, LSTM gates);
* ``backward_train`` produces every parameter gradient in ONE flat fp32 buffer through the
  backward kernels (input / weight gradient GEMMs on the same tcgen05 kernel as the forward,
  attention backward, LSTM BPTT cluster kernel, LayerNorm / bias / posterior / head kernels);
* ``EdaLossFn`` hangs both on torch.autograd so ``loss.backward()`` fills ``p.grad`` as usual;
* ``NoamAdam`` keeps parameters, gradients and Adam moments in flat buffers: one fused Adam kernel
  per step and, under ``torch.distributed``, one NCCL all-reduce of the whole gradient.

Dropout (``nn.TransformerEncoderLayer`` in train() mode, p = the constructor's ``dropout``): applied at
the layer's four sites (attention weights, after out_proj, after the ReLU, after linear2) by a counter-based
Philox function of (seed, site, row, column) that the forward epilogues and the backward kernels both
evaluate - no mask is stored.  The step's seed derives from ``torch.initial_seed()`` and a step counter
(``torch.manual_seed`` makes a run repeatable, the global generator is not consumed); the masks are NOT torch's own Philox stream, so with
dropout > 0 a step matches the reference in distribution, and exactly when the oracle is fed these masks
(``dropout_sites`` / ``ops_train.dropout_mask``).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np
import torch

from .. import _lib as L
from .. import ops
from .. import ops_train as OT
from ..sharding import allreduce_mean_

Tensor = torch.Tensor


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


# ----------------------------------------------------------------------------- weights --------
def prepared_train(model) -> dict:
    """Transposed split-bf16 copies of the weight matrices (operand W^T of the input-gradient
    GEMMs).  Cached next to the forward copies and rebuilt when a parameter changes."""
    w = model._prepared()
    if "wt_ih" not in w:
        t = lambda x: ops.split_bf16(x.t().contiguous())
        for lw in w["layers"]:
            lw["wt_qkv"], lw["wt_o"] = t(lw["w_qkv"]), t(lw["w_o"])
            lw["wt_1"], lw["wt_2"] = t(lw["w_1"]), t(lw["w_2"])
        e = model.eda._prepared()
        if "sw_ih" not in e:
            e["sw_ih"] = ops.split_bf16(e["w_ih"])
        w["wt_ih"] = t(e["w_ih"])
    return w


def param_layout(model):
    """[(name, parameter, offset, numel)] of the flat gradient buffer, offsets 16-byte aligned."""
    out, off = [], 0
    for name, p in model.named_parameters():
        out.append((name, p, off, p.numel()))
        off += _round_up(p.numel(), 4)
    return out, off


# --------------------------------------------------------------------------- workspace --------
def train_workspace(model, B: int, Tmax: int, n_dec: int, dev) -> dict:
    key = ("train", B, Tmax, n_dec, str(dev))
    ws = model._ws.get(key)
    if ws is not None:
        return ws
    # a training workspace is ~5 GB at 32 x 2000 frames: keep at most two shapes (a steady batch shape + the
    # ragged last batch of an epoch), oldest first out, independent of the inference arena
    old = [k for k in model._ws if k[0] == "train"]
    for k in old[: max(0, len(old) - 1)]:
        del model._ws[k]
    M, D, H, FF = B * Tmax, model.n_units, model.n_heads, model.dim_feedforward
    Tp = _round_up(Tmax, 8)
    f32 = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
    nl = model.n_layers
    layers = [dict(qk=ops.Split.empty(M, 2 * D, dev), vt=ops.Split.empty(B * H * (D // H), Tp, dev),
                   ctx=ops.Split.empty(M, D, dev), lse=f32(B * H, Tmax), y1_f=f32(M, D), y1_s=ops.Split.empty(M, D, dev),
                   ff=ops.Split.empty(M, FF, dev), rstd1=f32(M), rstd2=f32(M)) for _ in range(nl)]
    for lw in layers:                      # key columns >= T of v^T are never written by the QKV epilogue
        lw["vt"].hi.zero_(); lw["vt"].lo.zero_()
    zs = lambda r, c: ops.Split(torch.zeros((r, c), dtype=torch.bfloat16, device=dev),
                                torch.zeros((r, c), dtype=torch.bfloat16, device=dev))
    ws = dict(
        Tp=Tp, layers=layers, xin=ops.Split.empty(M, _round_up(model.in_size, 16), dev),
        x_f=[f32(M, D) for _ in range(nl + 1)], x_s=[ops.Split.empty(M, D, dev) for _ in range(nl + 1)], rstd0=f32(M),
        xproj=f32(M, 4 * D),
        # rows of padded frames are never written: they must stay finite (they meet zero gradients)
        sv=dict(gates=f32(M, 4 * D), c=f32(M, D), hprev=zs(M, D), dec_gates=f32(B * n_dec, 4 * D),
                dec_c=f32(B * n_dec, D), dec_hprev=zs(B * n_dec, D), hc=f32(B, 2, D)),
        # backward temporaries
        dy=[f32(M, D) for _ in range(2)], g_f=f32(M, D), g_s=ops.Split.empty(M, D, dev),
        dh=ops.Split.empty(M, FF, dev), dctx=ops.Split.empty(M, D, dev), dqkv=ops.Split.empty(M, 3 * D, dev),
        dq_acc=f32(M, D), delta=f32(M, H), dxproj=ops.Split.empty(M, 4 * D, dev),
        ddec=ops.Split.empty(B * n_dec, 4 * D, dev), demb=f32(M, D),
        work=torch.empty(32 << 20, dtype=torch.uint8, device=dev),
    )
    model._ws[key] = ws
    return ws


# ----------------------------------------------------------------------------- dropout --------
SITE_ATTN, SITE_OUT, SITE_FFN, SITE_FFN2 = 0, 1, 2, 3


def site_id(layer: int, which: int) -> int:
    return layer * 4 + which


def _step_seed(model, step: int) -> int:
    return (torch.initial_seed() * 0x9E3779B97F4A7C15 + step * 0xD1B54A32D192ED03) & (2 ** 63 - 1)


def _fill_keep_bits(model, bufs, pd: float, seed: int, B: int, H: int, Tmax: int):
    for i, buf in enumerate(bufs):
        OT.dropout_bits(pd, seed, site_id(i, SITE_ATTN), B * H * Tmax, Tmax, buf)


def _attention_keep_bits(model, ws: dict, pd: float, seed: int, B: int, H: int, Tmax: int, dev):
    """The attention-weight keep masks of this step as bits (1 bit per weight, 16 MB per layer at 32 x 2000 frames), one
    buffer per layer: evaluated once instead of in the softmax warps of both attention kernels (the other three sites
    stay recomputed in their epilogues).  Two sets alternate: the previous step evaluated this step's set on a side
    stream beside its LSTM recurrence (``_prefetch_keep_bits``); if it guessed the seed and shape right, only wait."""
    sets = ws.setdefault("bits", [[torch.empty((B * H * Tmax, _round_up(Tmax, 128) // 32), dtype=torch.int32, device=dev)
                                   for _ in range(model.n_layers)] for _ in range(2)])
    which = model._drop_step & 1
    ready = ws.get("bits_ready")
    if ready is not None and ready[0] == (seed, which):
        torch.cuda.current_stream(dev).wait_event(ready[1])
    else:
        _fill_keep_bits(model, sets[which], pd, seed, B, H, Tmax)
    ws["bits_ready"] = None
    return sets[which]


def _prefetch_keep_bits(model, ws: dict, pd: float, B: int, H: int, Tmax: int, dev, fork):
    nxt = model._drop_step + 1
    seed, which = _step_seed(model, nxt), nxt & 1
    side = ws.setdefault("side_stream", torch.cuda.Stream(device=dev))
    side.wait_event(fork)              # the other set was last read by the previous step's backward: long finished
    with torch.cuda.stream(side), L.use_stream(side):    # explicit: a caller's use_stream() pin must not redirect these
        _fill_keep_bits(model, ws["bits"][which], pd, seed, B, H, Tmax)
        done = torch.cuda.Event()
        done.record(side)
    ws["bits_ready"] = ((seed, which), done)


def dropout_sites(model, seed: int, B: int, Tmax: int, layer: int, device):
    """The four keep masks of one layer as torch.nn.TransformerEncoderLayer would apply them (bool tensors):
    attention weights (B, H, T, T), dropout1 (B, T, D), FFN dropout (B, T, FF), dropout2 (B, T, D)."""
    p, H, D, FF = model.dropout_p, model.n_heads, model.n_units, model.dim_feedforward
    Tc = _round_up(Tmax, 8)
    att = OT.dropout_mask(p, seed, site_id(layer, SITE_ATTN), B * H * Tmax, Tc, device)[:, :Tmax].view(B, H, Tmax, Tmax)
    d1 = OT.dropout_mask(p, seed, site_id(layer, SITE_OUT), B * Tmax, D, device).view(B, Tmax, D)
    ff = OT.dropout_mask(p, seed, site_id(layer, SITE_FFN), B * Tmax, FF, device).view(B, Tmax, FF)
    d2 = OT.dropout_mask(p, seed, site_id(layer, SITE_FFN2), B * Tmax, D, device).view(B, Tmax, D)
    return att.bool(), d1.bool(), ff.bool(), d2.bool()


# ----------------------------------------------------------------------------- forward --------
def forward_train(model, src: List[Tensor], ts: List[Tensor]):
    """models.py:236-295 with everything the backward pass needs kept in the workspace.
    Returns (loss (0-d), stats dict, saved dict)."""
    dev = model._device()
    P = model.precision
    if P == L.PREC_FP32:
        raise NotImplementedError("training runs on the tensor-core pipeline: use precision 'bf16x3' or 'bf16'")
    pd = float(model.dropout_p)
    # one seed per step from torch's seed and a step counter: repeatable under torch.manual_seed without consuming
    # the global generator (the frame shuffle must draw the same numbers as the reference's, models.py:229-232)
    model._drop_step = getattr(model, "_drop_step", 0) + 1
    seed = _step_seed(model, model._drop_step) if pd > 0 else 0
    model._last_drop_seed = seed
    drop = lambda layer, which: (pd, seed, site_id(layer, which)) if pd > 0 else OT.NO_DROP
    w = prepared_train(model)
    ilens = [int(x.shape[0]) for x in src]
    B, Tmax = len(ilens), max(ilens)
    S = int(ts[0].shape[1])
    n_dec = S + 1
    D, H, FF = model.n_units, model.n_heads, model.dim_feedforward
    M = B * Tmax
    ws = train_workspace(model, B, Tmax, n_dec, dev)
    Tp = ws["Tp"]
    off = np.zeros(B + 1, dtype=np.int64)
    np.cumsum(ilens, out=off[1:])
    meta = torch.from_numpy(np.concatenate([off, np.asarray(ilens, dtype=np.int64)])).to(dev, non_blocking=True)
    seq_off, lens_dev = meta[: B + 1], meta[B + 1:].to(torch.int32)
    flat = model._to_device_flat(src, dev)
    ops.pad_pack_split(flat, seq_off, lens_dev, B, Tmax, ws["xin"], -1.0)
    x_f, x_s = ws["x_f"], ws["x_s"]
    OT.gemm_split_train(ws["xin"], w["ws_in"], w["b_in"], M, D, ws["xin"].ld, L.EPI_BIAS_LN, P, ws["rstd0"],
                        gamma=w["g_in"], beta=w["be_in"], out=x_f[0], out_split=x_s[0])
    bits = _attention_keep_bits(model, ws, pd, seed, B, H, Tmax, dev) if pd > 0 else None
    for i, (lw, a) in enumerate(zip(w["layers"], ws["layers"])):
        ops.gemm_qkv(x_s[i], lw["sw_qkv"], lw["b_qkv"], a["qk"], a["vt"], B, Tmax, Tp, D, H, P)
        OT.mha_split_train(a["qk"], a["vt"], B, Tmax, Tp, H, model.attn_precision, a["ctx"], a["lse"],
                           drop=drop(i, SITE_ATTN), drop_bits=bits[i] if pd > 0 else None)
        OT.gemm_split_train(a["ctx"], lw["sw_o"], lw["b_o"], M, D, D, L.EPI_BIAS_RES_LN, P, a["rstd1"], resid=x_f[i],
                            gamma=lw["g1"], beta=lw["be1"], out=a["y1_f"], out_split=a["y1_s"], drop=drop(i, SITE_OUT))
        OT.gemm_split_train(a["y1_s"], lw["sw_1"], lw["b_1"], M, FF, D, L.EPI_BIAS_RELU, P, None, out_split=a["ff"],
                            drop=drop(i, SITE_FFN))
        OT.gemm_split_train(a["ff"], lw["sw_2"], lw["b_2"], M, D, FF, L.EPI_BIAS_RES_LN, P, a["rstd2"], resid=a["y1_f"],
                            gamma=lw["g2"], beta=lw["be2"], out=x_f[i + 1], out_split=x_s[i + 1], drop=drop(i, SITE_FFN2))
    emb, emb_s = x_f[-1], x_s[-1]
    label = ops.pad_pack(model._to_device_flat(list(ts), dev), seq_off, lens_dev, B, Tmax, S, -1.0).view(B, Tmax, S)
    perm = model._draw_perm(ilens, Tmax, dev)
    e = model.eda._prepared()
    ops.gemm_split(emb_s, e["sw_ih"], e["b_enc"], M, 4 * D, D, L.EPI_BIAS, P, out=ws["xproj"])
    order = torch.tensor(sorted(range(B), key=lambda b: -ilens[b]), dtype=torch.int32).to(dev, non_blocking=True)
    fork = torch.cuda.Event() if pd > 0 else None
    if pd > 0:
        fork.record()
    att = OT.lstm_attractors_train(ws["xproj"], perm, lens_dev, order, e["u_enc"], e["u_dec"], e["b_dec"], B, Tmax,
                                   n_dec, ws["sv"], P)
    if pd > 0:          # the recurrence leaves 84 SMs idle for ~2.5 ms: the NEXT step's masks are evaluated beside it
        _prefetch_keep_bits(model, ws, pd, B, H, Tmax, dev, fork)
    probs, loss_att = ops.attractor_head(att, e["w_lin"], e["b_lin"], S)
    pred, _, _ = ops.posterior(emb, att, B, Tmax, S, want_logits=True, want_post=False)
    loss_pit, min_idx, label_perm, _ = ops.pit_loss(pred, label, lens_dev)
    counts = ops.der_counters(pred, label_perm, lens_dev)
    loss = (model.diar_weight * loss_pit + model.attractor_weight * loss_att).reshape(())
    saved = dict(ws=ws, B=B, Tmax=Tmax, S=S, n_dec=n_dec, lens_dev=lens_dev, perm=perm, order=order, att=att,
                 probs=probs, pred=pred, label_perm=label_perm, sum_len=float(sum(ilens)), p_drop=pd, seed=seed,
                 bits=bits)
    model.last = dict(emb=emb.view(B, Tmax, -1), attractors=att, probs=probs, pred=pred, min_idx=min_idx,
                      label_permute=label_perm, counts=counts)
    return loss, loss_att, loss_pit, counts, saved


# ---------------------------------------------------------------------------- backward --------
def backward_train(model, saved: dict) -> Tensor:
    """d loss / d parameters for the step recorded in ``saved``: one flat fp32 buffer laid out by
    ``param_layout`` (upstream gradient 1; the caller scales)."""
    P = model.precision
    w = prepared_train(model)
    e = model.eda._prepared()
    ws, B, Tmax, S, n_dec = saved["ws"], saved["B"], saved["Tmax"], saved["S"], saved["n_dec"]
    D, H, FF = model.n_units, model.n_heads, model.dim_feedforward
    M = B * Tmax
    dev = saved["pred"].device
    layout, total = param_layout(model)
    gflat = torch.zeros(total, dtype=torch.float32, device=dev)
    G: Dict[str, Tensor] = {name: gflat[off: off + n].view(p.shape) for name, p, off, n in layout}
    work, sv = ws["work"], ws["sv"]
    lens_dev = saved["lens_dev"]
    emb, emb_s = ws["x_f"][-1], ws["x_s"][-1]

    # loss -> logits -> (emb, attractors)
    dpred = ops.pit_loss_bwd(saved["pred"], saved["label_perm"], lens_dev, model.diar_weight)
    datt = torch.zeros((B, n_dec, D), dtype=torch.float32, device=dev)
    OT.posterior_bwd(dpred, emb, saved["att"], B, Tmax, S, ws["demb"], datt)
    OT.attractor_head_bwd(saved["att"], saved["probs"], e["w_lin"], S, model.attractor_weight, datt,
                          G["eda.linear.weight"].view(-1), G["eda.linear.bias"])
    # EDA: BPTT, then the gate gradients become weight / bias / input gradients
    OT.lstm_attractors_bwd(saved["perm"], lens_dev, saved["order"], e["u_enc"], e["u_dec"], sv, datt, B, Tmax, n_dec,
                           ws["dxproj"], ws["ddec"])
    OT.gemm_wgrad(ws["dxproj"], sv["hprev"], M, 4 * D, D, G["eda.encoder.weight_hh_l0"], work, P)
    OT.gemm_wgrad(ws["dxproj"], emb_s, M, 4 * D, D, G["eda.encoder.weight_ih_l0"], work, P)
    OT.gemm_wgrad(ws["ddec"], sv["dec_hprev"], B * n_dec, 4 * D, D, G["eda.decoder.weight_hh_l0"], work, P)
    OT.colsum_split(ws["dxproj"], M, 4 * D, G["eda.encoder.bias_ih_l0"], work)
    OT.colsum_split(ws["ddec"], B * n_dec, 4 * D, G["eda.decoder.bias_ih_l0"], work)
    G["eda.encoder.bias_hh_l0"].copy_(G["eda.encoder.bias_ih_l0"])      # b_ih and b_hh enter as a sum
    G["eda.decoder.bias_hh_l0"].copy_(G["eda.decoder.bias_ih_l0"])      # (decoder.weight_ih sees zero input: grad 0)
    dy, dy_other = ws["dy"]
    OT.gemm_dgrad(ws["dxproj"], w["wt_ih"], M, D, 4 * D, P, resid=ws["demb"], out=dy)

    g_f, g_s = ws["g_f"], ws["g_s"]
    pd, seed = saved["p_drop"], saved["seed"]
    drop = lambda layer, which: (pd, seed, site_id(layer, which)) if pd > 0 else OT.NO_DROP
    keep_scale = 1.0 / (1.0 - pd) if pd > 0 else 1.0
    for i in range(model.n_layers - 1, -1, -1):
        lw, a = w["layers"][i], ws["layers"][i]
        pre = f"{model.NAMES['stack']}.layers.{i}."
        # norm2 / linear2 / ReLU / linear1
        OT.ln_bwd(dy, ws["x_f"][i + 1], lw["g2"], lw["be2"], a["rstd2"], work, G[pre + "norm2.weight"],
                  G[pre + "norm2.bias"], G[pre + "linear2.bias"], dx=g_f, dx_split=g_s, drop=drop(i, SITE_FFN2))
        OT.gemm_wgrad(g_s, a["ff"], M, D, FF, G[pre + "linear2.weight"], work, P)
        OT.gemm_dgrad(g_s, lw["wt_2"], M, FF, D, P, relu_mask=a["ff"], out_split=ws["dh"], mask_scale=keep_scale)
        OT.colsum_split(ws["dh"], M, FF, G[pre + "linear1.bias"], work)
        OT.gemm_wgrad(ws["dh"], a["y1_s"], M, FF, D, G[pre + "linear1.weight"], work, P)
        OT.gemm_dgrad(ws["dh"], lw["wt_1"], M, D, FF, P, resid=g_f, out=dy_other)
        # norm1 / out_proj / attention / in_proj
        OT.ln_bwd(dy_other, a["y1_f"], lw["g1"], lw["be1"], a["rstd1"], work, G[pre + "norm1.weight"],
                  G[pre + "norm1.bias"], G[pre + "self_attn.out_proj.bias"], dx=g_f, dx_split=g_s, drop=drop(i, SITE_OUT))
        OT.gemm_wgrad(g_s, a["ctx"], M, D, D, G[pre + "self_attn.out_proj.weight"], work, P)
        OT.gemm_dgrad(g_s, lw["wt_o"], M, D, D, P, out_split=ws["dctx"])
        OT.mha_delta(ws["dctx"], a["ctx"], M, D, ws["delta"])
        OT.mha_bwd(a["qk"], a["vt"], ws["Tp"], ws["dctx"], a["lse"], ws["delta"], B, Tmax, H, ws["dq_acc"],
                   ws["dqkv"], model.attn_precision, drop=drop(i, SITE_ATTN), drop_bits=saved["bits"][i] if pd > 0 else None)
        OT.scale_split(ws["dq_acc"], 0.125, ws["dqkv"], cols=D)
        OT.colsum_split(ws["dqkv"], M, 3 * D, G[pre + "self_attn.in_proj_bias"], work)
        OT.gemm_wgrad(ws["dqkv"], ws["x_s"][i], M, 3 * D, D, G[pre + "self_attn.in_proj_weight"], work, P)
        OT.gemm_dgrad(ws["dqkv"], lw["wt_qkv"], M, D, 3 * D, P, resid=g_f, out=dy)
    # encoder_norm / encoder (input projection)
    lin, nrm = model.NAMES["in_linear"], model.NAMES["in_norm"]
    OT.ln_bwd(dy, ws["x_f"][0], w["g_in"], w["be_in"], ws["rstd0"], work, G[nrm + ".weight"],
              G[nrm + ".bias"], G[lin + ".bias"], dx_split=g_s)
    OT.gemm_wgrad(g_s, ws["xin"], M, D, model.in_size, G[lin + ".weight"], work, P)
    return gflat


class EdaLossFn(torch.autograd.Function):
    """loss = EdaLossFn.apply(model, src, ts, *parameters): ties forward_train / backward_train to
    autograd so that ``loss.backward()`` behaves like the reference's (train_eda.py:184)."""

    @staticmethod
    def forward(ctx, model, src, ts, *params):
        loss, loss_att, loss_pit, counts, saved = forward_train(model, src, ts)
        ctx.model, ctx.saved = model, saved
        model._train_extras = (loss_att, loss_pit, counts)
        return loss

    @staticmethod
    def backward(ctx, gout):
        model = ctx.model
        gflat = backward_train(model, ctx.saved)
        OT.scale_dev(gout.detach().float().reshape(1).contiguous(), gflat)
        layout, _ = param_layout(model)
        sink = getattr(model, "_grad_sink", None)
        if sink is not None:               # NoamAdam owns flat gradients: add there, hand autograd nothing
            sink.add(gflat)
            return (None, None, None) + tuple(None for _ in layout)
        return (None, None, None) + tuple(gflat[off: off + n].view(p.shape) for _, p, off, n in layout)


# --------------------------------------------------------------------------- optimiser --------
def noam_lr(step: int, d_model: int, warmup: int, base_lr: float = 1.0) -> float:
    """lr_scheduler.py:29-33 (``last_epoch`` = number of scheduler.step() calls so far)."""
    s = max(1, step)
    return base_lr * d_model ** (-0.5) * min(s ** (-0.5), s * warmup ** (-1.5))


class NoamAdam:
    """``optim.Adam(model.parameters(), lr, betas=(0.9, 0.98), eps=1e-9)`` + ``NoamScheduler``
    (train_eda.py:131,137; stepped per optimiser step, :188-192) on flat device buffers.

    Construction moves the model's parameters into ONE flat fp32 buffer (each ``p.data`` becomes a
    view of it, values preserved) and binds ``p.grad`` to views of one flat gradient buffer that
    ``loss.backward()`` accumulates into directly.  ``step()`` = optional NCCL all-reduce (mean over
    ranks, what DDP does) + one fused Adam kernel."""

    def __init__(self, model, lr: float = 1.0, betas=(0.9, 0.98), eps: float = 1e-9, warmup_steps: int = 25000,
                 noam: bool = True, process_group=None):
        self.model, self.base_lr, self.betas, self.eps = model, lr, betas, eps
        self.warmup, self.noam, self.group = warmup_steps, noam, process_group
        self.layout, total = param_layout(model)
        dev = model._device()
        self.flat_p = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros(total, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(total, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(total, dtype=torch.float32, device=dev)
        for _, p, off, n in self.layout:
            self.flat_p[off: off + n].copy_(p.data.reshape(-1))
            p.data = self.flat_p[off: off + n].view(p.shape)
            p.grad = self.flat_g[off: off + n].view(p.shape)
        self.t = 0                      # optimiser steps taken
        self.sched_steps = 0            # scheduler.step() calls (NoamScheduler.last_epoch)
        model._grad_sink = self
        model._prep = None
        model.eda._prep = None

    # gradient sink protocol (EdaLossFn.backward)
    def add(self, gflat: Tensor) -> None:
        self.flat_g.add_(gflat)

    @property
    def lr(self) -> float:
        return noam_lr(self.sched_steps, self.model.n_units, self.warmup, self.base_lr) if self.noam else self.base_lr

    def zero_grad(self) -> None:
        self.flat_g.zero_()

    def step(self) -> None:
        allreduce_mean_(self.flat_g, self.group)
        self.t += 1
        OT.adam_step(self.flat_p, self.flat_g, self.exp_avg, self.exp_avg_sq, self.lr, self.betas[0], self.betas[1],
                     self.eps, self.t)
        self.sched_steps += 1           # the reference calls scheduler.step() right after optimizer.step()
        self.model._prep = None         # the kernels' split / transposed weight copies are stale now
        self.model.eda._prep = None


class DeviceBatches:
    """Iterate ``(src, ts)`` host batches (lists of (T_i, F) / (T_i, S) CPU tensors, what the reference's DataLoader
    + collate yields, train_eda.py:150-171) as lists of CUDA tensors, staging batch i+1 while step i runs: each batch is
    concatenated into a pinned buffer and sent with ONE H2D copy per side on a copy stream (what
    ``DataLoader(pin_memory=True)`` + ``.to(device, non_blocking=True)`` do for the reference); the tensors handed out
    are views into that device buffer.  ``depth`` staging sets alternate; a set is reused only after the copy that read
    it has completed and the step that consumed its device buffer has been enqueued."""

    def __init__(self, batches, device, depth: int = 2):
        self.batches, self.device, self.depth = batches, torch.device(device), max(2, depth)
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self._sets = [dict() for _ in range(self.depth + 1)]

    def _stage(self, k: int, seqs: List[Tensor], name: str):
        st = self._sets[k]
        total, cols = sum(int(s.shape[0]) for s in seqs), int(seqs[0].shape[1])
        key = name + "_dev"
        if key not in st or st[key].shape != (total, cols):
            st[key] = torch.empty((total, cols), dtype=torch.float32, device=self.device)
        views = list(st[key].split([int(s.shape[0]) for s in seqs], dim=0))
        if all(s.is_pinned() and s.dtype == torch.float32 for s in seqs):
            # what DataLoader(pin_memory=True) yields: one DMA per sequence straight into its slice, no host-side copy
            # (concatenating 88 MB on one host thread costs more than the whole GPU step)
            for v, s_ in zip(views, seqs):
                v.copy_(s_, non_blocking=True)
        else:
            hk = name + "_host"
            if hk not in st or st[hk].shape != (total, cols):
                st[hk] = torch.empty((total, cols), dtype=torch.float32, pin_memory=True)
            torch.cat([s_.float() for s_ in seqs], dim=0, out=st[hk])
            st[key].copy_(st[hk], non_blocking=True)
        return views

    def _submit(self, k: int, batch):
        st = self._sets[k]
        if "copied" in st:
            st["copied"].synchronize()                 # the pinned buffers of this set are free again
        src, ts = batch
        if "consumed" in st:
            self.copy_stream.wait_event(st["consumed"])    # the step that read this set's device buffers is enqueued and done
        with torch.cuda.stream(self.copy_stream):
            out = (self._stage(k, list(src), "src"), self._stage(k, list(ts), "ts"))
            st["copied"] = torch.cuda.Event()
            st["copied"].record(self.copy_stream)
        return out

    def __iter__(self):
        return self.iterate(self.batches)

    def iterate(self, batches):
        """Generator over ``batches`` (the staging buffers are kept between calls: a second pass allocates nothing)."""
        pending = []                                   # (set index, device batch)
        k = 0
        self._last_k = None
        for st in self._sets:
            st.pop("consumed", None)
        torch.cuda.current_stream(self.device).synchronize()
        for batch in batches:
            pending.append((k, self._submit(k, batch)))
            k = (k + 1) % len(self._sets)
            if len(pending) >= self.depth:
                yield self._hand_out(pending.pop(0))
        while pending:
            yield self._hand_out(pending.pop(0))

    def _hand_out(self, item):
        k, dev_batch = item
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self._sets[k]["copied"])
        # whatever the caller enqueues on the current stream before asking for the next batch reads this set
        prev = getattr(self, "_last_k", None)
        if prev is not None:
            ev = torch.cuda.Event()
            ev.record(cur)
            self._sets[prev]["consumed"] = ev
        self._last_k = k
        return dev_batch

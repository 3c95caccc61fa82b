"""CPU restatement (torch fp32) of the reference EEND-EDA model path.

TEST INFRASTRUCTURE (see oracle/__init__.py) — the checker, never the product.

Each function cites the reference lines it restates
(``R`` = /root/reference/speaker_diarization/eend_eda).  The restatement is
functional: it works on a plain ``state_dict`` and writes the math out with
``torch.nn.functional`` primitives instead of building the reference's modules, so a
misunderstanding of the module semantics (post-LN order, gate order, packed-sequence
final state, padded-frame attention) shows up as a golden mismatch rather than being
inherited.  ``library_lstm=True`` swaps the hand-written recurrence for
``torch._VF.lstm`` (what ``nn.LSTM`` calls) so the oracle is as fast as the reference
when it is timed as the CPU baseline.
"""
from __future__ import annotations

from itertools import permutations
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# --------------------------------------------------------------------------- encoder
def pad_batch(seqs: Sequence[Tensor], value: float = -1.0) -> Tuple[Tensor, List[int]]:
    """R/models.py:216-217 — ``pad_sequence(padding_value=-1, batch_first=True)``."""
    lens = [int(s.shape[0]) for s in seqs]
    out = seqs[0].new_full((len(seqs), max(lens)) + tuple(seqs[0].shape[1:]), value)
    for i, s in enumerate(seqs):
        out[i, : lens[i]] = s
    return out, lens


def encoder_layer(x: Tensor, sd: Dict[str, Tensor], prefix: str, n_heads: int, drop=None) -> Tensor:
    """One ``nn.TransformerEncoderLayer`` as the reference configures it
    (R/models.py:193: post-LN, ReLU, eps 1e-5, batch_first, no masks at all — padded
    frames are attended to like any other, models.py:223).

    ``drop`` = (p, (keep_attn (B,H,T,T), keep_out (B,T,D), keep_ffn (B,T,FF), keep_ffn2 (B,T,D))) restates
    train()-mode dropout with EXPLICIT keep masks at the layer's four sites (attention weights inside
    nn.MultiheadAttention, dropout1, the dropout after the activation, dropout2); kept values are
    divided by 1 - p like ``F.dropout``."""
    B, T, D = x.shape
    dh = D // n_heads
    dz = (lambda t, m: t * m.to(t.dtype) / (1.0 - drop[0])) if drop is not None else (lambda t, m: t)
    km = drop[1] if drop is not None else (None, None, None, None)
    qkv = F.linear(x, sd[prefix + "self_attn.in_proj_weight"], sd[prefix + "self_attn.in_proj_bias"])
    q, k, v = (t.reshape(B, T, n_heads, dh).transpose(1, 2) for t in qkv.split(D, dim=-1))
    att = dz(torch.softmax((q @ k.transpose(-1, -2)) * (dh ** -0.5), dim=-1), km[0]) @ v
    att = att.transpose(1, 2).reshape(B, T, D)
    att = dz(F.linear(att, sd[prefix + "self_attn.out_proj.weight"], sd[prefix + "self_attn.out_proj.bias"]), km[1])
    x = F.layer_norm(x + att, (D,), sd[prefix + "norm1.weight"], sd[prefix + "norm1.bias"], 1e-5)
    ff = dz(F.linear(x, sd[prefix + "linear1.weight"], sd[prefix + "linear1.bias"]).relu(), km[2])
    ff = dz(F.linear(ff, sd[prefix + "linear2.weight"], sd[prefix + "linear2.bias"]), km[3])
    return F.layer_norm(x + ff, (D,), sd[prefix + "norm2.weight"], sd[prefix + "norm2.bias"], 1e-5)


def n_layers_of(sd: Dict[str, Tensor]) -> int:
    n = 0
    while f"transformer_encoder.layers.{n}.norm1.weight" in sd:
        n += 1
    return n


def encode(padded: Tensor, sd: Dict[str, Tensor], n_heads: int, drop=None) -> Tensor:
    """R/models.py:220-223 — Linear → LayerNorm → L encoder layers (has_pos=False).
    ``drop`` = (p, [per-layer keep-mask 4-tuples]) for train()-mode dropout with explicit masks."""
    D = sd["encoder.weight"].shape[0]
    x = F.linear(padded, sd["encoder.weight"], sd["encoder.bias"])
    x = F.layer_norm(x, (D,), sd["encoder_norm.weight"], sd["encoder_norm.bias"], 1e-5)
    for l in range(n_layers_of(sd)):
        x = encoder_layer(x, sd, f"transformer_encoder.layers.{l}.", n_heads,
                          None if drop is None else (drop[0], drop[1][l]))
    return x


def draw_shuffle(lens: Sequence[int]) -> List[Tensor]:
    """R/models.py:229-232 — one ``torch.randperm(ilen)`` per sequence, in batch order,
    from the GLOBAL CPU generator (also in eval mode)."""
    return [torch.randperm(int(n)) for n in lens]


# ------------------------------------------------------------------------------- EDA
def lstm_cell(x_gates: Tensor, h: Tensor, c: Tensor, w_hh: Tensor) -> Tuple[Tensor, Tensor]:
    """PyTorch LSTM cell, gate order i,f,g,o.  ``x_gates`` already holds
    ``W_ih x + b_ih + b_hh``."""
    g = x_gates + h @ w_hh.t()
    i, f, gg, o = g.chunk(4, dim=-1)
    c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
    h = torch.sigmoid(o) * torch.tanh(c)
    return h, c


def eda(emb: Tensor, lens: Sequence[int], perms: Sequence[Tensor], n_dec: int,
        sd: Dict[str, Tensor], library_lstm: bool = False) -> Tuple[Tensor, Tensor]:
    """R/encoder_decoder_attractor.py:19-59 fed with the shuffled frames of
    R/models.py:228-232.

    Encoder LSTM runs over ``emb[b, perm_b[0]], emb[b, perm_b[1]], ...`` and stops at the
    sequence's own length (packed-sequence semantics, :36-39); its final (h, c) seed the
    decoder LSTM, which is stepped ``n_dec`` times on all-zero inputs (:40); the
    existence probability is ``sigmoid(Linear(attractor))`` (:50-54).
    Returns attractors (B, n_dec, D) and probs (B, n_dec).
    """
    B, T, D = emb.shape
    We, Ue = sd["eda.encoder.weight_ih_l0"], sd["eda.encoder.weight_hh_l0"]
    be = sd["eda.encoder.bias_ih_l0"] + sd["eda.encoder.bias_hh_l0"]
    Ud = sd["eda.decoder.weight_hh_l0"]
    bd = sd["eda.decoder.bias_ih_l0"] + sd["eda.decoder.bias_hh_l0"]
    h = emb.new_zeros((B, D))
    c = emb.new_zeros((B, D))
    if library_lstm:
        shuf = emb.clone()
        for b in range(B):
            shuf[b, : lens[b]] = emb[b, perms[b]]
        pack = torch.nn.utils.rnn.pack_padded_sequence(
            shuf, torch.as_tensor(list(lens)), batch_first=True, enforce_sorted=False)
        flat = [We, Ue, sd["eda.encoder.bias_ih_l0"], sd["eda.encoder.bias_hh_l0"]]
        hx = (h.unsqueeze(0)[:, pack.sorted_indices], c.unsqueeze(0)[:, pack.sorted_indices])
        _, hn, cn = torch._VF.lstm(pack.data, pack.batch_sizes, hx, flat, True, 1, 0.0, False, False)
        h, c = hn[0][pack.unsorted_indices], cn[0][pack.unsorted_indices]
    else:
        for b in range(B):
            xg = F.linear(emb[b, perms[b]], We, be)          # (len, 4D), hoisted input projection
            hb, cb = h[b:b + 1], c[b:b + 1]
            for t in range(int(lens[b])):
                hb, cb = lstm_cell(xg[t:t + 1], hb, cb, Ue)
            h[b], c[b] = hb[0], cb[0]
    atts = []
    for _ in range(n_dec):                                    # zero input ⇒ gates = b + U h
        h, c = lstm_cell(bd.unsqueeze(0), h, c, Ud)
        atts.append(h)
    att = torch.stack(atts, dim=1)
    logit = F.linear(att, sd["eda.linear.weight"], sd["eda.linear.bias"]).squeeze(-1)
    return att, torch.sigmoid(logit)


# ---------------------------------------------------------------------- loss pieces
def attractor_loss(probs: Tensor, n_spk: int) -> Tensor:
    """R/models.py:350-371 — BCE-with-logits applied to the ALREADY sigmoided existence
    probabilities against [1,...,1,0]; mean over S+1 then over B."""
    target = probs.new_zeros(probs.shape)
    target[:, :n_spk] = 1.0
    return F.binary_cross_entropy_with_logits(probs, target, reduction="none").mean(dim=1).mean()


def pit_loss(pred: Tensor, label: Tensor, lengths: Sequence[int]):
    """R/models.py:373-421.  For every permutation p (``itertools.permutations`` order)
    cost[b,p] = Σ_{t<len_b} mean_s BCE(pred[b,t,s], label[b,t,p[s]]); the first minimum
    wins; loss = Σ_b min / Σ_b len.  Returns (loss, min_idx int64 (B,), [np perms],
    label_permute float32 CPU (B,T,S))."""
    B, T, S = label.shape
    perms = [np.array(p) for p in permutations(range(S))]
    mask = (torch.arange(T).unsqueeze(0) < torch.as_tensor(list(lengths)).unsqueeze(1)).to(pred.dtype)
    costs = []
    for p in perms:
        bce = F.binary_cross_entropy_with_logits(pred, label[:, :, p], reduction="none")
        costs.append(((bce * mask.unsqueeze(-1)).mean(dim=2)).sum(dim=1))
    cost = torch.stack(costs, dim=1)
    best, idx = torch.min(cost, dim=1)
    loss = best.sum() / float(sum(int(l) for l in lengths))
    lab = torch.stack([label[b][:, perms[int(idx[b])]] for b in range(B)]).float().cpu()
    return loss, idx, perms, lab


def diarization_counts(pred: Tensor, label: Tensor, lengths: Sequence[int]):
    """R/models.py:423-463 — nine counters from decisions ``pred > 0`` under the length
    mask.  Returns the same 9-tuple of Python floats."""
    B, T, S = label.shape
    m = (np.arange(T)[None, :] < np.asarray(list(lengths))[:, None])[:, :, None]
    lab = label.detach().cpu().numpy().astype(int) * m
    dec = (pred.detach().cpu().numpy() > 0).astype(int) * m
    n_ref, n_sys = lab.sum(2), dec.sum(2)
    n_map = np.logical_and(lab == 1, dec == 1).sum(2)
    return (
        float(((lab == dec) * m).sum() / S),
        float(np.sum(list(lengths))),
        float((n_ref > 0).sum()),
        float(np.logical_and(n_ref > 0, n_sys == 0).sum()),
        float(np.logical_and(n_ref == 0, n_sys > 0).sum()),
        float(n_ref.sum()),
        float(np.maximum(n_ref - n_sys, 0).sum()),
        float(np.maximum(n_sys - n_ref, 0).sum()),
        float((np.minimum(n_ref, n_sys) - n_map).sum()),
    )


# ------------------------------------------------------------------ public entry points
class OracleEda:
    """Same call surface as the reference ``TransformerEdaModel`` (eval-mode math)."""

    def __init__(self, state_dict: Dict[str, Tensor], n_heads: int = 4,
                 diar_weight: float = 1.0, attractor_weight: float = 1.0,
                 library_lstm: bool = False):
        self.sd = {k: v.detach().float().cpu() for k, v in state_dict.items()}
        self.n_heads = n_heads
        self.diar_weight = diar_weight
        self.attractor_weight = attractor_weight
        self.library_lstm = library_lstm
        self.trace: Dict[str, object] = {}

    def forward_embedding(self, src: Sequence[Tensor]):
        padded, lens = pad_batch(src, -1.0)
        emb = encode(padded, self.sd, self.n_heads)
        perms = draw_shuffle(lens)
        return emb, lens, perms

    @torch.no_grad()
    def infer(self, src: Sequence[Tensor], infer_num_speakers: Optional[int] = None,
              max_n_speakers: int = 15, attractor_threshold: Optional[float] = 0.5) -> List[Tensor]:
        """R/models.py:297-347."""
        emb, lens, perms = self.forward_embedding(src)
        att, probs = eda(emb, lens, perms, max_n_speakers, self.sd, self.library_lstm)
        post = torch.sigmoid(emb @ att[:, :-1].transpose(1, 2))
        self.trace = dict(emb=emb, perms=perms, attractors=att, probs=probs, posteriors=post)
        out = []
        for b in range(len(lens)):
            if infer_num_speakers is not None:
                order = torch.sort(probs[b], descending=True)[1]
                out.append(post[b][:, order[:infer_num_speakers]])   # IndexError if order hits 14, like R:339
            elif attractor_threshold is not None:
                below = np.where(probs[b].numpy() < attractor_threshold)[0]
                out.append(post[b][:, : (int(below[0]) if below.size else None)])
        return out

    @torch.no_grad()
    def forward(self, src: Sequence[Tensor], ts: Sequence[Tensor]):
        """R/models.py:236-295 (eval-mode / dropout-free math)."""
        emb, lens, perms = self.forward_embedding(src)
        label, _ = pad_batch(ts, -1.0)
        S = label.shape[2]
        att, probs = eda(emb, lens, perms, S + 1, self.sd, self.library_lstm)
        pred = emb @ att[:, :-1].transpose(1, 2)
        l_att = attractor_loss(probs, S)
        l_pit, idx, perm_list, lab = pit_loss(pred, label, lens)
        loss = self.diar_weight * l_pit + self.attractor_weight * l_att
        (correct, nfr, sp_sc, sp_ms, sp_fa, spk_sc, spk_ms, spk_fa, spk_er) = \
            diarization_counts(pred, lab, lens)
        stats = dict(loss=loss, loss_att=l_att, loss_pit=l_pit,
                     sad_mr=sp_ms / sp_sc, sad_fr=sp_fa / sp_sc, mi=spk_ms / spk_sc,
                     fa=spk_fa / spk_sc, cf=spk_er / spk_sc, acc=correct / nfr,
                     der=(spk_ms + spk_fa + spk_er) / spk_sc)
        self.trace = dict(emb=emb, perms=perms, attractors=att, probs=probs, pred=pred,
                          min_idx=idx, label_permute=lab,
                          counts=(correct, nfr, sp_sc, sp_ms, sp_fa, spk_sc, spk_ms, spk_fa, spk_er))
        return loss, stats

    __call__ = forward


# ---------------------------------------------------------------- training step
def forward_backward(state_dict: Dict[str, Tensor], src: Sequence[Tensor], ts: Sequence[Tensor],
                     n_heads: int = 4, diar_weight: float = 1.0, attractor_weight: float = 1.0,
                     library_lstm: bool = True, drop=None):
    """R/models.py:236-263 followed by ``loss.backward()`` (R/train_eda.py:182-190), dropout-free.

    Returns (loss, {state_dict key: gradient}).  The random frame shuffle is drawn from the global
    generator exactly like ``forward``; the PIT permutation is chosen without gradient
    (R/models.py:401-403 picks ``min`` over the per-permutation losses, so only the winning
    permutation carries gradient)."""
    sd = {k: v.detach().float().cpu().clone().requires_grad_(True) for k, v in state_dict.items()}
    padded, lens = pad_batch(src, -1.0)
    with torch.enable_grad():
        emb = encode(padded, sd, n_heads, drop)        # drop = (p, per-layer explicit keep masks) or None
        perms = draw_shuffle(lens)
        label, _ = pad_batch(ts, -1.0)
        S = label.shape[2]
        att, probs = eda(emb, lens, perms, S + 1, sd, library_lstm)
        pred = emb @ att[:, :-1].transpose(1, 2)
        l_att = attractor_loss(probs, S)
        l_pit, idx, _, lab = pit_loss(pred, label, lens)
        loss = diar_weight * l_pit + attractor_weight * l_att
        loss.backward()
    grads = {k: (v.grad if v.grad is not None else torch.zeros_like(v)).detach() for k, v in sd.items()}
    return loss.detach(), grads, dict(emb=emb.detach(), pred=pred.detach(), min_idx=idx, label_permute=lab,
                                      attractors=att.detach(), probs=probs.detach())


def noam_lr(step: int, d_model: int, warmup: int, factor: float = 1.0) -> float:
    """R/lr_scheduler.py:6-33 (NoamScheduler): factor * d_model^-0.5 * min(step^-0.5, step * warmup^-1.5)."""
    step = max(step, 1)
    return factor * d_model ** (-0.5) * min(step ** (-0.5), step * warmup ** (-1.5))


def adam_step(params: Dict[str, Tensor], grads: Dict[str, Tensor], state: Dict[str, Tensor], step: int,
              lr: float, betas=(0.9, 0.98), eps: float = 1e-9):
    """torch.optim.Adam as R/train_eda.py:131 configures it for 'noam' (betas (0.9, 0.98), eps 1e-9),
    restated; ``step`` counts from 1.  (The reference's clip_grad_value_ at R/train_eda.py:194-199
    runs AFTER optimizer.zero_grad(), i.e. on empty gradients: it never changes an update.)"""
    b1, b2 = betas
    for k, p in params.items():
        g = grads[k]
        m = state.setdefault("m." + k, torch.zeros_like(p))
        v = state.setdefault("v." + k, torch.zeros_like(p))
        m.mul_(b1).add_(g, alpha=1 - b1)
        v.mul_(b2).addcmul_(g, g, value=1 - b2)
        denom = (v.sqrt() / (1 - b2 ** step) ** 0.5).add_(eps)
        p.addcdiv_(m, denom, value=-lr / (1 - b1 ** step))
    return params

"""CPU restatement (torch fp32) of the reference's Conformer variant:
``EendEdaModel(encoder_type="conformer")`` (R/models.py:465-700; R = /root/reference/speaker_diarization/eend_eda),
whose encoder is ``torchaudio.models.Conformer(input_dim, num_heads, ffn_dim, num_layers, 31)``
(R/models.py:501) applied with the sequence lengths as key_padding_mask (R/models.py:527-528).

TEST INFRASTRUCTURE (see oracle/__init__.py).  torchaudio is a third-party dependency of the reference
(requirements.txt; 2.11 in this image): its published layer structure (torchaudio/models/conformer.py:
_FeedForwardModule, _ConvolutionModule, ConformerLayer.forward) is written out functionally here on a plain
``state_dict`` and pinned against the unmodified reference class by oracle/make_golden_conformer.py.
Eval-mode math: dropout off, BatchNorm1d uses its running statistics.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F

from .eend_eda_oracle import attractor_loss, diarization_counts, draw_shuffle, eda, pad_batch, pit_loss

Tensor = torch.Tensor


def _ffn(x: Tensor, sd, p: str) -> Tensor:
    """_FeedForwardModule: LayerNorm -> Linear -> SiLU -> Linear."""
    D = x.shape[-1]
    y = F.layer_norm(x, (D,), sd[p + "sequential.0.weight"], sd[p + "sequential.0.bias"], 1e-5)
    y = F.silu(F.linear(y, sd[p + "sequential.1.weight"], sd[p + "sequential.1.bias"]))
    return F.linear(y, sd[p + "sequential.4.weight"], sd[p + "sequential.4.bias"])


def _conv(x: Tensor, sd, p: str) -> Tensor:
    """_ConvolutionModule on (B, T, D): LayerNorm -> pointwise (D -> 2D) -> GLU -> depthwise (k = 31, 'same' zero
    padding over the PADDED time axis) -> BatchNorm1d (running statistics) -> SiLU -> pointwise."""
    D = x.shape[-1]
    y = F.layer_norm(x, (D,), sd[p + "layer_norm.weight"], sd[p + "layer_norm.bias"], 1e-5).transpose(1, 2)
    y = F.glu(F.conv1d(y, sd[p + "sequential.0.weight"], sd[p + "sequential.0.bias"]), dim=1)
    w = sd[p + "sequential.2.weight"]
    y = F.conv1d(y, w, sd[p + "sequential.2.bias"], padding=(w.shape[-1] - 1) // 2, groups=D)
    y = F.batch_norm(y, sd[p + "sequential.3.running_mean"], sd[p + "sequential.3.running_var"],
                     sd[p + "sequential.3.weight"], sd[p + "sequential.3.bias"], False, 0.0, 1e-5)
    y = F.conv1d(F.silu(y), sd[p + "sequential.5.weight"], sd[p + "sequential.5.bias"])
    return y.transpose(1, 2)


def conformer_layer(x: Tensor, lens: Sequence[int], sd, p: str, n_heads: int) -> Tensor:
    """ConformerLayer.forward (convolution_first=False) on batch-first (B, T, D)."""
    B, T, D = x.shape
    dh = D // n_heads
    x = 0.5 * _ffn(x, sd, p + "ffn1.") + x
    y = F.layer_norm(x, (D,), sd[p + "self_attn_layer_norm.weight"], sd[p + "self_attn_layer_norm.bias"], 1e-5)
    qkv = F.linear(y, sd[p + "self_attn.in_proj_weight"], sd[p + "self_attn.in_proj_bias"])
    q, k, v = (t.reshape(B, T, n_heads, dh).transpose(1, 2) for t in qkv.split(D, dim=-1))
    s = (q @ k.transpose(-1, -2)) * (dh ** -0.5)
    pad = torch.arange(T).unsqueeze(0) >= torch.as_tensor(list(lens)).unsqueeze(1)          # key_padding_mask
    s = s.masked_fill(pad[:, None, None, :], float("-inf"))
    a = (torch.softmax(s, dim=-1) @ v).transpose(1, 2).reshape(B, T, D)
    x = F.linear(a, sd[p + "self_attn.out_proj.weight"], sd[p + "self_attn.out_proj.bias"]) + x
    x = x + _conv(x, sd, p + "conv_module.")
    x = 0.5 * _ffn(x, sd, p + "ffn2.") + x
    return F.layer_norm(x, (D,), sd[p + "final_layer_norm.weight"], sd[p + "final_layer_norm.bias"], 1e-5)


def n_conformer_layers(sd) -> int:
    n = 0
    while f"encoder.conformer_layers.{n}.final_layer_norm.weight" in sd:
        n += 1
    return n


def encode_conformer(padded: Tensor, lens: Sequence[int], sd, n_heads: int) -> Tensor:
    """R/models.py:521-528: Linear -> LayerNorm -> Conformer(src, ilens)."""
    D = sd["linear.weight"].shape[0]
    x = F.layer_norm(F.linear(padded, sd["linear.weight"], sd["linear.bias"]), (D,), sd["linear_norm.weight"],
                     sd["linear_norm.bias"], 1e-5)
    for l in range(n_conformer_layers(sd)):
        x = conformer_layer(x, lens, sd, f"encoder.conformer_layers.{l}.", n_heads)
    return x


class OracleEendConformer:
    """infer / forward of R/models.py:563-652 (eval-mode math) for encoder_type='conformer'."""

    def __init__(self, state_dict: Dict[str, Tensor], n_heads: int = 4, library_lstm: bool = True):
        self.sd = {k: v.detach().float().cpu() for k, v in state_dict.items() if v.is_floating_point()}
        self.n_heads, self.library_lstm = n_heads, library_lstm
        self.trace = {}

    def _emb(self, src):
        padded, lens = pad_batch(src, -1.0)
        emb = encode_conformer(padded, lens, self.sd, self.n_heads)
        return emb, lens, draw_shuffle(lens)

    @torch.no_grad()
    def infer(self, src, infer_num_speakers: Optional[int] = None, max_n_speakers: int = 15,
              attractor_threshold: Optional[float] = 0.5) -> List[Tensor]:
        emb, lens, perms = self._emb(src)
        att, probs = eda(emb, lens, perms, max_n_speakers, self.sd, self.library_lstm)
        post = torch.sigmoid(emb @ att[:, :-1].transpose(1, 2))
        self.trace = dict(emb=emb, attractors=att, probs=probs, posteriors=post)
        out = []
        for b in range(len(lens)):
            if infer_num_speakers is not None:
                out.append(post[b][:, :infer_num_speakers])                     # R/models.py:644
            else:
                below = np.where(probs[b].numpy() < attractor_threshold)[0]
                out.append(post[b][:, : (int(below[0]) if below.size else None)])
        return out

    @torch.no_grad()
    def forward(self, src, ts):
        emb, lens, perms = self._emb(src)
        label, _ = pad_batch(ts, -1.0)
        S = label.shape[2]
        att, probs = eda(emb, lens, perms, S + 1, self.sd, self.library_lstm)
        pred = emb @ att[:, :-1].transpose(1, 2)
        l_att = attractor_loss(probs, S)
        l_pit, idx, _, lab = pit_loss(pred, label, lens)
        c = diarization_counts(pred, lab, lens)
        self.trace = dict(emb=emb, pred=pred, min_idx=idx, label_permute=lab, counts=c)
        return l_pit + l_att, dict(der=(c[6] + c[7] + c[8]) / c[5])

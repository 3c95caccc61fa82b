"""LSTM encoder-decoder attractor, B200-native.

Drop-in for the reference ``LstmEncoderDedecoderAttractor``
(speaker_diarization/eend_eda/encoder_decoder_attractor.py:8-59): same constructor, same
``state_dict`` keys (``encoder.*_l0``, ``decoder.*_l0``, ``linear.*``), same
``forward(emb, emb_length, decoder_zero_input) -> (attractors, [probs])`` contract.  The
``nn.LSTM`` / ``nn.Linear`` members only OWN the parameters; the arithmetic runs in the
hand-written kernels behind ``eda_linear`` / ``eda_lstm_attractors`` / ``eda_attractor_head``.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.nn as nn

from .. import _lib as L
from .. import ops


class LstmEncoderDedecoderAttractor(nn.Module):
    def __init__(self, input_size, num_layers=1, dropout=0.1):
        super().__init__()
        if num_layers != 1:
            raise NotImplementedError("the reference only ever builds a 1-layer EDA (models.py:196)")
        self.encoder = nn.LSTM(input_size, input_size, num_layers, batch_first=True)
        self.decoder = nn.LSTM(input_size, input_size, num_layers, batch_first=True)
        self.linear = nn.Linear(input_size, 1)
        self.dropout_layer = torch.nn.Dropout(p=dropout)   # defined, never applied (reference :15)
        self.input_size = input_size
        self.precision = L.PREC_FP32
        self._prep = None
        self._prep_event = None
        self._order = {}

    # -- cached kernel-side views of the parameters ------------------------------------
    def _prepared(self, want_split: bool = False):
        key = tuple((p.data_ptr(), p._version) for p in self.parameters())
        if self._prep is None or self._prep[0] != key:
            e, d = self.encoder, self.decoder
            self._prep_event = None
            self._prep = (key, dict(
                w_ih=e.weight_ih_l0.detach().contiguous(),
                b_enc=(e.bias_ih_l0 + e.bias_hh_l0).detach().contiguous(),
                u_enc=e.weight_hh_l0.detach().contiguous(),
                u_dec=d.weight_hh_l0.detach().contiguous(),
                b_dec=(d.bias_ih_l0 + d.bias_hh_l0).detach().contiguous(),
                w_lin=self.linear.weight.detach().reshape(-1).contiguous(),
                b_lin=self.linear.bias.detach().contiguous(),
            ))
        w = self._prep[1]
        if want_split and "sw_ih" not in w:
            w["sw_ih"] = ops.split_bf16(w["w_ih"])
            self._prep_event = None
        if w["w_ih"].is_cuda:
            # the copies are made on whichever stream asks first: every other stream is ordered after it
            if self._prep_event is None:
                self._prep_event = torch.cuda.Event()
                self._prep_event.record(torch.cuda.current_stream(w["w_ih"].device))
            torch.cuda.current_stream(w["w_ih"].device).wait_event(self._prep_event)
        return w

    def attractors(self, emb2d: torch.Tensor, perm: torch.Tensor, lens_dev: torch.Tensor,
                   lens_host: List[int], B: int, Tmax: int, n_dec: int,
                   n_spk_label: int = -1, emb_split=None):
        """emb2d (B*Tmax, D) UNshuffled; perm (B, Tmax) int32 frame order per sequence.
        Returns attractors (B, n_dec, D), probs (B, n_dec), attractor loss (1,) or None."""
        use_split = emb_split is not None and self.precision != L.PREC_FP32
        w = self._prepared(use_split)
        if use_split:
            # the encoder left emb as a split-bf16 pair: the input projection reads it directly
            es, xproj = emb_split
            ops.gemm_split(es, w["sw_ih"], w["b_enc"], B * Tmax, 4 * self.input_size, self.input_size,
                           L.EPI_BIAS, self.precision, out=xproj)
        else:
            xproj = ops.linear(emb2d, w["w_ih"], w["b_enc"], L.EPI_BIAS, self.precision)
        okey = (tuple(lens_host), str(emb2d.device))
        order = self._order.get(okey)
        if order is None:                       # sequences co-scheduled by length; cached per geometry, blocking copy once
            if len(self._order) > 256:
                self._order.clear()
            order = self._order[okey] = torch.tensor(sorted(range(B), key=lambda b: -lens_host[b]), dtype=torch.int32).to(emb2d.device)
        att, _ = ops.lstm_attractors(xproj, perm, lens_dev, order, w["u_enc"], w["u_dec"], w["b_dec"],
                                     B, Tmax, n_dec, prec=self.precision)
        probs, loss_att = ops.attractor_head(att, w["w_lin"], w["b_lin"], n_spk_label)
        return att, probs, loss_att

    def forward(self, emb: torch.Tensor, emb_length: torch.Tensor,
                decoder_zero_input: torch.Tensor) -> Tuple[torch.Tensor, List[torch.Tensor]]:
        """Reference signature (:19).  ``emb`` (B,T,D) is consumed in the order given (the
        caller has already shuffled it); ``decoder_zero_input`` only conveys n_dec."""
        assert emb.size(2) == self.input_size
        if not emb.is_cuda:
            raise L.EdaError("speaker_diarization_b200 has no CPU path: move the model and inputs to cuda")
        if bool((decoder_zero_input != 0).any()):
            raise NotImplementedError("decoder input must be all zeros (models.py:247,321)")
        B, T, D = emb.shape
        lens_host = [int(x) for x in emb_length.tolist()]
        lens_dev = torch.tensor(lens_host, dtype=torch.int32, device=emb.device)
        perm = torch.arange(T, dtype=torch.int32, device=emb.device).repeat(B, 1).contiguous()
        att, probs, _ = self.attractors(emb.detach().float().contiguous().view(B * T, D), perm, lens_dev,
                                        lens_host, B, T, decoder_zero_input.size(1))
        return att, [p for p in probs]

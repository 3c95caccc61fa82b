"""B200-native ``TransformerEdaModel``.

Drop-in for the reference class of the same name
(speaker_diarization/eend_eda/models.py:160-463): same constructor arguments, same
``state_dict`` key set (checkpoints interchange with ``torch.load`` as in
infer_eda.py:88), same ``forward`` / ``infer`` / ``pit_loss`` / ``attractor_loss`` /
``calc_diarization_error`` / ``forward_embedding`` call signatures and return contracts.

The torch sub-modules are parameter CONTAINERS only (they give the reference's key names
and default initialisation for free); no arithmetic goes through them.  Everything on the
path runs in the hand-written sm_100a kernels of libeda_b200.so via ``..ops``.  There is
no CPU path: the model must live on a CUDA device.

Training: in ``train()`` mode with gradients enabled and a tensor-core precision,
``forward`` returns a loss with a ``grad_fn`` whose backward runs the hand-written backward kernels
(``training.py``); in ``eval()`` mode or under ``no_grad`` it is the inference-cost forward.

Not covered: ``has_pos=True`` (never enabled by the reference drivers, train_eda.py:81), gradients
under ``precision='fp32'``.
"""
from __future__ import annotations

import os
from itertools import permutations
from typing import List, Optional

import numpy as np
import torch
import torch.nn as nn
from torch import Tensor
from torch.nn import TransformerEncoder, TransformerEncoderLayer

from .. import _lib as L
from .. import ops
from .encoder_decoder_attractor import LstmEncoderDedecoderAttractor


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class TransformerEdaModel(nn.Module):
    def __init__(self, n_speakers, in_size, n_heads, n_units, n_layers, dim_feedforward=2048,
                 dropout=0.5, has_pos=False, diar_weight: float = 1.0, attractor_weight: float = 1.0,
                 precision: Optional[str] = None):
        super().__init__()
        if has_pos:
            raise NotImplementedError("has_pos=True is never used by the reference drivers (train_eda.py:81)")
        if n_units != 256 or n_units // n_heads != 64:
            raise NotImplementedError("kernels are specialised for n_units=256, head_dim=64 (conf/eend_eda/*.yaml)")
        self.n_speakers, self.in_size, self.n_heads = n_speakers, in_size, n_heads
        self.n_units, self.n_layers, self.has_pos = n_units, n_layers, has_pos
        self.dim_feedforward, self.dropout_p = dim_feedforward, dropout
        self.src_mask = None
        self._build_modules(in_size, n_units, n_heads, n_layers, dim_feedforward, dropout)
        self.eda = LstmEncoderDedecoderAttractor(input_size=n_units, num_layers=1, dropout=0.1)
        self.diar_weight, self.attractor_weight = diar_weight, attractor_weight
        self.set_precision(precision or os.environ.get("EDA_B200_PRECISION", "fp32"))
        self._prep = None
        self._ws = {}
        self._plist = None
        self._arange = None
        self.workspace_slot = 0     # callers with several batches in flight on different streams use one slot each
        self._perm_ring = {}
        self._emb_split = None
        self._prep_event = None
        self._geo = {}
        self._prep_gen = 0

    # state_dict names of the input projection, its LayerNorm and the encoder stack (models.py:185-193)
    NAMES = dict(in_linear="encoder", in_norm="encoder_norm", stack="transformer_encoder")

    def _build_modules(self, in_size, n_units, n_heads, n_layers, dim_feedforward, dropout):
        self.encoder = nn.Linear(in_size, n_units)
        self.encoder_norm = nn.LayerNorm(n_units)
        layer = TransformerEncoderLayer(n_units, n_heads, dim_feedforward, dropout, batch_first=True)
        self.transformer_encoder = TransformerEncoder(layer, n_layers, enable_nested_tensor=False)
        self.init_weights()

    def _in_linear(self) -> nn.Linear:
        return getattr(self, self.NAMES["in_linear"])

    def _in_norm(self) -> nn.LayerNorm:
        return getattr(self, self.NAMES["in_norm"])

    def _stack(self) -> TransformerEncoder:
        return getattr(self, self.NAMES["stack"])

    def set_precision(self, name: str) -> None:
        """'fp32' (CUDA-core FFMA), 'bf16x3' (tcgen05, split-bf16 operands) or 'bf16'."""
        if name not in L.PRECISIONS:
            raise ValueError(f"unknown precision {name!r}; choose from {sorted(L.PRECISIONS)}")
        self.precision_name, self.precision = name, L.PRECISIONS[name]
        self.eda.precision = self.precision
        self.attn_precision = self.precision

    def init_weights(self):
        """models.py:207-210."""
        self.encoder.bias.data.zero_()
        self.encoder.weight.data.uniform_(-0.1, 0.1)

    # ---------------------------------------------------------------- parameters --------
    def _prepared(self):
        if self._plist is None:                 # the Parameter objects are fixed after construction
            self._plist = list(self.parameters())
        key = tuple((p.data_ptr(), p._version) for p in self._plist)
        if self._prep is None or self._prep[0] != key:
            kpad = _round_up(self.in_size, 16)
            w = self._in_linear().weight.detach()
            w_in = torch.zeros((self.n_units, kpad), dtype=torch.float32, device=w.device)
            w_in[:, : self.in_size] = w
            layers = []
            for lyr in self._stack().layers:
                g = lambda t: t.detach().contiguous()
                layers.append(dict(
                    w_qkv=g(lyr.self_attn.in_proj_weight), b_qkv=g(lyr.self_attn.in_proj_bias),
                    w_o=g(lyr.self_attn.out_proj.weight), b_o=g(lyr.self_attn.out_proj.bias),
                    w_1=g(lyr.linear1.weight), b_1=g(lyr.linear1.bias),
                    w_2=g(lyr.linear2.weight), b_2=g(lyr.linear2.bias),
                    g1=g(lyr.norm1.weight), be1=g(lyr.norm1.bias),
                    g2=g(lyr.norm2.weight), be2=g(lyr.norm2.bias)))
            prep = dict(kpad=kpad, w_in=w_in, b_in=self._in_linear().bias.detach().contiguous(),
                        g_in=self._in_norm().weight.detach().contiguous(),
                        be_in=self._in_norm().bias.detach().contiguous(), layers=layers)
            if w.is_cuda:
                # split-bf16 copies of the weight matrices for the tensor-core pipeline (made once)
                prep["ws_in"] = ops.split_bf16(w_in)
                for lw in layers:
                    for k in ("w_qkv", "w_o", "w_1", "w_2"):
                        lw["s" + k] = ops.split_bf16(lw[k])
            self._prep = (key, prep)
            self._prep_gen += 1
            self._prep_event = self._record_event(w.device)
        self._wait_event(self._prep_event)
        return self._prep[1]

    @staticmethod
    def _record_event(dev):
        if dev.type != "cuda":
            return None
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(dev))
        return ev

    @staticmethod
    def _wait_event(ev):
        """Order the current stream after the stream that built the cached weight copies (several compute
        streams share them: InferencePipeline, the dropout side stream of the training step)."""
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)

    def _workspace(self, B: int, Tmax: int, dev):
        """Buffers of the tensor-core pipeline for a (B, Tmax) batch: views carved out of ONE arena per
        (device, stream slot) that only grows to the largest batch seen, so ragged chunk batches
        (infer_eda.py's per-recording loop: a different tail length per recording) neither pile up
        multi-GB workspaces per shape nor thrash the allocator.  A shape always gets the same pointers
        while the arena stands, which keeps libeda_b200's tensor-map cache (keyed on pointer + dims) warm."""
        akey = ("arena", str(dev), self.workspace_slot)
        arena = self._ws.get(akey)
        if arena is not None:
            ws = arena["views"].get((B, Tmax))
            if ws is not None:
                return ws
        M, D, H, FF = B * Tmax, self.n_units, self.n_heads, self.dim_feedforward
        Tp, kpad = _round_up(Tmax, 8), _round_up(self.in_size, 16)
        plan = [("xin", torch.bfloat16, (2, M, kpad)), ("xs0", torch.bfloat16, (2, M, D)), ("xs1", torch.bfloat16, (2, M, D)),
                ("xf0", torch.float32, (M, D)), ("xf1", torch.float32, (M, D)), ("qk", torch.bfloat16, (2, M, 2 * D)),
                ("vt", torch.bfloat16, (2, B * H * (D // H), Tp)), ("att", torch.bfloat16, (2, M, D)),
                ("ff", torch.bfloat16, (2, M, FF)), ("xproj", torch.float32, (M, 4 * D))]
        sizes = [_round_up(int(np.prod(shape)) * torch.empty((), dtype=dt).element_size(), 1024) for _, dt, shape in plan]
        need = sum(sizes)
        arena = self._ws.get(akey)
        if arena is None or arena["buf"].numel() < need:
            arena = dict(buf=torch.empty(need, dtype=torch.uint8, device=dev), views={})
            self._ws[akey] = arena
        ws = arena["views"].get((B, Tmax))
        if ws is None:
            if len(arena["views"]) >= 32:
                arena["views"].clear()
            t, pos = {}, 0
            for (name, dt, shape), nbytes in zip(plan, sizes):
                t[name] = arena["buf"][pos:pos + nbytes].view(dt)[: int(np.prod(shape))].view(shape)
                pos += nbytes
            sp = lambda x: ops.Split(x[0], x[1])
            # key columns [Tmax, Tp) of v^T are never written by the QKV epilogue and must not hold NaN bit patterns
            ws = dict(Tp=Tp, plans={}, vt_pad=t["vt"][:, :, Tmax:] if Tp != Tmax else None, xin=sp(t["xin"]), xs=[sp(t["xs0"]), sp(t["xs1"])], xf=[t["xf0"], t["xf1"]], qk=sp(t["qk"]),
                      vt=sp(t["vt"]), att=sp(t["att"]), ff=sp(t["ff"]), xproj=t["xproj"])
            arena["views"][(B, Tmax)] = ws
        return ws

    def _device(self) -> torch.device:
        dev = self._in_linear().weight.device
        if dev.type != "cuda":
            raise L.EdaError("speaker_diarization_b200 has no CPU path: call .cuda() on the model first")
        return dev

    def _check_mode(self):
        """The inference-cost forward has no dropout: it serves eval() mode (the reference's dev loop,
        train_eda.py:205) and train() mode only when dropout is 0; train-mode dropout lives in the
        training path (training.py), which needs gradients enabled and a tensor-core precision."""
        if self.training and self.dropout_p > 0:
            raise NotImplementedError(
                "train-mode dropout runs in the training path only (gradients enabled, precision 'bf16x3' or "
                "'bf16'); call .eval() for inference or build with dropout=0")

    # ------------------------------------------------------------------- encoder --------
    def _to_device_flat(self, seqs: List[Tensor], dev) -> Tensor:
        """Concatenate a list of (T_i, C) tensors into one device matrix with ONE H2D copy."""
        if seqs[0].is_cuda:
            return torch.cat([s.float() for s in seqs], dim=0).contiguous()
        total = sum(int(s.shape[0]) for s in seqs)
        stage = torch.empty((total, seqs[0].shape[1]), dtype=torch.float32, pin_memory=True)
        torch.cat([s.float() for s in seqs], dim=0, out=stage)
        return stage.to(dev, non_blocking=True)

    def _embed(self, src, ilens: Optional[List[int]] = None, independent: bool = False, offsets=None):
        """models.py:213-226 up to ``emb``; returns emb2d (B*Tmax, D) plus batch geometry.
        ``src`` is the reference's list of (T_i, F) tensors, or (with ``ilens``) ONE device
        matrix holding the sequences back to back (row stride may exceed F); ``offsets`` then gives each
        sequence's first row in it (default: back to back in the order of ``ilens``).
        ``independent``: attention masks keys >= ilen, so every sequence gets the result of a batch-of-one
        call - the reference's inference loop (infer_eda.py:102-113), where padding never exists."""
        dev = self._device()
        self._check_mode()
        w = self._prepared()
        flat_in = ilens is not None
        if not flat_in:
            ilens = [int(x.shape[0]) for x in src]
        B, Tmax = len(ilens), max(ilens)
        masked = independent and min(ilens) != Tmax
        seq_off, lens_dev = self._geometry(ilens, offsets, dev)
        flat = src if flat_in else self._to_device_flat(src, dev)
        P = self.precision
        if flat_in and P == L.PREC_FP32 and not flat.is_contiguous():
            flat = flat.contiguous()
        if P != L.PREC_FP32:
            return self._embed_tc(flat, w, seq_off, lens_dev, ilens, B, Tmax, dev, masked)
        self._emb_split = None
        x = ops.pad_pack(flat, seq_off, lens_dev, B, Tmax, w["kpad"], -1.0)
        x = ops.linear(x, w["w_in"], w["b_in"], L.EPI_BIAS_LN, P, gamma=w["g_in"], beta=w["be_in"])
        for lw in w["layers"]:
            qkv = ops.linear(x, lw["w_qkv"], lw["b_qkv"], L.EPI_BIAS, P)
            att = ops.mha(qkv, B, Tmax, self.n_heads, self.attn_precision, key_lens=lens_dev if masked else None)
            x = ops.linear(att, lw["w_o"], lw["b_o"], L.EPI_BIAS_RES_LN, P, resid=x, gamma=lw["g1"], beta=lw["be1"])
            ff = ops.linear(x, lw["w_1"], lw["b_1"], L.EPI_BIAS_RELU, P)
            x = ops.linear(ff, lw["w_2"], lw["b_2"], L.EPI_BIAS_RES_LN, P, resid=x, gamma=lw["g2"], beta=lw["be2"])
        return x, ilens, lens_dev, seq_off, B, Tmax

    def _geometry(self, ilens, offsets, dev):
        """(row offsets int64 (B+1), lengths int32 (B)) of a batch on the device; cached per geometry (the pipeline sees
        the same few geometries over and over), made with a blocking copy so that any stream may read them."""
        key = (tuple(ilens), None if offsets is None else tuple(int(o) for o in offsets), str(dev))
        hit = self._geo.get(key)
        if hit is None:
            if len(self._geo) > 256:
                self._geo.clear()
            B = len(ilens)
            off = np.zeros(B + 1, dtype=np.int64)
            if offsets is not None:
                off[:B] = np.asarray(offsets, dtype=np.int64)
                off[B] = off[B - 1] + ilens[-1]
            else:
                np.cumsum(ilens, out=off[1:])
            hit = self._geo[key] = (torch.from_numpy(off).to(dev), torch.tensor([int(n) for n in ilens], dtype=torch.int32).to(dev))
        return hit

    def _embed_tc(self, flat, w, seq_off, lens_dev, ilens, B, Tmax, dev, masked=False):
        """The encoder on the split-bf16 tensor-core pipeline: every GEMM epilogue writes the
        next kernel's operands (split bf16) and, where needed, the fp32 residual stream."""
        P, D, H = self.precision, self.n_units, self.n_heads
        M = B * Tmax
        ws = self._workspace(B, Tmax, dev)
        if ws["vt_pad"] is not None:
            ws["vt_pad"].zero_()
        emb = torch.empty((M, D), dtype=torch.float32, device=dev)      # returned to the caller: fresh
        self._emb_split = (ws["xs"][0], ws["xproj"])
        # The ~20 launches of the encoder are the same for every batch of a geometry except for five pointers: replay a
        # recording of them (made by the first call through the regular bindings) with those pointers patched.
        key = (masked, self._prep_gen, P, self.attn_precision, int(flat.stride(0)), int(flat.shape[1]))
        plan = ws["plans"].get(key) if L.PROFILE is None and L.RECORD is None else None
        var = dict(flat=flat.data_ptr(), seq_off=seq_off.data_ptr(), lens=lens_dev.data_ptr(), emb=emb.data_ptr())
        if plan is not None:
            st = L.stream()
            for fn, args, patches in plan:
                for pos, which in patches:
                    args[pos] = var[which]
                args[-1] = st
                rc = fn(*args)
                if rc != 0:
                    raise L.EdaError(f"{fn.__name__} failed ({rc}): {L.load().eda_last_error_string().decode()}")
            return emb, ilens, lens_dev, seq_off, B, Tmax
        record = L.PROFILE is None and L.RECORD is None
        if record:
            L.RECORD = []
        try:
            self._embed_tc_calls(flat, w, seq_off, lens_dev, B, Tmax, ws, emb, masked)
        finally:
            rec, L.RECORD = L.RECORD, None
        if record and len(set(var.values())) == len(var):
            lib, inv, plan = L.load(), {v: k for k, v in var.items()}, []
            for name, args in rec:
                patches = [(i, inv[a]) for i, a in enumerate(args[:-1]) if isinstance(a, int) and not isinstance(a, bool) and a in inv]
                plan.append((getattr(lib, name), args, patches))
            if len(ws["plans"]) > 8:
                ws["plans"].clear()
            ws["plans"][key] = plan
        return emb, ilens, lens_dev, seq_off, B, Tmax

    def _embed_tc_calls(self, flat, w, seq_off, lens_dev, B, Tmax, ws, emb, masked):
        P, D, H = self.precision, self.n_units, self.n_heads
        M = B * Tmax
        Tp = ws["Tp"]
        xin, (sa, sb), (xa, xb) = ws["xin"], ws["xs"], ws["xf"]
        ops.pad_pack_split(flat, seq_off, lens_dev, B, Tmax, xin, -1.0)
        n_layers = len(w["layers"])
        ops.gemm_split(xin, w["ws_in"], w["b_in"], M, D, xin.ld, L.EPI_BIAS_LN, P, gamma=w["g_in"], beta=w["be_in"],
                       out=xa if n_layers else emb, out_split=sa)
        for i, lw in enumerate(w["layers"]):
            last = i == n_layers - 1
            ops.gemm_qkv(sa, lw["sw_qkv"], lw["b_qkv"], ws["qk"], ws["vt"], B, Tmax, Tp, D, H, P)
            if masked:
                ops.mha_split_masked(ws["qk"], ws["vt"], B, Tmax, Tp, H, lens_dev, self.attn_precision, ws["att"])
            else:
                ops.mha_split(ws["qk"], ws["vt"], B, Tmax, Tp, H, self.attn_precision, out_split=ws["att"])
            ops.gemm_split(ws["att"], lw["sw_o"], lw["b_o"], M, D, D, L.EPI_BIAS_RES_LN, P, resid=xa,
                           gamma=lw["g1"], beta=lw["be1"], out=xb, out_split=sb)
            ops.gemm_split(sb, lw["sw_1"], lw["b_1"], M, self.dim_feedforward, D, L.EPI_BIAS_RELU, P,
                           out_split=ws["ff"])
            ops.gemm_split(ws["ff"], lw["sw_2"], lw["b_2"], M, D, self.dim_feedforward, L.EPI_BIAS_RES_LN, P,
                           resid=xb, gamma=lw["g2"], beta=lw["be2"], out=emb if last else xa, out_split=sa)

    def _perm_slot(self, B: int, Tmax: int):
        """A pinned (B, Tmax) int32 staging buffer prefilled with the identity order, from a small ring
        (an event per buffer guards the asynchronous H2D copy that last read it)."""
        ring = self._perm_ring.setdefault((B, Tmax), [])
        slot = next((r for r in ring if r[1].query()), None)
        if slot is None:
            slot = [torch.empty((B, Tmax), dtype=torch.int32, pin_memory=True), torch.cuda.Event()]
            if len(ring) < 8:
                ring.append(slot)
            if len(self._perm_ring) > 16:                 # ragged workloads: do not keep a ring per shape forever
                for k in list(self._perm_ring)[:8]:
                    if k != (B, Tmax):
                        del self._perm_ring[k]
        if self._arange is None or self._arange.numel() < Tmax:
            self._arange = torch.arange(max(Tmax, 4096), dtype=torch.int32)
        slot[0][:] = self._arange[:Tmax]
        return slot

    @staticmethod
    def _ship_perm(slot, dev) -> Tensor:
        out = slot[0].to(dev, non_blocking=True)
        slot[1].record(torch.cuda.current_stream(dev))
        return out

    def _draw_perm(self, ilens: List[int], Tmax: int, dev, perms=None) -> Tensor:
        """models.py:229-232: one ``torch.randperm(ilen)`` per sequence from the global CPU
        generator, in batch order; shipped as an int32 (B, Tmax) index (identity beyond ilen).
        ``perms`` (extension): use these frame orders instead of drawing (callers that batch
        chunks out of order draw them first, in the reference's order)."""
        slot = self._perm_slot(len(ilens), Tmax)
        perm = slot[0]
        if perms is None:
            rows = perm.data_ptr() + 4 * Tmax * np.arange(len(ilens), dtype=np.uint64)
            if not ops.host_randperm_rows(ilens, rows):
                for i, n in enumerate(ilens):   # int32 output draws the same numbers from the generator as the int64 default
                    torch.randperm(n, out=perm[i, :n])
        else:
            for i, n in enumerate(ilens):
                perm[i, :n] = perms[i].to(torch.int32)
        return self._ship_perm(slot, dev)

    def draw_perms(self, groups: List[List[int]], order, dev) -> List[Tensor]:
        """Frame shuffles for several batches at once, DRAWN in the order the reference's driver would draw
        them: ``groups[g]`` = sequence lengths of batch g, ``order`` = [(g, i)] in the driver's processing order
        (recording by recording, chunk by chunk, infer_eda.py:92-113).  Returns one (B_g, Tmax_g) int32 device
        index per batch."""
        slots = [self._perm_slot(len(g), max(g)) for g in groups]
        gi = np.asarray(order, dtype=np.int64).reshape(-1, 2)
        base = np.asarray([s[0].data_ptr() for s in slots], dtype=np.uint64)
        stride = np.asarray([4 * max(g) for g in groups], dtype=np.uint64)
        lens = [groups[g][i] for g, i in order]
        rows = base[gi[:, 0]] + stride[gi[:, 0]] * gi[:, 1].astype(np.uint64)
        if not ops.host_randperm_rows(lens, rows):
            for g, i in order:
                torch.randperm(groups[g][i], out=slots[g][0][i, : groups[g][i]])
        return [self._ship_perm(s, dev) for s in slots]

    def prepare_weights(self) -> None:
        """Build the kernel-side copies of the parameters now, on the current stream (they are otherwise made
        lazily by the first call; other streams wait on the event recorded here)."""
        self._prepared()
        self.eda._prepared(self.precision != L.PREC_FP32)

    def forward_embedding(self, src: List[Tensor]):
        """Reference signature (models.py:213): (encoder_out_shuffled, emb, ilens)."""
        x, ilens, _, _, B, Tmax = self._embed(src)
        emb = x.view(B, Tmax, self.n_units)
        perm = self._draw_perm(ilens, Tmax, emb.device)
        shuffled = torch.gather(emb, 1, perm.long().unsqueeze(-1).expand(-1, -1, self.n_units))
        return shuffled, emb, ilens

    # --------------------------------------------------------------------- forward -------
    def forward(self, src: List[Tensor], ts: List[Tensor], has_mask=False, activation=None):
        """models.py:236-295: (loss, stats).  ``loss`` is a 0-d CUDA tensor; it carries a ``grad_fn``
        in train() mode with gradients enabled (tensor-core precisions; one backward per forward)."""
        if self.training and torch.is_grad_enabled() and self.precision != L.PREC_FP32:
            from .training import EdaLossFn
            if self._plist is None:
                self._plist = list(self.parameters())
            loss = EdaLossFn.apply(self, list(src), list(ts), *self._plist)
            loss_att, loss_pit, counts = self._train_extras
            return loss, self._stats(loss, loss_att, loss_pit, counts, int(ts[0].shape[1]))
        x, ilens, lens_dev, seq_off, B, Tmax = self._embed(src)
        dev = x.device
        S = int(ts[0].shape[1])
        label = ops.pad_pack(self._to_device_flat(list(ts), dev), seq_off, lens_dev, B, Tmax, S, -1.0)
        label = label.view(B, Tmax, S)
        perm = self._draw_perm(ilens, Tmax, dev)
        att, probs, loss_att = self.eda.attractors(x, perm, lens_dev, ilens, B, Tmax, S + 1, n_spk_label=S,
                                                   emb_split=self._emb_split)
        pred, _, _ = ops.posterior(x, att, B, Tmax, S, want_logits=True, want_post=False)
        loss_pit, min_idx, label_perm, _ = ops.pit_loss(pred, label, lens_dev)
        counts = ops.der_counters(pred, label_perm, lens_dev)
        loss = (self.diar_weight * loss_pit + self.attractor_weight * loss_att).reshape(())
        self.last = dict(emb=x.view(B, Tmax, -1), attractors=att, probs=probs, pred=pred,
                         min_idx=min_idx, label_permute=label_perm, counts=counts)
        return loss, self._stats(loss, loss_att, loss_pit, counts, S)

    @staticmethod
    def _stats(loss, loss_att, loss_pit, counts, S):
        """models.py:274-294 from the nine DER counters."""
        c = counts.tolist()                       # the step's single device->host sync
        correct, num_frames = c[0] / S, float(c[1])
        sp_sc, sp_ms, sp_fa, spk_sc, spk_ms, spk_fa, spk_er = (float(v) for v in c[2:])
        return dict(
            loss=loss.detach(), loss_att=loss_att.reshape(()).detach(), loss_pit=loss_pit.reshape(()).detach(),
            sad_mr=sp_ms / sp_sc, sad_fr=sp_fa / sp_sc, mi=spk_ms / spk_sc, fa=spk_fa / spk_sc,
            cf=spk_er / spk_sc, acc=correct / num_frames, der=(spk_ms + spk_fa + spk_er) / spk_sc)

    # ----------------------------------------------------------------------- infer -------
    @torch.no_grad()
    def infer_device(self, feats: Tensor, ilens: List[int], max_n_speakers=15, activity_threshold=0.5, perms=None,
                     independent: bool = False, offsets=None, packed_out=None, perm_dev: Optional[Tensor] = None):
        """The device half of ``infer`` with NO host synchronisation: ``feats`` is one CUDA matrix
        holding the sequences back to back (as ``FrontEnd.features_flat`` returns it), or anywhere in it
        with ``offsets``.  Returns (posteriors (B, Tmax, max_n_speakers-1), attractor existence probs
        (B, max_n_speakers), uint8 activity labels) on the device; speaker counting / slicing is the
        caller's.  ``independent``: see ``infer``.  ``packed_out`` = (post, labels[, bit masks]) matrices with one row
        per row of ``feats``: results of frame t of sequence b land in row offsets[b] + t (the time-axis
        concatenation of infer_eda.py:121), nothing is written for padding, and (None, probs, None) is returned.
        ``perm_dev``: the (B, Tmax) int32 frame orders already on the device (else drawn / taken from ``perms``)."""
        x, ilens, lens_dev, seq_off, B, Tmax = self._embed(feats, ilens, independent, offsets)
        perm = perm_dev if perm_dev is not None else self._draw_perm(ilens, Tmax, x.device, perms)
        att, probs, _ = self.eda.attractors(x, perm, lens_dev, ilens, B, Tmax, max_n_speakers,
                                            emb_split=self._emb_split)
        if packed_out is not None:
            ops.posterior_packed(x, att, B, Tmax, max_n_speakers - 1, seq_off, lens_dev, activity_threshold,
                                 packed_out[0], packed_out[1], packed_out[2] if len(packed_out) > 2 else None)
            return None, probs, None
        _, post, labels = ops.posterior(x, att, B, Tmax, max_n_speakers - 1, threshold=activity_threshold,
                                        want_labels=True)
        return post, probs, labels

    @torch.no_grad()
    def infer(self, src: List[Tensor], infer_num_speakers=None, max_n_speakers=15, attractor_threshold=0.5,
              activity_threshold: Optional[float] = None, perms=None, independent: bool = False):
        """models.py:297-347: list of (Tmax, n_spk_b) posterior tensors.

        ``activity_threshold`` (extension): also keep the fused ``posterior > threshold``
        uint8 decisions of bin/make_rttm.py:29 in ``self.last['labels']``.
        ``independent`` (extension): every sequence is processed as if it were alone in the batch -
        attention sees only its own frames and the returned tensor has ITS length, (T_b, n_spk_b).  That is
        what the reference's driver computes (infer_eda.py:102-113 calls ``infer([chunk])`` once per chunk, so
        no padding exists there); the default reproduces ``infer`` on a padded batch, where padded frames are
        attended to (models.py:223 passes no mask) and every tensor has Tmax rows (models.py:333)."""
        x, ilens, lens_dev, _, B, Tmax = self._embed(src, independent=independent)
        perm = self._draw_perm(ilens, Tmax, x.device, perms)
        att, probs, _ = self.eda.attractors(x, perm, lens_dev, ilens, B, Tmax, max_n_speakers,
                                            emb_split=self._emb_split)
        n_att = max_n_speakers - 1
        _, post, labels = ops.posterior(x, att, B, Tmax, n_att,
                                        threshold=0.5 if activity_threshold is None else activity_threshold,
                                        want_labels=activity_threshold is not None)
        self.last = dict(emb=x.view(B, Tmax, -1), attractors=att, probs=probs, posteriors=post, labels=labels)
        rows = (lambda b: ilens[b]) if independent else (lambda b: Tmax)
        ys_active = []
        if infer_num_speakers is not None:
            order = torch.sort(probs, dim=1, descending=True, stable=False)[1][:, :infer_num_speakers]
            if bool((order >= n_att).any()):      # the reference indexes 14 columns with it (models.py:339)
                raise IndexError(f"index {n_att} is out of bounds for dimension 1 with size {n_att}")
            for b in range(B):
                ys_active.append(post[b][: rows(b), order[b]])
        elif attractor_threshold is not None:
            p_host = probs.cpu().numpy()          # the call's single device->host sync
            for b in range(B):
                silence = np.where(p_host[b] < attractor_threshold)[0]
                n_spk = int(silence[0]) if silence.size else None
                ys_active.append(post[b][: rows(b), :n_spk])
        return ys_active

    # ---------------------------------------------------------------- loss pieces --------
    def attractor_loss(self, att_prob, label: Tensor):
        """models.py:350-371 (list of (S+1,) probabilities, label (B,T,S))."""
        assert isinstance(label, Tensor), f"label: {label}"
        probs = torch.stack(list(att_prob)).float().contiguous().to(self._device())
        B, n_dec = probs.shape
        # the head kernel owns this formula; feed it logits whose sigmoid is `probs`
        att = torch.logit(probs.double()).float().reshape(B, n_dec, 1).contiguous()
        one = torch.ones(1, device=att.device)
        _, loss = ops.attractor_head(att, one, torch.zeros(1, device=att.device), int(label.size(2)))
        return loss.reshape(())

    def pit_loss(self, pred: Tensor, label: Tensor, lengths: List[int]):
        """models.py:389-413: (loss, min_idx, permute_list, label_permute[CPU float32])."""
        assert isinstance(lengths, List), f"length: {lengths}"
        dev = self._device()
        S = label.size(2)
        permute_list = [np.array(p) for p in permutations(range(S))]
        lens_dev = torch.tensor([int(v) for v in lengths], dtype=torch.int32, device=dev)
        loss, min_idx, label_perm, _ = ops.pit_loss(pred.detach().float().contiguous().to(dev),
                                                    label.detach().float().contiguous().to(dev), lens_dev)
        return loss.reshape(()), min_idx, permute_list, label_perm.cpu()

    def pit_loss_backward(self, pred: Tensor, label_permute: Tensor, lengths: List[int], grad_scale=1.0):
        """d pit_loss / d pred for the winning permutation."""
        dev = self._device()
        lens_dev = torch.tensor([int(v) for v in lengths], dtype=torch.int32, device=dev)
        return ops.pit_loss_bwd(pred.detach().float().contiguous().to(dev),
                                label_permute.float().contiguous().to(dev), lens_dev, grad_scale)

    def calc_diarization_error(self, pred: Tensor, label: Tensor, length: List[int]):
        """models.py:423-463: the 9-tuple of Python floats."""
        dev = self._device()
        lens_dev = torch.tensor([int(v) for v in length], dtype=torch.int32, device=dev)
        c = ops.der_counters(pred.detach().float().contiguous().to(dev),
                             label.detach().float().contiguous().to(dev), lens_dev).tolist()
        return (c[0] / label.size(2), float(c[1]), *(float(v) for v in c[2:]))


class EendEdaModel(TransformerEdaModel):
    """Drop-in for the reference's second model class (models.py:465-700): ``state_dict`` names
    ``linear`` / ``linear_norm`` / ``encoder`` (models.py:487,488,497-501), default PyTorch
    initialisation of the input projection (no ``init_weights``), ``infer`` keeping the FIRST
    ``infer_num_speakers`` columns (models.py:644) instead of the most probable ones.

    ``encoder_type="transformer"``: the arithmetic of ``TransformerEdaModel`` (training included).
    ``encoder_type="conformer"``: ``torchaudio.models.Conformer`` layers with the
    sequence lengths as key_padding_mask (models.py:525-528) on the tensor-core kernels; inference /
    eval-mode ``forward`` only (BatchNorm uses its running statistics), precision 'bf16x3' or 'bf16'."""

    NAMES = dict(in_linear="linear", in_norm="linear_norm", stack="encoder")

    def __init__(self, n_speakers, in_size, n_heads, n_units, n_layers, dim_feedforward=2048, dropout=0.5,
                 diar_weight: float = 1.0, attractor_weight: float = 1.0, encoder_type="transformer", eda_type="lstm",
                 precision: Optional[str] = None):
        if encoder_type not in ("transformer", "conformer"):
            raise NotImplementedError(f"encoder_type not support {encoder_type}!!!")     # models.py:503
        if eda_type != "lstm":
            raise NotImplementedError(f"eda_type not support {eda_type}!!!")             # models.py:509
        object.__setattr__(self, "encoder_type", encoder_type)       # _build_modules runs inside super().__init__
        super().__init__(n_speakers, in_size, n_heads, n_units, n_layers, dim_feedforward, dropout, False,
                         diar_weight, attractor_weight, precision)
        self.eda_type = eda_type

    def _build_modules(self, in_size, n_units, n_heads, n_layers, dim_feedforward, dropout):
        self.linear = nn.Linear(in_size, n_units)
        self.linear_norm = nn.LayerNorm(n_units)
        if self.encoder_type == "transformer":
            layer = TransformerEncoderLayer(n_units, n_heads, dim_feedforward, dropout, batch_first=True)
            self.encoder = TransformerEncoder(layer, n_layers, enable_nested_tensor=False)
        else:
            from torchaudio.models import Conformer      # parameter container only (models.py:501)
            self.encoder = Conformer(input_dim=n_units, num_heads=n_heads, ffn_dim=dim_feedforward,
                                     num_layers=n_layers, depthwise_conv_kernel_size=31)

    # ------------------------------------------------------------- Conformer encoder --------
    def _prepared(self):
        if self.encoder_type == "transformer":
            return super()._prepared()
        if self._plist is None:
            self._plist = list(self.parameters()) + list(self.buffers())
        key = tuple((p.data_ptr(), p._version) for p in self._plist)
        if self._prep is None or self._prep[0] != key:
            g = lambda t: t.detach().float().contiguous()
            sp = lambda t: ops.split_bf16(g(t))
            kpad = _round_up(self.in_size, 16)
            w = g(self.linear.weight)
            w_in = torch.zeros((self.n_units, kpad), dtype=torch.float32, device=w.device)
            w_in[:, : self.in_size] = w
            D = self.n_units
            layers = []
            for lyr in self.encoder.conformer_layers:
                d = {}
                for name, ffn in (("f1", lyr.ffn1), ("f2", lyr.ffn2)):
                    q = ffn.sequential
                    d[name] = dict(g=g(q[0].weight), be=g(q[0].bias), w1=sp(q[1].weight), b1=g(q[1].bias),
                                   w2=sp(q[4].weight), b2=g(q[4].bias))
                d["att"] = dict(g=g(lyr.self_attn_layer_norm.weight), be=g(lyr.self_attn_layer_norm.bias),
                                w_qkv=sp(lyr.self_attn.in_proj_weight), b_qkv=g(lyr.self_attn.in_proj_bias),
                                w_o=sp(lyr.self_attn.out_proj.weight), b_o=g(lyr.self_attn.out_proj.bias))
                cm = lyr.conv_module
                q = cm.sequential
                # GLU pairs channel c with c + D: interleave the pointwise rows in blocks of 32 (a-block, gate-block)
                # so that a pair lands in one thread of the GEMM epilogue (EDA_EPI_BIAS_GLU)
                idx = torch.arange(D, device=w.device).view(D // 32, 32)
                order = torch.stack([idx, idx + D], dim=1).reshape(-1)
                bn = q[3]
                scale = g(bn.weight) / torch.sqrt(g(bn.running_var) + bn.eps)
                d["conv"] = dict(g=g(cm.layer_norm.weight), be=g(cm.layer_norm.bias),
                                 w_pw1=sp(g(q[0].weight).view(2 * D, D)[order]), b_pw1=g(q[0].bias)[order].contiguous(),
                                 w_dw=g(q[2].weight).view(D, -1).contiguous(), scale=scale.contiguous(),
                                 shift=((g(q[2].bias) - g(bn.running_mean)) * scale + g(bn.bias)).contiguous(),
                                 w_pw2=sp(g(q[5].weight).view(D, D)), b_pw2=g(q[5].bias))
                d["fin"] = dict(g=g(lyr.final_layer_norm.weight), be=g(lyr.final_layer_norm.bias))
                layers.append(d)
            prep = dict(kpad=kpad, ws_in=ops.split_bf16(w_in), b_in=g(self.linear.bias), g_in=g(self.linear_norm.weight),
                        be_in=g(self.linear_norm.bias), layers=layers)
            self._prep = (key, prep)
            self._prep_event = self._record_event(w.device)
        self._wait_event(self._prep_event)
        return self._prep[1]

    def _embed(self, src, ilens: Optional[List[int]] = None, independent: bool = False, offsets=None):
        if self.encoder_type == "transformer":
            return super()._embed(src, ilens, independent, offsets)
        if offsets is not None or (independent and ilens is not None and min(ilens) != max(ilens)):
            raise NotImplementedError("the Conformer variant batches like the reference (key_padding_mask, convolution over "
                                      "the padded time axis, models.py:525-528): independent / scattered sequences are a "
                                      "transformer-path feature")
        dev = self._device()
        if self.training:
            raise NotImplementedError("the Conformer variant runs in eval() mode only (no backward kernels, BatchNorm "
                                      "uses its running statistics)")
        P = self.precision
        if P == L.PREC_FP32:
            raise NotImplementedError("the Conformer variant runs on the tensor-core pipeline: precision 'bf16x3' or 'bf16'")
        w = self._prepared()
        flat_in = ilens is not None
        if not flat_in:
            ilens = [int(x.shape[0]) for x in src]
        B, Tmax = len(ilens), max(ilens)
        off = np.zeros(B + 1, dtype=np.int64)
        np.cumsum(ilens, out=off[1:])
        meta = torch.from_numpy(np.concatenate([off, np.asarray(ilens, dtype=np.int64)])).to(dev, non_blocking=True)
        seq_off, lens_dev = meta[: B + 1], meta[B + 1:].to(torch.int32)
        flat = src if flat_in else self._to_device_flat(src, dev)
        M, D, H, FF = B * Tmax, self.n_units, self.n_heads, self.dim_feedforward
        key = ("conformer", B, Tmax, str(dev), self.workspace_slot)
        ws = self._ws.get(key)
        if ws is None:
            old = [k for k in self._ws if k[0] == "conformer"]
            for k in old[: max(0, len(old) - 3)]:       # at most 4 shapes stay resident (oldest go first)
                del self._ws[k]
            Tp = _round_up(Tmax, 8)
            f32 = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
            ws = dict(Tp=Tp, xin=ops.Split.empty(M, w["kpad"], dev), xa=f32(M, D), xb=f32(M, D), glu=f32(M, D),
                      sa=ops.Split.empty(M, D, dev), sb=ops.Split.empty(M, D, dev), ff=ops.Split.empty(M, FF, dev),
                      qk=ops.Split.empty(M, 2 * D, dev), vt=ops.Split.empty(B * H * (D // H), Tp, dev),
                      att=ops.Split.empty(M, D, dev), emb_s=ops.Split.empty(M, D, dev), xproj=f32(M, 4 * D))
            self._ws[key] = ws
        Tp, xa, xb, sa, sb, ff = ws["Tp"], ws["xa"], ws["xb"], ws["sa"], ws["sb"], ws["ff"]
        ops.pad_pack_split(flat, seq_off, lens_dev, B, Tmax, ws["xin"], -1.0)
        ops.gemm_split(ws["xin"], w["ws_in"], w["b_in"], M, D, ws["xin"].ld, L.EPI_BIAS_LN, P, gamma=w["g_in"],
                       beta=w["be_in"], out=xa)
        emb = torch.empty((M, D), dtype=torch.float32, device=dev)
        n_layers = len(w["layers"])
        for i, lw in enumerate(w["layers"]):
            f1, at, cv, f2, fin = lw["f1"], lw["att"], lw["conv"], lw["f2"], lw["fin"]
            # x = 0.5 * ffn1(x) + x
            ops.layernorm_split(xa, f1["g"], f1["be"], out_split=sa)
            ops.gemm_split_ex(sa, f1["w1"], f1["b1"], M, FF, D, L.EPI_BIAS_SILU, P, out_split=ff)
            ops.gemm_split_ex(ff, f1["w2"], f1["b2"], M, D, FF, L.EPI_BIAS_RES, P, resid=xa, out=xb, out_scale=0.5)
            # x = self_attn(LN(x), key_padding_mask) + x
            ops.layernorm_split(xb, at["g"], at["be"], out_split=sa)
            ops.gemm_qkv(sa, at["w_qkv"], at["b_qkv"], ws["qk"], ws["vt"], B, Tmax, Tp, D, H, P)
            ops.mha_split_masked(ws["qk"], ws["vt"], B, Tmax, Tp, H, lens_dev, self.attn_precision, ws["att"])
            ops.gemm_split_ex(ws["att"], at["w_o"], at["b_o"], M, D, D, L.EPI_BIAS_RES, P, resid=xb, out=xa)
            # x = x + conv_module(x)
            ops.layernorm_split(xa, cv["g"], cv["be"], out_split=sa)
            ops.gemm_split_ex(sa, cv["w_pw1"], cv["b_pw1"], M, 2 * D, D, L.EPI_BIAS_GLU, P, out=ws["glu"])
            ops.dwconv_bn_silu(ws["glu"], cv["w_dw"], cv["scale"], cv["shift"], B, Tmax, sb)
            ops.gemm_split_ex(sb, cv["w_pw2"], cv["b_pw2"], M, D, D, L.EPI_BIAS_RES, P, resid=xa, out=xb)
            # x = final_layer_norm(0.5 * ffn2(x) + x)
            ops.layernorm_split(xb, f2["g"], f2["be"], out_split=sa)
            ops.gemm_split_ex(sa, f2["w1"], f2["b1"], M, FF, D, L.EPI_BIAS_SILU, P, out_split=ff)
            last = i == n_layers - 1
            ops.gemm_split_ex(ff, f2["w2"], f2["b2"], M, D, FF, L.EPI_BIAS_RES_LN, P, resid=xb, gamma=fin["g"],
                              beta=fin["be"], out=emb if last else xa, out_split=ws["emb_s"] if last else None,
                              out_scale=0.5)
        if n_layers == 0:
            emb.copy_(xa)
            ops.split_bf16(emb, out=ws["emb_s"])
        self._emb_split = (ws["emb_s"], ws["xproj"])
        return emb, ilens, lens_dev, seq_off, B, Tmax

    def forward(self, src: List[Tensor], ts: List[Tensor], has_mask=False, activation=None):
        """models.py:563-599."""
        if self.encoder_type == "conformer" and self.training:
            raise NotImplementedError("the Conformer variant has no backward kernels: call .eval() (as the reference's dev "
                                      "loop does, train_eda.py:205)")
        return super().forward(src, ts, has_mask, activation)

    @torch.no_grad()
    def infer(self, src: List[Tensor], infer_num_speakers=None, max_n_speakers=15, attractor_threshold=0.5,
              activity_threshold: Optional[float] = None, perms=None, independent: bool = False):
        """models.py:601-652."""
        if infer_num_speakers is None:
            return super().infer(src, None, max_n_speakers, attractor_threshold, activity_threshold, perms, independent)
        ys = super().infer(src, None, max_n_speakers, -1.0, activity_threshold, perms, independent)  # all columns, one pass
        return [y[:, :infer_num_speakers] for y in ys]

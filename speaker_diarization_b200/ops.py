"""Tensor-level wrappers over the C ABI (one Python function per entry point).

Only unpacking happens here: shape bookkeeping, output allocation through torch's caching
allocator, and the ctypes call on torch's current CUDA stream.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib as L

Tensor = torch.Tensor


def _f32(t: Tensor) -> Tensor:
    assert t.dtype == torch.float32 and t.is_cuda and t.is_contiguous(), (t.dtype, t.device, t.is_contiguous())
    return t


_MT_STATE_BYTES = 5056      # torch's serialised CPU generator (32-bit Mersenne Twister): seed u64 | left i32 | seeded i32 | next u64 | state u64[624] | ...


def host_randperm_rows(lens, row_ptrs) -> bool:
    """``torch.randperm(n)`` for every n in ``lens`` (in order) from torch's GLOBAL CPU generator, written as int32 to
    the host addresses ``row_ptrs`` (numpy uint64 array) by ONE native call that advances the generator exactly like the
    individual calls would (models.py:229-232 draws one per sequence; 64 dispatcher round trips per batch were the
    largest item of the host path).  Returns False (nothing drawn) if the generator state has an unknown layout."""
    import numpy as np
    st = torch.get_rng_state()
    if st.numel() != _MT_STATE_BYTES:
        return False
    a = st.numpy()
    left, nxt, state = a[8:12].view(np.int32), a[16:24].view(np.uint64), a[24:24 + 624 * 8].view(np.uint64)
    if not (1 <= int(left[0]) <= 624 and int(nxt[0]) <= 624 and int(a[12:16].view(np.int32)[0]) == 1):
        return False
    lens_a = np.ascontiguousarray(lens, dtype=np.int32)
    rows = np.ascontiguousarray(row_ptrs, dtype=np.uint64)
    rc = L.load().eda_host_randperm_batch(left.ctypes.data, nxt.ctypes.data, state.ctypes.data, lens_a.ctypes.data,
                                          int(lens_a.size), rows.ctypes.data)
    if rc != 0:
        raise L.EdaError(f"eda_host_randperm_batch failed ({rc}): {L.load().eda_last_error_string().decode()}")
    torch.set_rng_state(st)
    return True


def pad_pack(flat: Tensor, seq_off: Tensor, lens: Tensor, B: int, Tmax: int, ld_dst: int, pad: float) -> Tensor:
    """flat (sum T_i, C) -> (B*Tmax, ld_dst) padded with `pad` (models.py:217)."""
    _f32(flat)
    out = torch.empty((B * Tmax, ld_dst), dtype=torch.float32, device=flat.device)
    L.call("eda_pad_pack", L.ptr(flat), L.ptr(seq_off), L.ptr(lens), B, Tmax, flat.shape[1], float(pad),
           L.ptr(out), ld_dst, L.stream())
    return out


def linear(a: Tensor, w: Tensor, bias: Optional[Tensor], epilogue: int = L.EPI_BIAS, prec: int = L.PREC_FP32,
           resid: Optional[Tensor] = None, gamma: Optional[Tensor] = None, beta: Optional[Tensor] = None,
           out: Optional[Tensor] = None) -> Tensor:
    """nn.Linear + fused epilogue.  a (M,K), w (N,K) -> (M,N)."""
    _f32(a), _f32(w)
    M, K = a.shape
    N = w.shape[0]
    assert w.shape[1] == K, (a.shape, w.shape)
    if out is None:
        out = torch.empty((M, N), dtype=torch.float32, device=a.device)
    L.call("eda_linear", L.ptr(a), a.stride(0), L.ptr(w), w.stride(0), L.ptr(bias),
           L.ptr(resid), resid.stride(0) if resid is not None else 0, L.ptr(gamma), L.ptr(beta),
           L.ptr(out), out.stride(0), M, N, K, epilogue, prec, L.stream())
    return out


def mha(qkv: Tensor, B: int, T: int, n_heads: int, prec: int = L.PREC_FP32, key_lens: Optional[Tensor] = None) -> Tensor:
    _f32(qkv)
    D = qkv.shape[1] // 3
    out = torch.empty((B * T, D), dtype=torch.float32, device=qkv.device)
    if key_lens is not None:
        assert key_lens.dtype == torch.int32
        L.call("eda_mha_fwd_masked", L.ptr(qkv), qkv.stride(0), L.ptr(out), out.stride(0), B, T, n_heads,
               D // n_heads, L.ptr(key_lens), prec, L.stream())
        return out
    L.call("eda_mha_fwd", L.ptr(qkv), qkv.stride(0), L.ptr(out), out.stride(0), B, T, n_heads,
           D // n_heads, prec, L.stream())
    return out


def lstm_attractors(xproj: Tensor, perm: Tensor, lens: Tensor, order: Optional[Tensor], w_hh_enc: Tensor,
                    w_hh_dec: Tensor, dec_bias: Tensor, B: int, Tmax: int, n_dec: int, want_hc: bool = False,
                    prec: int = L.PREC_FP32):
    D = w_hh_enc.shape[1]
    att = torch.empty((B, n_dec, D), dtype=torch.float32, device=xproj.device)
    hc = torch.empty((B, 2, D), dtype=torch.float32, device=xproj.device) if want_hc else None
    assert perm.dtype == torch.int32 and lens.dtype == torch.int32
    L.call("eda_lstm_attractors", L.ptr(_f32(xproj)), L.ptr(perm), L.ptr(lens), L.ptr(order),
           L.ptr(_f32(w_hh_enc)), L.ptr(_f32(w_hh_dec)), L.ptr(_f32(dec_bias)), B, Tmax, D, n_dec,
           L.ptr(att), L.ptr(hc), prec, L.stream())
    return att, hc


def attractor_head(att: Tensor, w: Tensor, bias: Tensor, n_spk_label: int = -1):
    B, n_dec, D = att.shape
    probs = torch.empty((B, n_dec), dtype=torch.float32, device=att.device)
    loss = torch.zeros((1,), dtype=torch.float32, device=att.device) if n_spk_label >= 0 else None
    L.call("eda_attractor_head", L.ptr(_f32(att)), L.ptr(_f32(w)), L.ptr(_f32(bias)), B, n_dec, D,
           L.ptr(probs), n_spk_label, L.ptr(loss), L.stream())
    return probs, loss


def posterior(emb: Tensor, att: Tensor, B: int, Tmax: int, n_att: int, threshold: float = 0.5,
              want_logits: bool = False, want_post: bool = True, want_labels: bool = False):
    """emb (B*Tmax, D), att (B, rows>=n_att, D) -> logits/post (B,Tmax,n_att) fp32, labels uint8."""
    D = emb.shape[1]
    mk = lambda dt: torch.empty((B, Tmax, n_att), dtype=dt, device=emb.device)
    logits = mk(torch.float32) if want_logits else None
    post = mk(torch.float32) if want_post else None
    labels = mk(torch.uint8) if want_labels else None
    L.call("eda_posterior", L.ptr(_f32(emb)), emb.stride(0), L.ptr(_f32(att)), att.shape[1], B, Tmax, D,
           n_att, float(threshold), L.ptr(logits), L.ptr(post), L.ptr(labels), L.stream())
    return logits, post, labels


def posterior_packed(emb: Tensor, att: Tensor, B: int, Tmax: int, n_att: int, out_off: Tensor, lens: Tensor,
                     threshold: float, post: Optional[Tensor], labels: Optional[Tensor], mask16: Optional[Tensor] = None) -> None:
    """eda_posterior into (total rows, n_att) matrices: frame t of sequence b -> row out_off[b] + t; ``mask16`` (total rows,)
    int16 = the thresholded decisions of a frame as bits."""
    D = emb.shape[1]
    assert out_off.dtype == torch.int64 and lens.dtype == torch.int32
    assert post is None or (post.dtype == torch.float32 and post.is_contiguous() and post.shape[1] == n_att)
    assert labels is None or (labels.dtype == torch.uint8 and labels.is_contiguous() and labels.shape[1] == n_att)
    L.call("eda_posterior_packed", L.ptr(_f32(emb)), emb.stride(0), L.ptr(_f32(att)), att.shape[1], B, Tmax, D,
           n_att, float(threshold), L.ptr(out_off), L.ptr(lens), None, L.ptr(post), L.ptr(labels), L.ptr(mask16), L.stream())


def pit_loss(pred: Tensor, label: Tensor, lens: Tensor):
    B, T, S = pred.shape
    dev = pred.device
    loss = torch.empty((1,), dtype=torch.float32, device=dev)
    min_idx = torch.empty((B,), dtype=torch.int64, device=dev)
    label_perm = torch.empty((B, T, S), dtype=torch.float32, device=dev)
    perm = torch.empty((B, S), dtype=torch.int32, device=dev)
    L.call("eda_pit_loss", L.ptr(_f32(pred)), L.ptr(_f32(label)), L.ptr(lens), B, T, S, L.ptr(loss),
           L.ptr(min_idx), L.ptr(label_perm), L.ptr(perm), L.stream())
    return loss, min_idx, label_perm, perm


def pit_loss_bwd(pred: Tensor, label_perm: Tensor, lens: Tensor, grad_scale: float = 1.0) -> Tensor:
    B, T, S = pred.shape
    dpred = torch.empty_like(pred)
    L.call("eda_pit_loss_bwd", L.ptr(_f32(pred)), L.ptr(_f32(label_perm)), L.ptr(lens), B, T, S,
           float(grad_scale), L.ptr(dpred), L.stream())
    return dpred


def der_counters(pred: Tensor, label_perm: Tensor, lens: Tensor) -> Tensor:
    B, T, S = pred.shape
    counts = torch.empty((9,), dtype=torch.int64, device=pred.device)
    L.call("eda_der_counters", L.ptr(_f32(pred)), L.ptr(_f32(label_perm)), L.ptr(lens), B, T, S,
           L.ptr(counts), L.stream())
    return counts


def rttm_segments(src: Tensor, rec_off: Tensor, n_spk: int, threshold: float = 0.5, median: int = 1,
                  cap: Optional[int] = None, n_spk_rec: Optional[Tensor] = None):
    """bin/make_rttm.py:29-35 on the device.  ``src``: (rows, ld >= n_spk) fp32 posteriors or uint8
    0/1 decisions, recordings stacked row-wise, ``rec_off`` int64 (n_rec + 1) row offsets.
    Returns (segments (n_rec, n_spk, cap, 2) int32, counts (n_rec, n_spk) int32)."""
    assert src.dim() == 2 and src.stride(1) == 1 and src.dtype in (torch.float32, torch.uint8)
    n_rec = rec_off.numel() - 1
    if cap is None:
        cap = src.shape[0] // 2 + 1          # worst case: alternating frames
    seg = torch.empty((n_rec, n_spk, cap, 2), dtype=torch.int32, device=src.device)
    counts = torch.empty((n_rec, n_spk), dtype=torch.int32, device=src.device)
    is_post = src.dtype == torch.float32
    L.call("eda_rttm_segments", L.ptr(src) if is_post else None, None if is_post else L.ptr(src), src.stride(0),
           L.ptr(rec_off), n_rec, L.ptr(n_spk_rec) if n_spk_rec is not None else None, n_spk, float(threshold),
           int(median), int(cap), L.ptr(seg), L.ptr(counts), L.stream())
    return seg, counts


# ------------------------------------------------------- split-bf16 tensor-core pipeline ----
class Split:
    """A matrix held as a split-bf16 pair (x = hi + lo): two (rows, ld) bf16 CUDA tensors."""
    __slots__ = ("hi", "lo")

    def __init__(self, hi: Tensor, lo: Tensor):
        assert hi.dtype == torch.bfloat16 and lo.dtype == torch.bfloat16 and hi.shape == lo.shape
        assert hi.is_contiguous() and lo.is_contiguous() and hi.is_cuda
        self.hi, self.lo = hi, lo

    @classmethod
    def empty(cls, rows: int, cols: int, device) -> "Split":
        buf = torch.empty((2, rows, cols), dtype=torch.bfloat16, device=device)
        return cls(buf[0], buf[1])

    @property
    def rows(self) -> int:
        return self.hi.shape[0]

    @property
    def ld(self) -> int:
        return self.hi.shape[1]

    def float(self) -> Tensor:
        return self.hi.float() + self.lo.float()


def split_bf16(x: Tensor, ldk: int = 0, out: Optional[Split] = None) -> Split:
    """fp32 (rows, K) -> Split (rows, ldk), zero padded; ldk defaults to K rounded up to 8."""
    _f32(x)
    rows, K = x.shape
    ldk = ldk or (K + 7) // 8 * 8
    out = out or Split.empty(rows, ldk, x.device)
    L.call("eda_split_bf16", L.ptr(x), x.stride(0), rows, K, L.ptr(out.hi), L.ptr(out.lo), ldk, L.stream())
    return out


def pad_pack_split(flat: Tensor, seq_off: Tensor, lens: Tensor, B: int, Tmax: int, out: Split, pad: float) -> Split:
    """flat (sum T_i, C) [row stride may exceed C] -> Split (B*Tmax, out.ld) padded with `pad`."""
    assert flat.dtype == torch.float32 and flat.is_cuda and flat.stride(1) == 1
    L.call("eda_pad_pack_split", flat.data_ptr(), flat.stride(0), L.ptr(seq_off), L.ptr(lens), B, Tmax,
           flat.shape[1], float(pad), L.ptr(out.hi), L.ptr(out.lo), out.ld, L.stream())
    return out


def gemm_split(a: Split, w: Split, bias: Optional[Tensor], M: int, N: int, K: int, epilogue: int, prec: int,
               resid: Optional[Tensor] = None, gamma: Optional[Tensor] = None, beta: Optional[Tensor] = None,
               out: Optional[Tensor] = None, out_split: Optional[Split] = None) -> None:
    """nn.Linear + fused epilogue on split operands; writes fp32 `out` and/or `out_split`."""
    L.call("eda_gemm_split", L.ptr(a.hi), L.ptr(a.lo), a.ld, L.ptr(w.hi), L.ptr(w.lo), w.ld, L.ptr(bias),
           L.ptr(resid), resid.stride(0) if resid is not None else 0, L.ptr(gamma), L.ptr(beta),
           L.ptr(out), out.stride(0) if out is not None else 0,
           L.ptr(out_split.hi) if out_split else None, L.ptr(out_split.lo) if out_split else None,
           out_split.ld if out_split else 0, M, N, K, epilogue, prec, L.stream())


def gemm_qkv(a: Split, w: Split, bias: Tensor, qk: Split, vt: Split, B: int, T: int, Tp: int, D: int,
             n_heads: int, prec: int) -> None:
    L.call("eda_gemm_qkv", L.ptr(a.hi), L.ptr(a.lo), a.ld, L.ptr(w.hi), L.ptr(w.lo), w.ld, L.ptr(bias),
           L.ptr(qk.hi), L.ptr(qk.lo), qk.ld, L.ptr(vt.hi), L.ptr(vt.lo), B, T, Tp, D, n_heads, prec, L.stream())


def mha_split(qk: Split, vt: Split, B: int, T: int, Tp: int, n_heads: int, prec: int,
              out: Optional[Tensor] = None, out_split: Optional[Split] = None) -> None:
    L.call("eda_mha_split", L.ptr(qk.hi), L.ptr(qk.lo), qk.ld, L.ptr(vt.hi), L.ptr(vt.lo), Tp, B, T, n_heads,
           L.ptr(out), out.stride(0) if out is not None else 0,
           L.ptr(out_split.hi) if out_split else None, L.ptr(out_split.lo) if out_split else None,
           out_split.ld if out_split else 0, prec, L.stream())


# ------------------------------------------------------------------ Conformer variant --------
def gemm_split_ex(a: Split, w: Split, bias: Optional[Tensor], M: int, N: int, K: int, epilogue: int, prec: int,
                  resid: Optional[Tensor] = None, gamma: Optional[Tensor] = None, beta: Optional[Tensor] = None,
                  out: Optional[Tensor] = None, out_split: Optional[Split] = None, out_scale: float = 1.0) -> None:
    """gemm_split + the SiLU / GLU / residual epilogues and a scale on the GEMM term of the residual ones."""
    L.call("eda_gemm_split_ex", L.ptr(a.hi), L.ptr(a.lo), a.ld, L.ptr(w.hi), L.ptr(w.lo), w.ld, L.ptr(bias),
           L.ptr(resid), resid.stride(0) if resid is not None else 0, L.ptr(gamma), L.ptr(beta),
           L.ptr(out), out.stride(0) if out is not None else 0,
           L.ptr(out_split.hi) if out_split else None, L.ptr(out_split.lo) if out_split else None,
           out_split.ld if out_split else 0, float(out_scale), M, N, K, epilogue, prec, L.stream())


def layernorm_split(x: Tensor, gamma: Tensor, beta: Tensor, out_split: Optional[Split] = None,
                    out: Optional[Tensor] = None) -> None:
    M, D = x.shape
    L.call("eda_layernorm_split", L.ptr(_f32(x)), x.stride(0), L.ptr(gamma), L.ptr(beta), M, D,
           L.ptr(out), out.stride(0) if out is not None else 0,
           L.ptr(out_split.hi) if out_split else None, L.ptr(out_split.lo) if out_split else None,
           out_split.ld if out_split else 0, L.stream())


def dwconv_bn_silu(x: Tensor, w: Tensor, scale: Tensor, shift: Tensor, B: int, T: int, out_split: Split) -> None:
    """x (B*T, 256) fp32 -> silu(scale * depthwise_conv31(x) + shift) as a split pair."""
    C, K = w.shape
    L.call("eda_dwconv_bn_silu", L.ptr(_f32(x)), x.stride(0), L.ptr(_f32(w)), L.ptr(scale), L.ptr(shift), B, T, C, K,
           L.ptr(out_split.hi), L.ptr(out_split.lo), out_split.ld, L.stream())


def mha_split_masked(qk: Split, vt: Split, B: int, T: int, Tp: int, n_heads: int, key_lens: Tensor, prec: int,
                     out_split: Split) -> None:
    assert key_lens.dtype == torch.int32
    L.call("eda_mha_split_masked", L.ptr(qk.hi), L.ptr(qk.lo), qk.ld, L.ptr(vt.hi), L.ptr(vt.lo), Tp, B, T, n_heads,
           L.ptr(key_lens), L.ptr(out_split.hi), L.ptr(out_split.lo), out_split.ld, prec, L.stream())

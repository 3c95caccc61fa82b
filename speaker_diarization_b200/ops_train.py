"""Tensor-level wrappers over the training-step entry points of the C ABI (include/eda_b200.h,
"training step" section).  Like ``ops``: unpacking only, no arithmetic on the Python side."""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib as L
from .ops import Split, _f32

Tensor = torch.Tensor


def _sp(s: Optional[Split]):
    return (L.ptr(s.hi), L.ptr(s.lo), s.ld) if s is not None else (None, None, 0)


NO_DROP = (0.0, 0, 0)      # (p, seed, site): dropout off


def gemm_split_train(a: Split, w: Split, bias: Tensor, M: int, N: int, K: int, epilogue: int, prec: int,
                     rstd: Optional[Tensor], resid: Optional[Tensor] = None, gamma: Optional[Tensor] = None,
                     beta: Optional[Tensor] = None, out: Optional[Tensor] = None, out_split: Optional[Split] = None,
                     drop=NO_DROP):
    """``drop`` = (p, seed, site): dropout on relu(..) (EPI_BIAS_RELU) or on A.W^T + bias before the residual."""
    oh, ol, old = _sp(out_split)
    L.call("eda_gemm_split_train", L.ptr(a.hi), L.ptr(a.lo), a.ld, L.ptr(w.hi), L.ptr(w.lo), w.ld, L.ptr(bias),
           L.ptr(resid), resid.stride(0) if resid is not None else 0, L.ptr(gamma), L.ptr(beta),
           L.ptr(out), out.stride(0) if out is not None else 0, oh, ol, old, L.ptr(rstd), float(drop[0]), int(drop[1]),
           int(drop[2]), M, N, K, epilogue, prec, L.stream())


def gemm_dgrad(dy: Split, wt: Split, M: int, K_in: int, N_out: int, prec: int, resid: Optional[Tensor] = None,
               relu_mask: Optional[Split] = None, out: Optional[Tensor] = None, out_split: Optional[Split] = None,
               mask_scale: float = 1.0):
    """dX (M, K_in) = dY (M, N_out) . W with wt = W^T (K_in, N_out) split; + resid or ReLU mask."""
    oh, ol, old = _sp(out_split)
    L.call("eda_gemm_dgrad", L.ptr(dy.hi), L.ptr(dy.lo), dy.ld, L.ptr(wt.hi), L.ptr(wt.lo), wt.ld,
           L.ptr(resid), resid.stride(0) if resid is not None else 0,
           L.ptr(relu_mask.hi) if relu_mask is not None else None, relu_mask.ld if relu_mask is not None else 0,
           float(mask_scale), L.ptr(out), out.stride(0) if out is not None else 0, oh, ol, old, M, K_in, N_out, prec, L.stream())


def gemm_wgrad(dy: Split, x: Split, M: int, N_out: int, K_in: int, dW: Tensor, work: Tensor, prec: int,
               accumulate: bool = False):
    """dW (N_out, K_in) (+)= dY^T X; ``dW`` is a 2-D fp32 view (row stride may exceed K_in)."""
    assert dW.dtype == torch.float32 and dW.is_cuda and dW.stride(1) == 1 and dW.shape == (N_out, K_in)
    L.call("eda_gemm_wgrad", L.ptr(dy.hi), L.ptr(dy.lo), dy.ld, L.ptr(x.hi), L.ptr(x.lo), x.ld, M, N_out, K_in,
           dW.data_ptr(), dW.stride(0), L.ptr(work), work.numel() * work.element_size(), int(accumulate), prec,
           L.stream())


def ln_bwd(dy: Tensor, y: Tensor, gamma: Tensor, beta: Tensor, rstd: Tensor, work: Tensor, dgamma: Tensor,
           dbeta: Tensor, dcolsum: Optional[Tensor] = None, dx: Optional[Tensor] = None,
           dx_split: Optional[Split] = None, accumulate: bool = False, drop=NO_DROP):
    M, D = dy.shape
    oh, ol, old = _sp(dx_split)
    L.call("eda_ln_bwd", L.ptr(_f32(dy)), dy.stride(0), L.ptr(_f32(y)), y.stride(0), L.ptr(gamma), L.ptr(beta),
           L.ptr(rstd), M, D, L.ptr(dx), dx.stride(0) if dx is not None else 0, oh, ol, old, L.ptr(dgamma),
           L.ptr(dbeta), L.ptr(dcolsum), L.ptr(work), work.numel() * work.element_size(), int(accumulate),
           float(drop[0]), int(drop[1]), int(drop[2]), L.stream())


def colsum_split(x: Split, M: int, N: int, out: Tensor, work: Tensor, accumulate: bool = False):
    L.call("eda_colsum_split", L.ptr(x.hi), L.ptr(x.lo), M, N, x.ld, L.ptr(out), L.ptr(work),
           work.numel() * work.element_size(), int(accumulate), L.stream())


def mha_split_train(qk: Split, vt: Split, B: int, T: int, Tp: int, n_heads: int, prec: int, out_split: Split,
                    lse: Tensor, drop=NO_DROP, drop_bits: Optional[Tensor] = None):
    """``drop_bits``: (B*n_heads*T, ld) int32 keep bits of this site from ``dropout_bits`` (None = evaluated in the kernel)."""
    L.call("eda_mha_split_train", L.ptr(qk.hi), L.ptr(qk.lo), qk.ld, L.ptr(vt.hi), L.ptr(vt.lo), Tp, B, T, n_heads,
           L.ptr(out_split.hi), L.ptr(out_split.lo), out_split.ld, L.ptr(lse), float(drop[0]), int(drop[1]),
           int(drop[2]), L.ptr(drop_bits), drop_bits.stride(0) if drop_bits is not None else 0, prec, L.stream())


def mha_delta(do: Split, o: Split, M: int, D: int, delta: Tensor):
    L.call("eda_mha_delta", L.ptr(do.hi), L.ptr(do.lo), do.ld, L.ptr(o.hi), L.ptr(o.lo), o.ld, M, D, L.ptr(delta),
           L.stream())


def mha_bwd(qk: Split, vt: Split, Tp: int, do: Split, lse: Tensor, delta: Tensor, B: int, T: int, n_heads: int,
            dq_acc: Tensor, dqkv: Split, prec: int, drop=NO_DROP, drop_bits: Optional[Tensor] = None):
    L.call("eda_mha_bwd", L.ptr(qk.hi), L.ptr(qk.lo), qk.ld, L.ptr(vt.hi), L.ptr(vt.lo), Tp, L.ptr(do.hi),
           L.ptr(do.lo), do.ld, L.ptr(lse), L.ptr(delta), B, T, n_heads, L.ptr(dq_acc), L.ptr(dqkv.hi),
           L.ptr(dqkv.lo), dqkv.ld, float(drop[0]), int(drop[1]), int(drop[2]), L.ptr(drop_bits),
           drop_bits.stride(0) if drop_bits is not None else 0, prec, L.stream())


def scale_split(src: Tensor, scale: float, dst: Split, cols: Optional[int] = None):
    rows = src.shape[0]
    L.call("eda_scale_split", L.ptr(_f32(src)), src.stride(0), rows, cols or src.shape[1], float(scale),
           L.ptr(dst.hi), L.ptr(dst.lo), dst.ld, L.stream())


def lstm_attractors_train(xproj: Tensor, perm: Tensor, lens: Tensor, order: Optional[Tensor], w_hh_enc: Tensor,
                          w_hh_dec: Tensor, dec_bias: Tensor, B: int, Tmax: int, n_dec: int, sv: dict,
                          prec: int = L.PREC_FP32) -> Tensor:
    """``sv``: the save buffers (gates, c, hprev: Split, dec_gates, dec_c, dec_hprev: Split, hc)."""
    D = w_hh_enc.shape[1]
    att = torch.empty((B, n_dec, D), dtype=torch.float32, device=xproj.device)
    L.call("eda_lstm_attractors_train", L.ptr(_f32(xproj)), L.ptr(perm), L.ptr(lens), L.ptr(order),
           L.ptr(_f32(w_hh_enc)), L.ptr(_f32(w_hh_dec)), L.ptr(_f32(dec_bias)), B, Tmax, D, n_dec, L.ptr(att),
           L.ptr(sv["hc"]), L.ptr(sv["gates"]), L.ptr(sv["c"]), L.ptr(sv["hprev"].hi), L.ptr(sv["hprev"].lo),
           L.ptr(sv["dec_gates"]), L.ptr(sv["dec_c"]), L.ptr(sv["dec_hprev"].hi), L.ptr(sv["dec_hprev"].lo),
           prec, L.stream())
    return att


def lstm_attractors_bwd(perm: Tensor, lens: Tensor, order: Optional[Tensor], w_hh_enc: Tensor, w_hh_dec: Tensor,
                        sv: dict, datt: Tensor, B: int, Tmax: int, n_dec: int, dxproj: Split, ddec: Split):
    D = w_hh_enc.shape[1]
    L.call("eda_lstm_attractors_bwd", L.ptr(perm), L.ptr(lens), L.ptr(order), L.ptr(_f32(w_hh_enc)),
           L.ptr(_f32(w_hh_dec)), L.ptr(sv["gates"]), L.ptr(sv["c"]), L.ptr(sv["dec_gates"]), L.ptr(sv["dec_c"]),
           L.ptr(sv["hc"]), L.ptr(_f32(datt)), B, Tmax, D, n_dec, L.ptr(dxproj.hi), L.ptr(dxproj.lo),
           L.ptr(ddec.hi), L.ptr(ddec.lo), L.stream())


def posterior_bwd(dpred: Tensor, emb: Tensor, att: Tensor, B: int, Tmax: int, n_att: int, demb: Tensor, datt: Tensor):
    D = emb.shape[1]
    L.call("eda_posterior_bwd", L.ptr(_f32(dpred)), L.ptr(_f32(emb)), emb.stride(0), L.ptr(_f32(att)), att.shape[1],
           B, Tmax, D, n_att, L.ptr(demb), demb.stride(0), L.ptr(datt), L.stream())


def attractor_head_bwd(att: Tensor, probs: Tensor, w: Tensor, n_spk_label: int, scale: float, datt: Tensor,
                       dw: Tensor, dbias: Tensor, accumulate: bool = False):
    B, n_dec, D = att.shape
    L.call("eda_attractor_head_bwd", L.ptr(_f32(att)), L.ptr(_f32(probs)), L.ptr(w), B, n_dec, D, n_spk_label,
           float(scale), L.ptr(datt), L.ptr(dw), L.ptr(dbias), int(accumulate), L.stream())


def scale_dev(scale: Tensor, x: Tensor):
    L.call("eda_scale_dev", L.ptr(scale), L.ptr(x), x.numel(), L.stream())


def adam_step(p: Tensor, g: Tensor, m: Tensor, v: Tensor, lr: float, beta1: float, beta2: float, eps: float, step: int):
    L.call("eda_adam_step", L.ptr(p), L.ptr(g), L.ptr(m), L.ptr(v), p.numel(), float(lr), float(beta1), float(beta2),
           float(eps), int(step), L.stream())


def dropout_bits(p: float, seed: int, site: int, rows: int, cols: int, out: Tensor):
    """Keep bits of dropout site ``site`` into ``out`` (rows, >= ceil(cols / 32)) int32: the attention kernels' ``drop_bits``."""
    assert out.dtype == torch.int32 and out.is_cuda and out.stride(1) == 1 and out.shape[0] >= rows
    L.call("eda_dropout_bits", float(p), int(seed), int(site), rows, cols, L.ptr(out), out.stride(0), L.stream())
    return out


def dropout_mask(p: float, seed: int, site: int, rows: int, cols: int, device) -> Tensor:
    """The keep mask (uint8 rows x cols) the kernels apply at dropout site ``site`` (tests)."""
    keep = torch.empty((rows, cols), dtype=torch.uint8, device=device)
    L.call("eda_dropout_mask", float(p), int(seed), int(site), rows, cols, L.ptr(keep), L.stream())
    return keep

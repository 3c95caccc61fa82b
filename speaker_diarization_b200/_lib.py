"""ctypes binding of libeda_b200.so (the C ABI declared in include/eda_b200.h).

Thin by design: tensors are unpacked to raw device pointers + sizes here and nothing else
happens on the Python side.  There is NO fallback: if the shared library is missing or a
call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libeda_b200.so")

EPI_BIAS, EPI_BIAS_RELU, EPI_BIAS_LN, EPI_BIAS_RES_LN = 0, 1, 2, 3
EPI_BIAS_RES, EPI_RELU_MASK, EPI_BIAS_SILU, EPI_BIAS_GLU = 5, 6, 7, 8
PREC_FP32, PREC_BF16X3, PREC_BF16 = 0, 1, 2
PRECISIONS = {"fp32": PREC_FP32, "bf16x3": PREC_BF16X3, "bf16": PREC_BF16}

_p, _i, _i64, _f, _u64 = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_uint64

# name -> (restype, argtypes); must list every symbol include/eda_b200.h declares
SIGNATURES = {
    "eda_abi_version": (_i, []),
    "eda_last_error_string": (C.c_char_p, []),
    "eda_launch_count": (_i64, []),
    "eda_frontend_plan_create": (_i, [_i, _i, _i, _i, C.POINTER(_p)]),
    "eda_frontend_plan_destroy": (_i, [_p]),
    "eda_frontend_num_frames": (_i64, [_i64, _i]),
    "eda_frontend_logmel": (_i, [_p, _p, _i, _p, _p, _i, _i64, _p, _p, _p]),
    "eda_frontend_transform": (_i, [_p, _p, _i, _p, _p, _i, _i64, _i, _p, _p, _p]),
    "eda_frontend_logmel_ranges": (_i, [_p, _p, _i, _p, _p, _p, _i, _i64, _p, _p, _p]),
    "eda_frontend_width": (_i, [_p, _i]),
    "eda_frontend_splice_subsample_w": (_i, [_p, _i, _p, _p, _p, _i, _i64, _i, _i, _p, _i, _p]),
    "eda_frontend_splice_subsample": (_i, [_p, _p, _p, _p, _p, _i, _i64, _i, _i, _p, _i, _p]),
    "eda_host_randperm_batch": (_i, [_p, _p, _p, _p, _i, _p]),
    "eda_pad_pack": (_i, [_p, _p, _p, _i, _i, _i, _f, _p, _i, _p]),
    "eda_linear": (_i, [_p, _i, _p, _i, _p, _p, _i, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "eda_mha_fwd": (_i, [_p, _i, _p, _i, _i, _i, _i, _i, _i, _p]),
    "eda_mha_fwd_masked": (_i, [_p, _i, _p, _i, _i, _i, _i, _i, _p, _i, _p]),
    "eda_split_bf16": (_i, [_p, _i, _i, _i, _p, _p, _i, _p]),
    "eda_pad_pack_split": (_i, [_p, _i, _p, _p, _i, _i, _i, _f, _p, _p, _i, _p]),
    "eda_gemm_split": (_i, [_p, _p, _i, _p, _p, _i, _p, _p, _i, _p, _p, _p, _i, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "eda_gemm_qkv": (_i, [_p, _p, _i, _p, _p, _i, _p, _p, _p, _i, _p, _p, _i, _i, _i, _i, _i, _i, _p]),
    "eda_mha_split": (_i, [_p, _p, _i, _p, _p, _i, _i, _i, _i, _p, _i, _p, _p, _i, _i, _p]),
    "eda_lstm_attractors": (_i, [_p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _p, _i, _p]),
    "eda_lstm_max_active_clusters": (_i, []),
    "eda_attractor_head": (_i, [_p, _p, _p, _i, _i, _i, _p, _i, _p, _p]),
    "eda_posterior": (_i, [_p, _i, _p, _i, _i, _i, _i, _i, _f, _p, _p, _p, _p]),
    "eda_posterior_packed": (_i, [_p, _i, _p, _i, _i, _i, _i, _i, _f, _p, _p, _p, _p, _p, _p, _p]),
    "eda_pit_loss": (_i, [_p, _p, _p, _i, _i, _i, _p, _p, _p, _p, _p]),
    "eda_pit_loss_static": (_i, [_p, _p, _p, _i, _i, _i, _p, _p, _p, _p, _p]),
    "eda_pit_loss_bwd": (_i, [_p, _p, _p, _i, _i, _i, _f, _p, _p]),
    "eda_der_counters": (_i, [_p, _p, _p, _i, _i, _i, _p, _p]),
    "eda_frame_labels": (_i, [_p, _p, _p, _p, _i64, _p, _p, _p, _p, _i, _i64, _i, C.c_double, _i, _i, _p, _i, _i, _p]),
    "eda_rttm_segments": (_i, [_p, _p, _i, _p, _i, _p, _i, _f, _i, _i, _p, _p, _p]),
    # ---- training step
    "eda_dropout_mask": (_i, [_f, _u64, _i, _i64, _i, _p, _p]),
    "eda_dropout_bits": (_i, [_f, _u64, _i, _i64, _i, _p, _i, _p]),
    "eda_gemm_split_train": (_i, [_p, _p, _i, _p, _p, _i, _p, _p, _i, _p, _p, _p, _i, _p, _p, _i, _p, _f, _u64, _i, _i, _i, _i, _i, _i, _p]),
    "eda_gemm_dgrad": (_i, [_p, _p, _i, _p, _p, _i, _p, _i, _p, _i, _f, _p, _i, _p, _p, _i, _i, _i, _i, _i, _p]),
    "eda_gemm_wgrad": (_i, [_p, _p, _i, _p, _p, _i, _i, _i, _i, _p, _i, _p, _i64, _i, _i, _p]),
    "eda_ln_bwd": (_i, [_p, _i, _p, _i, _p, _p, _p, _i, _i, _p, _i, _p, _p, _i, _p, _p, _p, _p, _i64, _i, _f, _u64, _i, _p]),
    "eda_colsum_split": (_i, [_p, _p, _i, _i, _i, _p, _p, _i64, _i, _p]),
    "eda_mha_split_train": (_i, [_p, _p, _i, _p, _p, _i, _i, _i, _i, _p, _p, _i, _p, _f, _u64, _i, _p, _i, _i, _p]),
    "eda_mha_delta": (_i, [_p, _p, _i, _p, _p, _i, _i, _i, _p, _p]),
    "eda_mha_bwd": (_i, [_p, _p, _i, _p, _p, _i, _p, _p, _i, _p, _p, _i, _i, _i, _p, _p, _p, _i, _f, _u64, _i, _p, _i, _i, _p]),
    "eda_scale_split": (_i, [_p, _i, _i, _i, _f, _p, _p, _i, _p]),
    "eda_lstm_attractors_train": (_i, [_p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _p]),
    "eda_lstm_attractors_bwd": (_i, [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _p, _p, _p, _p]),
    "eda_posterior_bwd": (_i, [_p, _p, _i, _p, _i, _i, _i, _i, _i, _p, _i, _p, _p]),
    "eda_attractor_head_bwd": (_i, [_p, _p, _p, _i, _i, _i, _i, _f, _p, _p, _p, _i, _p]),
    "eda_scale_dev": (_i, [_p, _p, _i64, _p]),
    "eda_adam_step": (_i, [_p, _p, _p, _p, _i64, _f, _f, _f, _f, _i, _p]),
    # ---- Conformer variant
    "eda_gemm_split_ex": (_i, [_p, _p, _i, _p, _p, _i, _p, _p, _i, _p, _p, _p, _i, _p, _p, _i, _f, _i, _i, _i, _i, _i, _p]),
    "eda_layernorm_split": (_i, [_p, _i, _p, _p, _i, _i, _p, _i, _p, _p, _i, _p]),
    "eda_dwconv_bn_silu": (_i, [_p, _i, _p, _p, _p, _i, _i, _i, _i, _p, _p, _i, _p]),
    "eda_mha_split_masked": (_i, [_p, _p, _i, _p, _p, _i, _i, _i, _i, _p, _p, _p, _i, _i, _p]),
}

_lib = None


class EdaError(RuntimeError):
    pass


def load() -> C.CDLL:
    """Load the shared library (no GPU needed to load; needed to call compute entry points)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise EdaError(
                f"{LIB_PATH} is missing: build it with `python -m speaker_diarization_b200.build` "
                "(there is no CPU or PyTorch fallback for this path)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        if lib.eda_abi_version() != 1:
            raise EdaError("libeda_b200.so ABI version mismatch")
        _lib = lib
    return _lib


def ptr(t):
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "device-resident contiguous tensor required"
    return t.data_ptr()


_STREAM_OVERRIDE = None


def stream() -> int:
    if _STREAM_OVERRIDE is not None:
        return _STREAM_OVERRIDE
    return torch.cuda.current_stream().cuda_stream


class use_stream:
    """Pin the stream handle the bindings launch on while torch's current stream is already set to
    it (``with torch.cuda.stream(s), L.use_stream(s): ...``): saves one torch.cuda.current_stream()
    lookup (~10 us) per C-ABI call on launch-heavy paths."""

    def __init__(self, s: "torch.cuda.Stream"):
        self.handle = s.cuda_stream

    def __enter__(self):
        global _STREAM_OVERRIDE
        self.prev, _STREAM_OVERRIDE = _STREAM_OVERRIDE, self.handle

    def __exit__(self, *exc):
        global _STREAM_OVERRIDE
        _STREAM_OVERRIDE = self.prev


# Optional per-entry-point device timing (bench.py's roofline leg): when PROFILE is a dict,
# every call is bracketed by CUDA events on the launching stream; read with profile_ms().
PROFILE = None


# When RECORD is a list, every call is also appended to it as (name, [args]): the model replays such a recording of
# its encoder with a handful of pointers patched instead of re-deriving ~30 argument lists per batch (host time per
# batch is what limits 8 ranks sharing one host).
RECORD = None


def call(name: str, *args) -> None:
    if RECORD is not None:
        RECORD.append((name, list(args)))
    if PROFILE is None:
        rc = getattr(load(), name)(*args)
    else:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = getattr(load(), name)(*args)
        e1.record()
        PROFILE.setdefault(name, []).append((e0, e1, args))
    if rc != 0:
        raise EdaError(f"{name} failed ({rc}): {load().eda_last_error_string().decode()}")


def profile_ms():
    """{entry point: [(milliseconds, args), ...]} for the calls recorded in PROFILE."""
    torch.cuda.synchronize()
    return {k: [(a.elapsed_time(b), args) for a, b, args in v] for k, v in (PROFILE or {}).items()}


def launch_count() -> int:
    return int(load().eda_launch_count())

"""Thin torch-tensor wrappers over single C-ABI kernels (used by the numerics tests and micro-benches)."""
from __future__ import annotations

import ctypes

import torch

from . import _lib

ACT = {"none": 0, "relu": 1, "quickgelu": 2, "sigmoid": 3}


def gemm_bf16(A: torch.Tensor, B: torch.Tensor, bias=None, resid=None, act: str = "none", alpha: float = 1.0,
              out_dtype=torch.float32, trans: bool = False, out: torch.Tensor = None) -> torch.Tensor:
    """out[b] = act(alpha * A[b] @ B[b]^T + bias) (+ resid).  A [batch,M,K] / [M,K] bf16; B [batch,N,K] / [N,K] bf16."""
    _lib.require_device()
    squeeze = A.dim() == 2
    if squeeze:
        A = A.unsqueeze(0)
    batch, M, K = A.shape
    b_bs = 0
    if B.dim() == 2:
        N = B.shape[0]
        ldb = B.stride(0)
    else:
        N = B.shape[1]
        ldb = B.stride(1)
        b_bs = B.stride(0)
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16 and A.stride(-1) == 1 and B.stride(-1) == 1
    if out is None:
        shape = (batch, N, M) if trans else (batch, M, N)
        out = torch.empty(shape, dtype=out_dtype, device=A.device)
    o = out if out.dim() == 3 else out.unsqueeze(0)
    rc = _lib.lib().wc_gemm_bf16(
        _lib.ptr(A), _lib.ptr(B), _lib.ptr(o), _lib.ptr(bias), _lib.ptr(resid), M, N, K, batch,
        ctypes.c_long(A.stride(1)), ctypes.c_long(ldb), ctypes.c_long(o.stride(1)), ctypes.c_long(A.stride(0)),
        ctypes.c_long(b_bs), ctypes.c_long(o.stride(0)), int(out.dtype == torch.bfloat16), int(trans), ACT[act],
        ctypes.c_float(alpha), _lib.stream_ptr())
    _lib.check(rc, "wc_gemm_bf16")
    return out[0] if squeeze and out.dim() == 3 else out

"""Thin torch-tensor wrappers over single C-ABI kernels (used by the numerics tests and micro-benches)."""
from __future__ import annotations

import ctypes
import math

import torch

from . import _lib

ACT = {"none": 0, "relu": 1, "quickgelu": 2, "sigmoid": 3}


def gemm_f16(A: torch.Tensor, B: torch.Tensor, bias=None, resid=None, act: str = "none", alpha: float = 1.0,
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
    assert A.dtype == torch.float16 and B.dtype == torch.float16 and A.stride(-1) == 1 and B.stride(-1) == 1
    if out is None:
        shape = (batch, N, M) if trans else (batch, M, N)
        out = torch.empty(shape, dtype=out_dtype, device=A.device)
    o = out if out.dim() == 3 else out.unsqueeze(0)
    rc = _lib.lib().wc_gemm_f16(
        _lib.ptr(A), _lib.ptr(B), _lib.ptr(o), _lib.ptr(bias), _lib.ptr(resid), M, N, K, batch,
        ctypes.c_long(A.stride(1)), ctypes.c_long(ldb), ctypes.c_long(o.stride(1)), ctypes.c_long(A.stride(0)),
        ctypes.c_long(b_bs), ctypes.c_long(o.stride(0)), int(out.dtype == torch.float16), int(trans), ACT[act],
        ctypes.c_float(alpha), _lib.stream_ptr())
    _lib.check(rc, "wc_gemm_f16")
    return out[0] if squeeze and out.dim() == 3 else out


def resize_bilinear(x: torch.Tensor, size=None, scale_factor: float = None) -> torch.Tensor:
    """F.interpolate(x, size=|scale_factor=, mode='bilinear', align_corners=False) for fp32 [N,C,H,W]
    (test_msc_flip_voc.py:63,80)."""
    _lib.require_device()
    x = x.contiguous().float()
    N, C, Hi, Wi = x.shape
    if size is not None:
        Ho, Wo, sf = int(size[0]), int(size[1]), 0.0
    else:
        Ho, Wo, sf = int(math.floor(Hi * scale_factor)), int(math.floor(Wi * scale_factor)), float(scale_factor)
    out = torch.empty(N, C, Ho, Wo, dtype=torch.float32, device=x.device)
    rc = _lib.lib().wc_resize_bilinear(_lib.ptr(x), _lib.ptr(out), N * C, Hi, Wi, Ho, Wo, ctypes.c_float(sf),
                                       ctypes.c_float(sf), _lib.stream_ptr())
    _lib.check(rc, "wc_resize_bilinear")
    return out


def flip_concat(x: torch.Tensor) -> torch.Tensor:
    """torch.cat([x, x.flip(-1)], dim=0) for fp32 [B,C,H,W] (test_msc_flip_voc.py:68,82)."""
    _lib.require_device()
    x = x.contiguous().float()
    n, C, H, W = x.shape
    out = torch.empty(2 * n, C, H, W, dtype=torch.float32, device=x.device)
    rc = _lib.lib().wc_flip_concat(_lib.ptr(x), _lib.ptr(out), n * C, H, W, _lib.stream_ptr())
    _lib.check(rc, "wc_flip_concat")
    return out


def msc_accumulate(seg2: torch.Tensor, acc: torch.Tensor, weight: float, first: bool) -> torch.Tensor:
    """acc[B,C,h,w] (+)= weight * (R(seg2[:B]) + flip(R(seg2[B:]))) / 2 with R the bilinear resize to acc's size."""
    _lib.require_device()
    seg2 = seg2.contiguous()
    n2, C, hs, ws = seg2.shape                  # [2B,C,hs,ws]: B plain images followed by their B flipped twins
    B = n2 // 2
    assert n2 == 2 * B and acc.is_contiguous() and tuple(acc.shape[:2]) == (B, C)
    rc = _lib.lib().wc_msc_accumulate(_lib.ptr(seg2), _lib.ptr(acc), B, C, hs, ws, acc.shape[2], acc.shape[3],
                                      ctypes.c_float(weight), int(first), _lib.stream_ptr())
    _lib.check(rc, "wc_msc_accumulate")
    return acc


def upsample_argmax_hist(logits: torch.Tensor, size, gt: torch.Tensor = None, hist: torch.Tensor = None,
                         want_pred: bool = True):
    """argmax over classes of the bilinear resize of logits [C,h,w] to `size`; optional confusion-matrix update."""
    _lib.require_device()
    logits = logits.contiguous()
    squeeze = logits.dim() == 3
    if squeeze:
        logits = logits.unsqueeze(0)
    B, C, h, w = logits.shape
    Ho, Wo = int(size[0]), int(size[1])
    pred = torch.empty(B, Ho, Wo, dtype=torch.int32, device=logits.device) if want_pred else None
    if gt is not None:
        assert gt.dtype == torch.int16 and gt.is_contiguous() and gt.numel() == B * Ho * Wo
        assert hist.dtype == torch.int64 and hist.is_contiguous() and hist.numel() == C * C
    rc = _lib.lib().wc_upsample_argmax_hist(_lib.ptr(logits), B, C, h, w, Ho, Wo, _lib.ptr(pred), _lib.ptr(gt),
                                            _lib.ptr(hist), _lib.stream_ptr())
    _lib.check(rc, "wc_upsample_argmax_hist")
    return pred[0] if (squeeze and pred is not None) else pred


def confusion_hist(pred: torch.Tensor, gt: torch.Tensor, hist: torch.Tensor) -> torch.Tensor:
    """hist[gt*C+pred] += 1 for valid gt (utils/evaluate.py:9-15); pred int64, gt int16, hist int64 [C,C]."""
    _lib.require_device()
    assert pred.dtype == torch.int64 and gt.dtype == torch.int16 and hist.dtype == torch.int64
    pred, gt = pred.contiguous(), gt.contiguous()
    assert pred.numel() == gt.numel()
    rc = _lib.lib().wc_confusion_hist(_lib.ptr(pred), _lib.ptr(gt), ctypes.c_long(pred.numel()), hist.shape[0],
                                      _lib.ptr(hist), _lib.stream_ptr())
    _lib.check(rc, "wc_confusion_hist")
    return hist


def attention_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, want_avg: bool = True):
    """Fused attention (head_dim 64 or 32).  q, k, v fp16 [N, heads, L, hd] (q already scaled by 1/sqrt(hd)).
    Returns (o fp16 [N, L, heads*hd], attn_avg fp32 [N, L, L] | None; the map needs head_dim 64)."""
    _lib.require_device()
    N, heads, L, hd = q.shape
    assert hd in (64, 32) and q.dtype == torch.float16
    Lp = (L + 7) // 8 * 8
    if hd == 64:
        vt = v.contiguous()          # head_dim 64: V itself (MN-major tensor-core operand), no transposed copy
    else:
        vt = torch.full((N, heads, hd, Lp), float("nan"), dtype=torch.float16, device=q.device)   # pad columns must not matter
        vt[..., :L] = v.transpose(-1, -2)
    o = torch.empty(N, L, heads * hd, dtype=torch.float16, device=q.device)
    lse = torch.empty(N, heads, L, dtype=torch.float32, device=q.device)
    avg = torch.empty(N, L, L, dtype=torch.float32, device=q.device) if want_avg else None
    rc = _lib.lib().wc_attention_fwd(_lib.ptr(q.contiguous()), _lib.ptr(k.contiguous()), _lib.ptr(vt), _lib.ptr(o),
                                     _lib.ptr(lse), _lib.ptr(avg), N, heads, L, Lp, hd, _lib.stream_ptr())
    _lib.check(rc, "wc_attention_fwd")
    return o, avg


def par_affinity(img: torch.Tensor, dilations, dtype=torch.float32, raw: bool = False) -> torch.Tensor:
    """wc_par_affinity (PAR.py:70-86): img fp32 [B,3,H,W] -> affinity [B,NN,H,W] (``raw``: the kernel's own blocked
    [B,H,Wb,NN,64] tensor, the layout wc_par_propagate streams)."""
    _lib.require_device()
    L = _lib.lib()
    img = img.contiguous().float()
    B, _, H, W = img.shape
    nn_ = 8 * len(dilations)
    wa = L.wc_par_aff_pitch(W)
    aff = torch.zeros(B, H, wa // 64, nn_, 64, dtype=dtype, device=img.device)
    rc = L.wc_par_affinity(_lib.ptr(img), _lib.ptr(aff), int(dtype == torch.float16), B, H, W, _lib.int_array(dilations),
                           len(dilations), _lib.stream_ptr())
    _lib.check(rc, "wc_par_affinity")
    if raw:
        return aff
    return aff.permute(0, 3, 1, 2, 4).reshape(B, nn_, H, wa)[..., :W].contiguous()


def par_propagate(aff_raw: torch.Tensor, masks: torch.Tensor, ch_off, dilations, num_iter: int) -> torch.Tensor:
    """wc_par_propagate (PAR.py:88-90) on a blocked affinity tensor from par_affinity(raw=True); masks packed [sumKc,H,W]."""
    _lib.require_device()
    L = _lib.lib()
    masks = masks.contiguous().float()
    sum_kc, H, W = masks.shape
    out = torch.empty_like(masks)
    scratch = torch.empty(2 * sum_kc * H * L.wc_par_aff_pitch(W), dtype=torch.float32, device=masks.device)
    rc = L.wc_par_propagate(_lib.ptr(aff_raw), int(aff_raw.dtype == torch.float16), _lib.ptr(masks), _lib.ptr(out),
                            _lib.ptr(scratch), _lib.int_array(ch_off), len(ch_off) - 1, H, W, _lib.int_array(dilations),
                            len(dilations), int(num_iter), _lib.stream_ptr())
    _lib.check(rc, "wc_par_propagate")
    return out

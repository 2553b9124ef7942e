"""Training-time label prep and losses on the hot path's outputs, mirroring the reference names:

    get_mask_by_radius(h, w, radius)                       scripts/dist_clip_voc.py:116-133
    cams_to_affinity_label(cam_label, mask, ignore_index)  utils/camutils.py:226-247
    get_aff_loss(inputs, targets)                          utils/losses.py:11-22
    get_seg_loss(pred, label, ignore_index)                utils/losses.py:24-33

plus ``aff_loss_from_labels`` — the fused form of ``get_aff_loss(attn_pred, cams_to_affinity_label(cam, mask))`` that
never materialises the [B,P,P] int64 label tensor.  ctypes glue over ``labelprep.cu``: no arithmetic here, no CPU path.
The losses return tensors attached to torch autograd (the CUDA pass emits d loss / d input alongside the value), so
``loss.backward()`` in the reference training loop (scripts/dist_clip_voc.py:264) keeps working.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _lib


def _workspace(B: int, H: int, W: int, device) -> torch.Tensor:
    n = int(_lib.lib().wc_labelprep_workspace_bytes(B, H, W))
    return torch.empty(n, dtype=torch.uint8, device=device)


def radius_mask(h: int = 20, w: int = 20, radius: int = 8, device="cuda") -> torch.Tensor:
    """Device form of get_mask_by_radius: uint8 [h*w, h*w]."""
    _lib.require_device()
    mask = torch.empty(h * w, h * w, dtype=torch.uint8, device=device)
    rc = _lib.lib().wc_radius_mask(h, w, radius, _lib.ptr(mask), _lib.stream_ptr())
    _lib.check(rc, "wc_radius_mask")
    return mask


def get_mask_by_radius(h: int = 20, w: int = 20, radius: int = 8) -> np.ndarray:
    """Reference signature and return type (float64 numpy [h*w, h*w]); computed on the device."""
    return radius_mask(h, w, radius).cpu().numpy().astype(np.float64)


def _as_mask_tensor(mask, device):
    if mask is None:
        return None
    if isinstance(mask, np.ndarray):
        mask = torch.from_numpy(np.ascontiguousarray(mask))
    return (mask != 0).to(device=device, dtype=torch.uint8).contiguous()


def cams_to_affinity_label(cam_label: torch.Tensor, mask=None, ignore_index: int = 255) -> torch.Tensor:
    """cam_label int64 [B,H,W] -> aff_label int64 [B,P,P], P = (H//16)*(W//16)."""
    _lib.require_device()
    assert cam_label.dim() == 3
    cam_label = cam_label.to(torch.int64).contiguous()
    B, H, W = cam_label.shape
    P = (H // 16) * (W // 16)
    m = _as_mask_tensor(mask, cam_label.device)
    assert m is None or tuple(m.shape) == (P, P)
    out = torch.empty(B, P, P, dtype=torch.int64, device=cam_label.device)
    ws = _workspace(B, H, W, cam_label.device)
    rc = _lib.lib().wc_affinity_label(_lib.ptr(cam_label), B, H, W, _lib.ptr(m), int(ignore_index), _lib.ptr(out),
                                      _lib.ptr(ws), ctypes.c_size_t(ws.numel()), _lib.stream_ptr())
    _lib.check(rc, "wc_affinity_label")
    return out


class _GradOfInput(torch.autograd.Function):
    """loss with a precomputed d loss / d input (both produced by the same CUDA pass)."""

    @staticmethod
    def forward(ctx, inp, loss, grad):
        ctx.save_for_backward(grad)
        return loss.clone()

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * g, None, None


def get_aff_loss(inputs: torch.Tensor, targets: torch.Tensor):
    """-> (loss, pos_count, neg_count) as the reference returns them (0-dim tensors)."""
    _lib.require_device()
    x = inputs.float().contiguous()
    t = targets.to(torch.int64).contiguous()
    assert x.numel() == t.numel()
    need_grad = inputs.requires_grad and torch.is_grad_enabled()
    loss = torch.empty(1, dtype=torch.float32, device=x.device)
    counts = torch.empty(2, dtype=torch.int64, device=x.device)
    grad = torch.empty_like(x) if need_grad else None
    ws = torch.empty(64, dtype=torch.uint8, device=x.device)
    rc = _lib.lib().wc_aff_loss(_lib.ptr(x), _lib.ptr(t), ctypes.c_long(x.numel()), _lib.ptr(loss), _lib.ptr(counts),
                                _lib.ptr(grad), _lib.ptr(ws), ctypes.c_size_t(64), _lib.stream_ptr())
    _lib.check(rc, "wc_aff_loss")
    out = _GradOfInput.apply(inputs, loss[0], grad.view_as(inputs)) if need_grad else loss[0]
    return out, counts[0], counts[1]


def aff_loss_from_labels(attn_pred: torch.Tensor, cam_label: torch.Tensor, radius: int = 8, ignore_index: int = 255):
    """Fused get_aff_loss(attn_pred, cams_to_affinity_label(cam_label, get_mask_by_radius(h, w, radius))).
    radius < 0 means no radius mask."""
    _lib.require_device()
    x = attn_pred.float().contiguous()
    lab = cam_label.to(torch.int64).contiguous()
    B, H, W = lab.shape
    P = (H // 16) * (W // 16)
    assert tuple(x.shape) == (B, P, P)
    need_grad = attn_pred.requires_grad and torch.is_grad_enabled()
    loss = torch.empty(1, dtype=torch.float32, device=x.device)
    counts = torch.empty(2, dtype=torch.int64, device=x.device)
    grad = torch.empty_like(x) if need_grad else None
    ws = _workspace(B, H, W, x.device)
    rc = _lib.lib().wc_aff_loss_fused(_lib.ptr(x), _lib.ptr(lab), B, H, W, int(radius), int(ignore_index),
                                      _lib.ptr(loss), _lib.ptr(counts), _lib.ptr(grad), _lib.ptr(ws),
                                      ctypes.c_size_t(ws.numel()), _lib.stream_ptr())
    _lib.check(rc, "wc_aff_loss_fused")
    out = _GradOfInput.apply(attn_pred, loss[0], grad.view_as(attn_pred)) if need_grad else loss[0]
    return out, counts[0], counts[1]


def get_seg_loss(pred: torch.Tensor, label: torch.Tensor, ignore_index: int = 255) -> torch.Tensor:
    """pred fp32 [B,C,H,W] (already at label resolution, scripts/dist_clip_voc.py:248), label [B,H,W]."""
    _lib.require_device()
    x = pred.float().contiguous()
    lab = label.to(torch.int64).contiguous()
    B, C, H, W = x.shape
    assert tuple(lab.shape) == (B, H, W)
    need_grad = pred.requires_grad and torch.is_grad_enabled()
    loss = torch.empty(1, dtype=torch.float32, device=x.device)
    grad = torch.empty_like(x) if need_grad else None
    ws = torch.empty(64, dtype=torch.uint8, device=x.device)
    rc = _lib.lib().wc_seg_loss(_lib.ptr(x), _lib.ptr(lab), B, C, H, W, int(ignore_index), _lib.ptr(loss),
                                _lib.ptr(grad), _lib.ptr(ws), ctypes.c_size_t(64), _lib.stream_ptr())
    _lib.check(rc, "wc_seg_loss")
    return _GradOfInput.apply(pred, loss[0], grad) if need_grad else loss[0]

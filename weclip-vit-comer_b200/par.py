"""PAR — pixel-adaptive refinement module, CUDA (sm_100a) drop-in for WeCLIP_model/PAR.py:26-92.

Same constructor and forward signature, same ``kernel`` buffer in the state dict (PAR.py:33); the
arithmetic runs in ``wc_par_fwd`` (csrc/par.cu).  No CPU path.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib


def get_kernel() -> torch.Tensor:
    """The one-hot 3x3 neighbour selectors of PAR.py:10-24 (kept only for state_dict compatibility)."""
    weight = torch.zeros(8, 1, 3, 3)
    for i, (r, c) in enumerate([(0, 0), (0, 1), (0, 2), (1, 0), (1, 2), (2, 0), (2, 1), (2, 2)]):
        weight[i, 0, r, c] = 1
    return weight


class PAR(nn.Module):
    def __init__(self, dilations, num_iter, affinity_dtype=torch.float32):
        """``affinity_dtype``: storage type of the materialised affinity (fp32 = exact, within 2e-5 of PAR.py;
        fp16 = the WeCLIP pipeline's setting, half the bytes per propagation step)."""
        super().__init__()
        assert affinity_dtype in (torch.float32, torch.float16)
        self.affinity_dtype = affinity_dtype
        self.dilations = list(dilations)
        self.num_iter = int(num_iter)
        self.register_buffer("kernel", get_kernel())
        self.dim = 2
        self.w1 = 0.3
        self.w2 = 0.01
        self._ws = {}

    def _workspace(self, nbytes: int, device) -> torch.Tensor:
        # under CUDA-graph capture: a private allocation from the graph's pool (fixed address for every replay)
        if device.type == "cuda" and torch.cuda.is_current_stream_capturing():
            return torch.empty(nbytes, dtype=torch.uint8, device=device)
        # one scratch arena per (device, CUDA stream): callers may run independent batches on separate streams
        key = (str(device), torch.cuda.current_stream(device).cuda_stream if device.type == "cuda" else 0)
        ws = self._ws.get(key)
        if ws is None or ws.numel() < nbytes:
            ws = torch.empty(nbytes, dtype=torch.uint8, device=device)
            self._ws[key] = ws
        return ws

    def refine_ragged(self, imgs: torch.Tensor, masks: torch.Tensor, ch_off) -> torch.Tensor:
        """imgs [B,3,Hi,Wi]; masks packed [sumKc,H,W] with image b owning channels ch_off[b]:ch_off[b+1]."""
        _lib.require_device()
        L = _lib.lib()
        imgs = imgs.contiguous().float()
        masks = masks.contiguous().float()
        B, _, Hi, Wi = imgs.shape
        sum_kc, H, W = masks.shape
        assert len(ch_off) == B + 1 and ch_off[-1] == sum_kc
        out = torch.empty_like(masks)
        nbytes = L.wc_par_workspace_bytes(B, sum_kc, H, W, len(self.dilations))
        ws = self._workspace(nbytes, imgs.device)
        rc = L.wc_par_fwd(_lib.ptr(imgs), Hi, Wi, _lib.ptr(masks), _lib.ptr(out), _lib.int_array(ch_off), B, H, W,
                          _lib.int_array(self.dilations), len(self.dilations), self.num_iter,
                          int(self.affinity_dtype == torch.float16), _lib.ptr(ws), _lib.c_size_t(ws.numel()), _lib.stream_ptr())
        _lib.check(rc, "wc_par_fwd")
        return out

    def forward(self, imgs: torch.Tensor, masks: torch.Tensor) -> torch.Tensor:
        """PAR.py:64-92: imgs [B,3,Hi,Wi], masks [B,Kc,H,W] -> refined masks [B,Kc,H,W]."""
        B, Kc, H, W = masks.shape
        out = self.refine_ragged(imgs, masks.reshape(B * Kc, H, W), [b * Kc for b in range(B + 1)])
        return out.reshape(B, Kc, H, W)


def argmax_lut(masks: torch.Tensor, lut: torch.Tensor, ch_off) -> torch.Tensor:
    """model_attn_aff_voc.py:54-57: per-image argmax over channels then valid_key lookup -> int64 [B,H,W]."""
    _lib.require_device()
    sum_kc, H, W = masks.shape
    B = len(ch_off) - 1
    labels = torch.empty(B, H, W, dtype=torch.int64, device=masks.device)
    lut = lut.to(device=masks.device, dtype=torch.int64).contiguous()
    rc = _lib.lib().wc_argmax_lut(_lib.ptr(masks.contiguous()), _lib.ptr(lut), _lib.ptr(labels),
                                  _lib.int_array(ch_off), B, H, W, _lib.stream_ptr())
    _lib.check(rc, "wc_argmax_lut")
    return labels

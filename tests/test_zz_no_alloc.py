"""wc_aff_loss_fused_static: the allocation-free variant of wc_aff_loss_fused (static __device__ scratch, labelprep.cu) through
the C ABI against the existing entry point — loss, counts and gradient, on two calls in a row with different inputs and
sizes (the shared scratch keeps the previous call's contents), plus the size limit.  Tolerance as tests/test_training_gpu.py
(1e-5 relative; the two entry points run the same kernels, so the results are in fact identical).  Collected last on purpose."""
import ctypes

import pytest
import torch

from weclip_b200 import _lib

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _inputs(B, H, W, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    P = (H // 16) * (W // 16)
    cam = torch.randint(0, 4, (B, H // 16, W // 16), device="cuda", generator=g)
    cam = torch.where(cam == 3, torch.full_like(cam, 255), cam)                    # some ignored patches
    cam = cam.repeat_interleave(16, 1).repeat_interleave(16, 2).contiguous().to(torch.int64)
    attn = torch.rand(B, P, P, device="cuda", generator=g)
    return attn, cam, P


def _existing(attn, cam, B, H, W, radius):
    L = _lib.lib()
    loss = torch.empty(1, dtype=torch.float32, device="cuda")
    counts = torch.empty(2, dtype=torch.int64, device="cuda")
    grad = torch.empty_like(attn)
    ws = torch.empty(int(L.wc_labelprep_workspace_bytes(B, H, W)), dtype=torch.uint8, device="cuda")
    rc = L.wc_aff_loss_fused(_lib.ptr(attn), _lib.ptr(cam), B, H, W, radius, 255, _lib.ptr(loss), _lib.ptr(counts),
                             _lib.ptr(grad), _lib.ptr(ws), ctypes.c_size_t(ws.numel()), _lib.stream_ptr())
    _lib.check(rc, "wc_aff_loss_fused")
    return loss, counts, grad


def _static(attn, cam, B, H, W, radius, loss, counts, grad):
    L = _lib.lib()
    return L.wc_aff_loss_fused_static(_lib.ptr(attn), _lib.ptr(cam), B, H, W, radius, 255, _lib.ptr(loss), _lib.ptr(counts),
                                      _lib.ptr(grad), _lib.stream_ptr())


def test_static_scratch_entry_point_matches_the_workspace_one_on_two_calls_in_a_row():
    _lib.require_device()
    # outputs are allocated up front: the entry point itself must not need any allocation
    cases = [(3, 160, 192, 8, 5), (2, 96, 112, 2, 6)]          # second call is smaller: stale labels stay behind it in the scratch
    outs = [(torch.full((1,), -1.0, device="cuda"), torch.zeros(2, dtype=torch.int64, device="cuda")) for _ in cases]
    for (B, H, W, radius, seed), (loss, counts) in zip(cases, outs):
        attn, cam, P = _inputs(B, H, W, seed)
        grad = torch.full_like(attn, 7.0)
        rc = _static(attn, cam, B, H, W, radius, loss, counts, grad)
        _lib.check(rc, "wc_aff_loss_fused_static")
        loss_e, counts_e, grad_e = _existing(attn, cam, B, H, W, radius)
        torch.cuda.synchronize()
        assert torch.isfinite(loss).all() and float(loss) > 0
        assert abs(float(loss) - float(loss_e)) <= RTOL * abs(float(loss_e))
        assert counts.tolist() == counts_e.tolist() and counts[0] > 1 and counts[1] > 1
        assert torch.allclose(grad, grad_e, rtol=RTOL, atol=1e-10)
        assert (grad != 7.0).any()


def test_static_scratch_entry_point_rejects_sizes_beyond_its_maximum():
    _lib.require_device()
    L = _lib.lib()
    L.wc_aff_loss_static_max_patches.restype = ctypes.c_long
    cap = int(L.wc_aff_loss_static_max_patches())
    assert cap == 65536
    B, H, W = 65, 512, 512                                      # 65 * 1024 patches > 65536
    assert B * (H // 16) * (W // 16) > cap
    dummy = torch.zeros(16, device="cuda")                      # rejected before any pointer is touched
    lab = torch.zeros(16, dtype=torch.int64, device="cuda")
    rc = L.wc_aff_loss_fused_static(_lib.ptr(dummy), _lib.ptr(lab), B, H, W, 8, 255, _lib.ptr(dummy), _lib.ptr(lab),
                                    ctypes.c_void_p(0), _lib.stream_ptr())
    assert rc == -3                                             # WC_ERR_WORKSPACE
    msg = L.wc_last_error().decode()
    assert "exceed the static scratch" in msg and "65536" in msg
    # exactly at the limit is accepted (64 crops of 512 x 512)
    B = 64
    attn, cam, P = _inputs(B, H, W, 9)
    assert B * P == cap
    loss = torch.empty(1, device="cuda")
    counts = torch.empty(2, dtype=torch.int64, device="cuda")
    assert _static(attn, cam, B, H, W, 8, loss, counts, None) == 0
    loss_e, counts_e, _ = _existing(attn, cam, B, H, W, 8)
    torch.cuda.synchronize()
    assert abs(float(loss) - float(loss_e)) <= RTOL * abs(float(loss_e)) and counts.tolist() == counts_e.tolist()

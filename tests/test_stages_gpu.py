"""CUDA stage kernels fed DIRECTLY with the reference's golden vectors (tests/golden/, written by oracle/make_golden.py from
the real reference): box masks (a11, integer work: bit-exact), compute_trans_mat in its factored form (a10) and the
cv2-resize CAM up-sampling (a12).  Every call goes through the C ABI (wc_affinity_refine / wc_cam_upsample_bg)."""
import ctypes
import os

import numpy as np
import pytest
import torch

from weclip_b200 import _lib

pytestmark = pytest.mark.gpu


def _refine(attn_pp: np.ndarray, cams: np.ndarray, h: int, w: int, thr: float):
    """wc_affinity_refine for ONE image whose fused affinity is the [P,P] matrix attn_pp (use_rfm=0, nl=1: B = the single
    map's [1:,1:] block) and K = cams.shape[0] classes.  Returns (refined [K,P], box masks [K,P])."""
    L = _lib.lib()
    P, K = h * w, cams.shape[0]
    attn = torch.zeros(1, 1, P + 1, P + 1)
    attn[0, 0, 1:, 1:] = torch.from_numpy(attn_pp)
    attn, cams_d = attn.cuda(), torch.from_numpy(np.ascontiguousarray(cams, np.float32)).reshape(K, P).cuda()
    out, masks = torch.empty(K, P, device="cuda"), torch.empty(K, P, device="cuda")
    ws = torch.empty(L.wc_affinity_workspace_bytes(1, h, w, K, 1), dtype=torch.uint8, device="cuda")
    rc = L.wc_affinity_refine(_lib.ptr(attn), 1, None, None, 1, h, w, 0, _lib.ptr(cams_d), _lib.int_array([0, K]),
                              ctypes.c_double(thr), _lib.ptr(out), _lib.ptr(masks), None, _lib.ptr(ws),
                              ctypes.c_size_t(ws.numel()), _lib.stream_ptr())
    _lib.check(rc, "wc_affinity_refine")
    return out.cpu().numpy(), masks.cpu().numpy()


def test_box_masks_bit_exact_vs_cv2_reference(golden_dir):
    """scoremap2bbox (clip/utils.py:115-142: uint8 quantisation, strict threshold, cv2.findContours + boundingRect,
    min(x+w, W-1) clip) + the exclusive-upper slice fill of clip_tool.py:181-183 — 24 reference cases incl. all-zero,
    all-one and a component confined to the last row.  Integer work: every cell must match."""
    g = np.load(os.path.join(golden_dir, "bbox.npz"))
    rng = np.random.default_rng(0)
    n = 0
    for cam, mask, (h, w, thr) in zip(g["cams"], g["masks"], g["meta"]):
        h, w = int(h), int(w)
        a = rng.random((h * w, h * w)).astype(np.float32) + 0.05
        _, got = _refine(a, cam[None, :h, :w], h, w, float(thr))
        assert np.array_equal(got.reshape(h, w), mask[:h, :w]), f"case {n} ({h}x{w}, thr {thr})"
        n += 1
    assert n == 24


def test_box_threshold_is_evaluated_in_float64():
    """int(0.7 * 250) = 175 in float64 (the reference, numpy) but 174 with a float32 0.7f: a cell at uint8 175 must NOT
    pass the strict '>' test (box_thr crosses the ABI as a double)."""
    h, w = 8, 8
    cam = np.zeros((h, w), np.float32)
    cam[1, 1] = 250.4 / 255.0          # uint8 250 = the maximum
    cam[4:6, 4:6] = 175.4 / 255.0      # uint8 175 == int(0.7 * 250): not above the threshold
    assert int(np.float64(0.7) * 250) == 175 and int(np.float64(np.float32(0.7)) * 250) == 174   # (double)0.7f * 250
    a = np.random.default_rng(1).random((h * w, h * w)).astype(np.float32) + 0.05
    _, got = _refine(a, cam[None], h, w, 0.7)
    want = np.zeros((h, w), np.float32)
    want[1:2, 1:2] = 1                  # only the maximum cell's component: box [1,2) x [1,2)
    assert np.array_equal(got.reshape(h, w), want)


def test_trans_mat_factored_form_vs_reference(golden_dir):
    """compute_trans_mat (clip/clip_tool.py:64-80) is never materialised on the device: cam' = T_s (T_s (mask * cam)).
    Against the reference's own T (30x30 golden): refined == T @ (mask * cam) within 1e-5 relative."""
    g = np.load(os.path.join(golden_dir, "transmat.npz"))
    a, T = g["a"], g["out"].astype(np.float64)
    h, w = 5, 6
    rng = np.random.default_rng(3)
    cams = rng.random((8, h, w)).astype(np.float32)
    refined, masks = _refine(a, cams, h, w, 0.4)
    want = (T @ (masks.astype(np.float64) * cams.reshape(8, -1)).T).T
    assert masks.sum() > 0
    assert np.abs(refined - want).max() <= 1e-5 * np.abs(want).max()


@pytest.mark.parametrize("tag,size", [("up", (112, 96)), ("odd", (90, 70)), ("down", (5, 4))])
def test_cam_upsample_vs_cv2_reference(golden_dir, tag, size):
    """generate_cam_label (clip/clip_tool.py:202-216): per-class min-max then cv2.resize(INTER_LINEAR) to (w,h); plus the
    background channel 1 - max_k (model_attn_aff_voc.py:162-163)."""
    g = np.load(os.path.join(golden_dir, "camlabel.npz"))
    refined = torch.from_numpy(g["refined"]).cuda().contiguous()       # [3,6,7]
    K, h, w = refined.shape
    Wc, Hc = size
    out = torch.empty(K + 1, Hc, Wc, device="cuda")
    stats = torch.empty(2 * K, device="cuda")
    rc = _lib.lib().wc_cam_upsample_bg(_lib.ptr(refined), _lib.int_array([0, K]), 1, h, w, Hc, Wc, ctypes.c_float(1.0),
                                       _lib.ptr(out), _lib.ptr(stats), _lib.stream_ptr())
    _lib.check(rc, "wc_cam_upsample_bg")
    out = out.cpu().numpy()
    assert np.abs(out[1:] - g[tag]).max() <= 2e-6
    assert np.abs(out[0] - (1.0 - g[tag].max(0))).max() <= 2e-6

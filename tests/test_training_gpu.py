"""CUDA label prep + losses (labelprep.cu through the C ABI) against the reference's golden outputs and the oracle.
Integer outputs are compared exactly; loss values within 1e-5 relative (fp64 block sums vs torch's fp32 tree sums),
gradients within 1e-5 relative."""
import os

import numpy as np
import pytest
import torch

from oracle import weclip_oracle as O
from oracle.make_golden import training_case
from weclip_b200 import losses as L

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _g(golden_dir):
    return np.load(os.path.join(golden_dir, "training.npz"), allow_pickle=False)


def test_radius_mask_golden(golden_dir):
    g = _g(golden_dir)
    for r in (2, 8):
        np.testing.assert_array_equal(L.radius_mask(6, 7, r).cpu().numpy(), g[f"mask_r{r}"])
    m = L.get_mask_by_radius(20, 20, 8)
    assert m.dtype == np.float64 and m.shape == (400, 400)
    np.testing.assert_array_equal(m, O.mask_by_radius(20, 20, 8))


def test_affinity_label_golden(golden_dir):
    g = _g(golden_dir)
    cam, _, _ = training_case()
    camd = cam.cuda()
    for r in (2, 8):
        got = L.cams_to_affinity_label(camd, mask=g[f"mask_r{r}"].astype(np.float64), ignore_index=255)
        assert got.dtype == torch.int64
        np.testing.assert_array_equal(got.cpu().numpy(), g[f"aff_r{r}"].astype(np.int64))
    got = L.cams_to_affinity_label(camd, mask=None, ignore_index=255)
    np.testing.assert_array_equal(got.cpu().numpy(), g["aff_nomask"].astype(np.int64))


def test_aff_loss_golden_value_counts_grad(golden_dir):
    g = _g(golden_dir)
    _, attn, _ = training_case()
    for r in (2, 8):
        t = torch.from_numpy(g[f"aff_r{r}"].astype(np.int64)).cuda()
        loss, pc, nc = L.get_aff_loss(attn.cuda(), t)
        np.testing.assert_allclose(float(loss), g[f"affloss_r{r}"][0], rtol=RTOL)
        assert (int(pc), int(nc)) == tuple(int(v) for v in g[f"affloss_r{r}"][1:])
    x = attn.cuda().requires_grad_(True)
    loss, _, _ = L.get_aff_loss(x, torch.from_numpy(g["aff_r2"].astype(np.int64)).cuda())
    (2.0 * loss).backward()
    np.testing.assert_allclose(x.grad.cpu().numpy(), 2.0 * g["affgrad_r2"], rtol=RTOL, atol=1e-10)


def test_aff_loss_fused_equals_two_step(golden_dir):
    g = _g(golden_dir)
    cam, attn, _ = training_case()
    for r in (2, 8):
        loss, pc, nc = L.aff_loss_from_labels(attn.cuda(), cam.cuda(), radius=r, ignore_index=255)
        np.testing.assert_allclose(float(loss), g[f"affloss_r{r}"][0], rtol=RTOL)
        assert (int(pc), int(nc)) == tuple(int(v) for v in g[f"affloss_r{r}"][1:])
    x = attn.cuda().requires_grad_(True)
    loss, _, _ = L.aff_loss_from_labels(x, cam.cuda(), radius=2)
    loss.backward()
    np.testing.assert_allclose(x.grad.cpu().numpy(), g["affgrad_r2"], rtol=RTOL, atol=1e-10)


def test_seg_loss_golden_value_and_grad(golden_dir):
    g = _g(golden_dir)
    cam, _, pred = training_case()
    p = pred.cuda().requires_grad_(True)
    loss = L.get_seg_loss(p, cam.cuda(), ignore_index=255)
    np.testing.assert_allclose(float(loss), g["segloss"][0], rtol=RTOL)
    loss.backward()
    np.testing.assert_allclose(p.grad[:, :, ::7, ::5].cpu().numpy(), g["seggrad_sample"], rtol=1e-4, atol=1e-9)
    with torch.no_grad():
        np.testing.assert_allclose(float(L.get_seg_loss(pred.cuda(), cam.cuda())), g["segloss"][0], rtol=RTOL)


def test_seg_loss_empty_foreground_is_nan_like_reference():
    pred = torch.randn(1, 5, 16, 16, device="cuda")
    lab = torch.zeros(1, 16, 16, dtype=torch.int64, device="cuda")
    assert torch.isnan(L.get_seg_loss(pred, lab))
    assert torch.isnan(O.seg_loss(pred.cpu(), lab.cpu()))


@pytest.mark.parametrize("B,H,W,C", [(32, 320, 320, 21), (4, 512, 512, 81)])
def test_full_size_against_oracle(B, H, W, C):
    """cfg-2 sized step (32 crops at 320x320) and a COCO-sized one: labels exact, losses / gradients vs the oracle."""
    g = torch.Generator().manual_seed(5)
    cam = torch.randint(0, 3, (B, H // 32, W // 32), generator=g).repeat_interleave(32, 1).repeat_interleave(32, 2) * 7
    cam[:, -40:, :] = 255
    P = (H // 16) * (W // 16)
    attn = torch.rand(B, P, P, generator=g)
    pred = torch.randn(B, C, H, W, generator=g)
    mask = O.mask_by_radius(H // 16, W // 16, 8)
    want = O.affinity_label(cam, mask, 255)
    got = L.cams_to_affinity_label(cam.cuda(), mask=L.radius_mask(H // 16, W // 16, 8), ignore_index=255)
    assert torch.equal(got.cpu(), want)
    wl, wp, wn = O.aff_loss(attn, want)
    x = attn.cuda().requires_grad_(True)
    gl, gp, gn = L.aff_loss_from_labels(x, cam.cuda(), radius=8)
    assert (int(gp), int(gn)) == (int(wp), int(wn))
    np.testing.assert_allclose(float(gl), float(wl), rtol=RTOL)
    gl.backward()
    gref = torch.where(want == 1, -0.5 / wp.double(), torch.where(want == 0, 0.5 / wn.double(), 0.0)).float()
    np.testing.assert_allclose(x.grad.cpu().numpy(), gref.numpy(), rtol=RTOL, atol=1e-12)
    p = pred.cuda().requires_grad_(True)
    sl = L.get_seg_loss(p, cam.cuda())
    np.testing.assert_allclose(float(sl), float(O.seg_loss(pred, cam)), rtol=RTOL)
    sl.backward()
    po = pred[:2].clone().requires_grad_(True)      # oracle gradient on a slice (denominators are global: rescale)
    full = O.seg_loss(pred, cam)
    lab = cam[:2]
    valid = lab != 255
    n_bg, n_fg = ((cam == 0).sum(), ((cam != 0) & (cam != 255)).sum())
    logp = torch.log_softmax(po.double(), 1)
    nll = -logp.gather(1, lab.clamp(0, C - 1).unsqueeze(1)).squeeze(1)
    (0.5 * nll[valid & (lab == 0)].sum() / n_bg + 0.5 * nll[valid & (lab != 0)].sum() / n_fg).backward()
    np.testing.assert_allclose(p.grad[:2].cpu().numpy(), po.grad.numpy(), rtol=1e-4, atol=1e-11)
    assert torch.isfinite(full)

"""Pin the oracle's label-prep / loss restatements against outputs of the real reference
(tests/golden/training.npz, written by oracle/make_golden.py::golden_training).  CPU only."""
import os

import numpy as np
import torch

from oracle import weclip_oracle as O
from oracle.make_golden import training_case


def _g(golden_dir):
    return np.load(os.path.join(golden_dir, "training.npz"), allow_pickle=False)


def test_radius_mask_matches_reference(golden_dir):
    g = _g(golden_dir)
    for r in (2, 8):
        np.testing.assert_array_equal(O.mask_by_radius(6, 7, r), g[f"mask_r{r}"].astype(np.float64))


def test_affinity_label_matches_reference(golden_dir):
    g = _g(golden_dir)
    cam, _, _ = training_case()
    for r in (2, 8):
        got = O.affinity_label(cam, O.mask_by_radius(6, 7, r), 255)
        np.testing.assert_array_equal(got.numpy(), g[f"aff_r{r}"].astype(np.int64))
    np.testing.assert_array_equal(O.affinity_label(cam, None, 255).numpy(), g["aff_nomask"].astype(np.int64))


def test_losses_match_reference(golden_dir):
    g = _g(golden_dir)
    cam, attn, pred = training_case()
    for r in (2, 8):
        loss, pc, nc = O.aff_loss(attn, torch.from_numpy(g[f"aff_r{r}"].astype(np.int64)))
        np.testing.assert_allclose([float(loss)], g[f"affloss_r{r}"][:1], rtol=2e-6)
        assert (int(pc), int(nc)) == tuple(int(v) for v in g[f"affloss_r{r}"][1:])
    np.testing.assert_allclose(float(O.seg_loss(pred, cam, 255)), g["segloss"][0], rtol=2e-6)


def test_oracle_loss_gradients_match_reference(golden_dir):
    g = _g(golden_dir)
    cam, attn, pred = training_case()
    x = attn.clone().requires_grad_(True)
    O.aff_loss(x, torch.from_numpy(g["aff_r2"].astype(np.int64)))[0].backward()
    np.testing.assert_allclose(x.grad.numpy(), g["affgrad_r2"], rtol=1e-5, atol=1e-9)
    p = pred.clone().requires_grad_(True)
    O.seg_loss(p, cam, 255).backward()
    np.testing.assert_allclose(p.grad[:, :, ::7, ::5].numpy(), g["seggrad_sample"], rtol=1e-4, atol=1e-9)

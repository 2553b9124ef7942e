"""Shared parity criteria of the GPU tests (BASELINE.json north_star: >= 99.5 % pixel agreement of argmax
pseudo-labels / predicted masks; 2e-2 max-abs after normalisation for the 16-bit tensor-core path).

Two decisions on the path are DISCONTINUOUS in their inputs and cannot be held to a fixed pixel quota on a
single small image when the inputs carry 16-bit rounding noise (SURVEY.md §7.2 "Mixed-precision parity"):
  * argmax over classes where the reference's own top-2 margin is inside the numeric tolerance (a tie), and
  * scoremap2bbox's uint8 threshold (clip/utils.py:118-121): a Grad-CAM cell whose value sits within one uint8
    step of int(thr*max) may land on either side, which moves a whole bounding box.
The helpers below keep the 99.5 % bar for everything else: ties are excluded only where the REFERENCE margin is
below the tolerance, and a box flip is accepted only if every differing cell is within one uint8 step of the
threshold (checked cell by cell on the thresholded maps) — in which case the labels must still agree >= 99.5 % with
the oracle re-run on the same boxes.  The full-size tests (tests/test_fullsize_gpu.py) and the aggregate test use the
raw 99.5 % with no exclusions.
"""
import numpy as np
import torch

from oracle import weclip_oracle as O

AGREE = 0.995
SEG_TOL = 2e-2


def seg_checks(seg: torch.Tensor, seg_ref: torch.Tensor):
    """Normalised max-abs error < 2e-2 and argmax agreement >= 99.5 % over cells the reference decides by more than
    the tolerance (ties inside the tolerance are numerically undecidable); never below 97 % overall."""
    seg, seg_ref = seg.float().cpu(), seg_ref.float().cpu()
    scale = seg_ref.abs().max()
    err = ((seg - seg_ref).abs().max() / scale).item()
    assert err < SEG_TOL, f"seg max-abs error after normalisation {err}"
    top2 = seg_ref.topk(2, dim=1).values
    decided = (top2[:, 0] - top2[:, 1]) > 2 * err * scale
    same = seg.argmax(1) == seg_ref.argmax(1)
    assert same[decided].float().mean().item() >= AGREE
    assert same.float().mean().item() >= 0.97
    return err


def label_checks(model, cam: torch.Tensor, cam_ref, oracle_call, thr: float):
    """cam: our pseudo-labels [N,Hc,Wc]; cam_ref: the reference's / oracle's.  ``oracle_call(box_masks)`` re-runs the
    oracle forward with the box masks held fixed and returns (labels, intermediates).  The model must have run with
    keep_intermediates=True."""
    cam, cam_ref = cam.cpu(), torch.as_tensor(np.asarray(cam_ref)).long()
    agree = (cam == cam_ref).float().mean().item()
    if agree >= AGREE:
        return agree
    last = model.last
    _, inter = oracle_call(None)
    ours = last["box_masks"].cpu().numpy()
    k_off = last["k_off"]
    flipped = False
    cams_g = last["cams"].cpu().numpy()
    for i, m_o in enumerate(inter["box_masks"]):
        m = ours[k_off[i]:k_off[i + 1]]
        for k in range(m.shape[0]):
            if np.array_equal(m[k], m_o[k]):
                continue
            flipped = True
            # the flip must be numerically justified cell by cell: the binarisations u8 > int(thr * max(u8))
            # (clip/utils.py:118-121) of our Grad-CAM and of the oracle's may only differ in cells whose oracle value lies
            # within (the measured CAM error + one uint8 step) of the threshold — not "some cell somewhere in the map"
            c = inter["cams"][i][k].numpy() if hasattr(inter["cams"][i][k], "numpy") else np.asarray(inter["cams"][i][k])
            g = cams_g[k_off[i] + k].reshape(c.shape)
            err = float(np.abs(g - c).max())
            assert err < SEG_TOL, f"Grad-CAM error {err}"
            u8_o, u8_g = np.floor(c * 255.0), np.floor(g * 255.0)
            t_o, t_g = int(thr * u8_o.max()), int(thr * u8_g.max())
            differ = (u8_o > t_o) != (u8_g > t_g)
            assert differ.any(), "box mask differs although the thresholded maps are identical"
            near = np.abs(c * 255.0 - (t_o + 1)) <= 255.0 * err + 1.0
            assert near[differ].all(), "a thresholded cell differs that is not within the numeric error of the threshold"
    assert flipped, f"pseudo-label agreement {agree} < {AGREE} with identical box masks"
    masks = [ours[k_off[i]:k_off[i + 1]] for i in range(len(inter["box_masks"]))]
    cam_fixed, _ = oracle_call(masks)
    agree2 = (cam == cam_fixed).float().mean().item()
    assert agree2 >= AGREE, f"pseudo-label agreement {agree2} with the box decision held fixed (raw {agree})"
    return agree2

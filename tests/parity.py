"""Shared parity criteria of the GPU tests (BASELINE.json north_star: >= 99.5 % pixel agreement of argmax
pseudo-labels / predicted masks; 2e-2 max-abs after normalisation for the 16-bit tensor-core path).

Three decisions on the path are DISCONTINUOUS in their inputs and cannot be held to a fixed pixel quota on a
single small image when the inputs carry 16-bit rounding noise (SURVEY.md §7.2 "Mixed-precision parity"):
  * argmax over classes where the reference's own top-2 margin is inside the numeric tolerance (a tie), and
  * scoremap2bbox's uint8 threshold (clip/utils.py:118-121): a Grad-CAM cell whose value sits within one uint8
    step of int(thr*max) may land on either side, which moves a whole bounding box, and
  * the RFM layer selection (clip/clip_tool.py:158-165): ``diff_l <= mean(diff)`` on fp32 sums over P^2 terms whose
    differences between layers are a few fp32 ulps of the sums at P >= 1024 — the reference's own decision depends on
    torch's summation order there.
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


def label_checks(model, cam: torch.Tensor, cam_ref, oracle_call, thr: float, per_image: bool = False):
    """cam: our pseudo-labels [N,Hc,Wc]; cam_ref: the reference's / oracle's.  ``oracle_call(box_masks, rfm_masks=None)``
    re-runs the oracle forward with the discontinuous decisions held fixed and returns (labels, intermediates).  The model
    must have run with keep_intermediates=True.

    Raw agreement >= 99.5 % passes outright.  Otherwise every image below the bar must show a discontinuous decision that
    flipped INSIDE its numerical noise, proven cell by cell / layer by layer:
      * box mask (clip/utils.py:118-121): the thresholded uint8 maps of our Grad-CAM and the oracle's differ only in cells
        whose oracle value lies within (measured CAM error + one uint8 step) of the threshold;
      * RFM layer selection (clip_tool.py:158-165, ``diff <= mean(diff)`` on fp32 sums of P^2 ~ 1e5..2.5e6 terms): a layer's
        decision differs only where its fp64 margin |d_l - mean| is within the resolution of the reference's own fp32 sum
        (8 ulp of the sum's magnitude + its distance from the fp64 sum + 4e-5 * P for the attention maps' own tolerance);
    and the labels must then agree >= 99.5 % with the oracle re-run on OUR decisions.
    per_image=True applies the bar image by image (aggregate otherwise).  Returns (raw agreement, images with a flip)."""
    cam, cam_ref = cam.cpu(), torch.as_tensor(np.asarray(cam_ref)).long()
    n_img = cam.shape[0]
    agree_img = (cam == cam_ref).float().flatten(1).mean(1)
    agree = (cam == cam_ref).float().mean().item()
    if (per_image and bool((agree_img >= AGREE).all())) or (not per_image and agree >= AGREE):
        return agree, 0
    last = model.last
    _, inter = oracle_call(None)
    ours = last["box_masks"].cpu().numpy()
    k_off = last["k_off"]
    cams_g = last["cams"].cpu().numpy()
    sel = last.get("rfm_sel")
    sel = sel.cpu().numpy() if sel is not None else None
    flipped = set()
    rfm_fixed = None
    if "rfm" in inter and sel is not None and inter["rfm"] and inter["rfm"][0]:
        rfm_fixed = []
        for i, info in enumerate(inter["rfm"]):
            m_o, m_g = info["mask"].numpy(), sel[i, 0]
            rfm_fixed.append(m_g.copy())
            if np.array_equal(m_o, m_g):
                continue
            flipped.add(i)
            d64 = info["diff64"].numpy()
            # what the decision can resolve: the reference's fp32 sums (8 ulp at their magnitude + their measured distance from
            # the fp64 sums) plus the difference of OUR maps' sums: rows of an attention map sum to 1, so sum(W[1:,1:]) only
            # moves with the class-token column, P entries, each within the measured 4e-5 of the oracle's
            resolution = 8 * np.spacing(np.float32(np.abs(d64).max())) + np.abs(info["diff32"].numpy() - d64).max() + \
                4e-5 * info["P"]
            differ = m_o != m_g
            margin = np.abs(info["margin64"].numpy())
            assert (margin[differ] <= resolution).all(), \
                f"image {i}: RFM layer decision differs outside the fp32 resolution of the reference's sums: " \
                f"margins {margin[differ]}, resolution {resolution}"
    for i, m_o in enumerate(inter["box_masks"]):
        m = ours[k_off[i]:k_off[i + 1]]
        for k in range(m.shape[0]):
            if np.array_equal(m[k], m_o[k]):
                continue
            flipped.add(i)
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
    low = [i for i in range(n_img) if agree_img[i] < AGREE] if per_image else list(range(n_img))
    assert flipped, f"pseudo-label agreement {agree} < {AGREE} with identical box masks and RFM layer selections"
    if per_image:
        assert set(low) <= flipped, f"images {sorted(set(low) - flipped)} are below {AGREE} without any flipped decision"
    masks = [ours[k_off[i]:k_off[i + 1]] for i in range(len(inter["box_masks"]))]
    cam_fixed, _ = oracle_call(masks, rfm_fixed)
    cam_fixed = torch.as_tensor(np.asarray(cam_fixed)).long()
    fixed_img = (cam == cam_fixed).float().flatten(1).mean(1)
    agree2 = (cam == cam_fixed).float().mean().item()
    if per_image:
        assert bool((fixed_img >= AGREE).all()), f"per-image agreement with the decisions held fixed: {fixed_img.tolist()}"
    assert agree2 >= AGREE, f"pseudo-label agreement {agree2} with the box / RFM decisions held fixed (raw {agree})"
    return agree, len(flipped)

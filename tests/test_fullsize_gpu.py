"""Full-size parity on the BASELINE.json configs: the CUDA path against the CPU oracle at the sizes the
bench runs — cfg-3 (msc+flip VOC, 512x512 base -> 384/512/640 pairs, L = 577/1025/1601: CTA-pair GEMMs, 13-tile attention),
cfg-4 (COCO, 81 classes) and cfg-2 (32 crops of 320x320, mode='train', both affinity regimes).

Criteria are the raw north_star ones: logits within 2e-2 max-abs after normalisation, argmax masks agreeing on >= 99.5 % of
pixels with no tie exclusions.  Pseudo-labels: >= 99.5 % raw; an image may fall below that only if one of the path's
discontinuous decisions (box threshold, RFM layer selection) provably flipped inside its numerical resolution, and must
then agree >= 99.5 % with the oracle re-run on those decisions (tests/parity.py::label_checks, per image).  The oracle runs
on the host cores (seconds to ~1 min per case)."""
import os
import tempfile

import numpy as np
import pytest
import torch

from oracle import weclip_oracle as O
from weclip_b200 import synth
from weclip_b200.evaluate import MscFlipEvaluator
from weclip_b200.model import WeCLIP
from tests.parity import label_checks

pytestmark = pytest.mark.gpu

AGREE, TOL = 0.995, 2e-2


def _setup(dataset, n_img, H, W, seed, label_hw=None):
    ncls, n_fg, n_bg, sub = (21, 20, 25, "SegmentationClassAug") if dataset == "voc" else \
        (81, 80, 23, os.path.join("SegmentationClass", "train"))
    sd = synth.make_state_dict(0, ncls)
    fg, bg = synth.make_text_features(7, n_fg, n_bg)
    root = tempfile.mkdtemp()
    rng = np.random.default_rng(seed)
    names, labs = [], []
    for i in range(n_img):
        lab = synth.make_label_map(rng, *(label_hw or (H, W)), synth.draw_label_set(rng, n_fg))
        synth.write_label_png(root, sub, f"f{seed}_{i}", lab)
        names.append(f"f{seed}_{i}"); labs.append(lab)
    imgs = synth.make_images(500 + seed, n_img, H, W)
    model = WeCLIP(num_classes=ncls, dataset_root_path=root, device="cuda", dataset=dataset, state_dict=sd,
                   fg_text_features=fg, bg_text_features=bg)
    model.keep_intermediates = True
    return model, sd, fg, bg, names, labs, imgs, ncls


def _oracle_call(img, labs, sd, fg, bg, mode, it):
    def call(box_masks, rfm_masks=None):
        out = O.weclip_forward(img, list(labs), sd, fg, bg, mode=mode, iter_num=it + 1, return_intermediates=True,
                               box_masks=box_masks, rfm_masks=rfm_masks)
        return out[1], out[3]
    return call


def _norm_err(got, want):
    return ((got.float().cpu() - want).abs().max() / want.abs().max()).item()

def _cam_labels_follow_oracle(model, cam, cam_o, imgs, names, labs, sd, fg, bg, size):
    """The driver's CAM labels (scale-1.0 pass, plain image) against the oracle's: >= 99.5 % raw, or — when one of the path's
    discontinuous decisions (RFM layer selection, box threshold) flipped — proven by label_checks on that pass to have flipped
    inside its numerical noise, with the labels following the oracle given the decision."""
    if (cam.cpu() == cam_o).float().mean().item() >= AGREE:
        return
    from weclip_b200 import ops
    pair = ops.flip_concat(ops.resize_bilinear(imgs.cuda(), size=size))
    _, cam2, _ = model(pair, names + names, mode="val")
    assert (cam2[0].cpu() == cam.cpu().reshape(cam2[0].shape)).float().mean().item() >= 0.998     # same path (fp32 atomics apart)
    pair_h = pair.cpu()
    _, cam_o2, _ = O.weclip_forward(pair_h, labs + labs, sd, fg, bg, mode="val", iter_num=1)
    raw, n_flipped = label_checks(model, cam2, cam_o2, _oracle_call(pair_h, labs + labs, sd, fg, bg, "val", 0), 0.4,
                                  per_image=True)
    print(f"scale-1.0 pass: raw pseudo-label agreement {raw:.4f}, images of the pair with a flipped decision: {n_flipped}")



def test_cfg3_msc_flip_voc_512_matches_oracle():
    """One dataset image of BASELINE configs[2] through the evaluation driver with the reference's call pattern
    (names = name+name: CAM / PAR for both images of every pair at 384 / 512 / 640) against oracle.msc_flip_image."""
    torch.set_num_threads(os.cpu_count() or 1)
    model, sd, fg, bg, names, labs, imgs, C = _setup("voc", 1, 512, 512, seed=41)
    ev = MscFlipEvaluator(model, C, (1.0, 0.75, 1.25), resize_long=512, cam_passes="reference")
    gt = torch.from_numpy(labs[0].astype(np.int16)).cuda()
    seg_pred, msc_pred, cam, msc = ev.process(names[0], imgs.cuda(), gt, want_preds=True)
    seg_o, msc_o, cam_o, msc_segs_o = O.msc_flip_image(imgs, labs[0], sd, fg, bg, (1.0, 0.75, 1.25), 512, "voc")
    assert _norm_err(msc, msc_segs_o[0]) < TOL
    assert (seg_pred.cpu() == seg_o).float().mean().item() >= AGREE
    assert (msc_pred.cpu() == msc_o).float().mean().item() >= AGREE
    _cam_labels_follow_oracle(model, cam, cam_o, imgs, names, labs, sd, fg, bg, (512, 512))
    h = ev.hist.cpu().numpy()
    assert np.array_equal(h[2], O.fast_hist(labs[0], cam.cpu().numpy(), C))        # integer work: exact


def test_cfg3_voc_sized_image_short_side_not_a_multiple_of_16_matches_oracle_eager_and_graph():
    """A dataset image of a real VOC size (333 x 500): the loop resizes it to 340 x 512 and runs 255 x 384 / 340 x 512 /
    425 x 640 pairs — none of the short sides is a multiple of the 16-pixel patch (the conv drops the remainder,
    clip/model.py:267) — with labels at the ORIGINAL 333 x 500.  Eager call against oracle.msc_flip_image, then the same image
    through the CUDA-graph path twice (capture, replay): the same histograms."""
    torch.set_num_threads(os.cpu_count() or 1)
    model, sd, fg, bg, names, labs, imgs, C = _setup("voc", 1, 333, 500, seed=53)
    ev = MscFlipEvaluator(model, C, (1.0, 0.75, 1.25), resize_long=512, cam_passes="reference")
    gt = torch.from_numpy(labs[0].astype(np.int16)).cuda()
    seg_pred, msc_pred, cam, msc = ev.process(names[0], imgs.cuda(), gt, want_preds=True)
    seg_o, msc_o, cam_o, msc_segs_o = O.msc_flip_image(imgs, labs[0], sd, fg, bg, (1.0, 0.75, 1.25), 512, "voc")
    assert tuple(msc.shape[-2:]) == (21, 32) and tuple(seg_pred.shape[-2:]) == (333, 500)
    assert _norm_err(msc, msc_segs_o[0]) < TOL
    assert (seg_pred.cpu() == seg_o).float().mean().item() >= AGREE
    assert (msc_pred.cpu() == msc_o).float().mean().item() >= AGREE
    _cam_labels_follow_oracle(model, cam, cam_o, imgs, names, labs, sd, fg, bg, (340, 512))
    h = ev.hist.cpu().numpy()
    assert np.array_equal(h[2], O.fast_hist(labs[0], cam.cpu().numpy(), C))
    evg = MscFlipEvaluator(model, C, (1.0, 0.75, 1.25), resize_long=512, cam_passes="reference", use_graphs=True,
                           graph_slots=1)
    for _ in range(2):
        evg.process(names[0], imgs.cuda(), gt)
    evg.join()
    torch.cuda.synchronize()
    assert evg.graph_captures == 1 and evg.graph_replays == 2
    hg = evg.hist.cpu().numpy()
    assert np.array_equal(hg, 2 * h)           # bit-identical, CAM pseudo-labels included: no floating-point atomics on the path


def test_cfg4_msc_flip_coco_512_matches_oracle():
    """BASELINE configs[3]: 81 classes; COCO val returns before the CAM path (model_attn_aff_coco.py:131-132)."""
    torch.set_num_threads(os.cpu_count() or 1)
    model, sd, fg, bg, names, labs, imgs, C = _setup("coco", 1, 512, 512, seed=43)
    ev = MscFlipEvaluator(model, C, (1.0, 0.75, 1.25), resize_long=512)
    gt = torch.from_numpy(labs[0].astype(np.int16)).cuda()
    seg_pred, msc_pred, cam, msc = ev.process(names[0], imgs.cuda(), gt, want_preds=True)
    seg_o, msc_o, cam_o, msc_segs_o = O.msc_flip_image(imgs, labs[0], sd, fg, bg, (1.0, 0.75, 1.25), 512, "coco")
    assert cam is None and cam_o is None
    assert _norm_err(msc, msc_segs_o[0]) < TOL
    assert (seg_pred.cpu() == seg_o).float().mean().item() >= AGREE
    assert (msc_pred.cpu() == msc_o).float().mean().item() >= AGREE


@pytest.mark.parametrize("it", [0, 20000])
def test_cfg2_pseudo_label_step_32x320_matches_oracle(it):
    """BASELINE configs[1]: batch 32 at 320x320, mode='train', iter_num <= 15000 (mean of the last 8 maps) and > 15000
    (RFM branch).  Agreement is over all pixels of the batch, no exclusions."""
    torch.set_num_threads(os.cpu_count() or 1)
    model, sd, fg, bg, names, labs, imgs, C = _setup("voc", 32, 320, 320, seed=47 + (it > 0))
    model.iter_num = it
    seg, cam, ap = model(imgs.cuda(), names, mode="train")
    seg_o, cam_o, ap_o = O.weclip_forward(imgs, labs, sd, fg, bg, mode="train", iter_num=it + 1)
    assert _norm_err(seg, seg_o) < TOL
    assert (ap.cpu() - ap_o).abs().max().item() < TOL
    assert (seg.cpu().argmax(1) == seg_o.argmax(1)).float().mean().item() >= AGREE
    raw, n_flipped = label_checks(model, cam, cam_o, _oracle_call(imgs, labs, sd, fg, bg, "train", it), 0.4, per_image=True)
    print(f"cfg2 it={it}: raw pseudo-label agreement {raw:.4f}, images with a flipped box / RFM decision: {n_flipped} of 32")
    # Every flipped decision has been proven, image by image, to lie inside the reference's own numerical noise (label_checks) and
    # the labels of those images agree >= 99.5 % with the oracle re-run on our decisions.  How many of the 32 crops have such a
    # near-tie is a property of the data and of the summation order on either side: 6 to 8 have been observed across kernel
    # revisions that only re-ordered fp32 sums.  The cap guards against a systematic error, which would flip most images.
    assert n_flipped <= 10


@pytest.mark.parametrize("HW", [384, 640])
def test_cfg3_scale_passes_match_oracle(HW):
    """The 0.75x / 1.25x passes of cfg-3 as the reference runs them: an [image, flipped image] pair with names = name+name,
    mode='val', so CAM + affinity refinement + PAR run for both images at L = 577 / 1601 (their labels are discarded by
    the driver, test_msc_flip_voc.py:84, but the reference computes them)."""
    torch.set_num_threads(os.cpu_count() or 1)
    model, sd, fg, bg, names, labs, imgs, C = _setup("voc", 1, HW, HW, seed=41, label_hw=(512, 512))
    pair = torch.cat([imgs, imgs.flip(-1)])
    seg, cam, ap = model(pair.cuda(), names + names, mode="val")
    seg_o, cam_o, ap_o = O.weclip_forward(pair, labs + labs, sd, fg, bg, mode="val", iter_num=1)
    assert _norm_err(seg, seg_o) < TOL
    assert (ap.cpu() - ap_o).abs().max().item() < TOL
    assert (seg.cpu().argmax(1) == seg_o.argmax(1)).float().mean().item() >= AGREE
    raw, n_flipped = label_checks(model, cam, cam_o, _oracle_call(pair, labs + labs, sd, fg, bg, "val", 0), 0.4, per_image=True)
    print(f"cfg3 pass at {HW}: raw pseudo-label agreement {raw:.4f}, images with a flipped decision: {n_flipped} of 2")


@pytest.mark.parametrize("HW", [320, 512, 640])
def test_gradcam_reproduces_the_reference_fp16_backward(HW):
    """The reference's out-projection runs in fp16 on every device (clip/myAtt.py:321) and so does its autograd backward:
    gradients of ~1e-7 are rounded to fp16 SUBNORMALS (2-3 significant bits).  wc_gradcam reproduces the two roundings
    (power-of-two scaled, exact): CAMs within 2e-3 of the reference-exact oracle at every size, where the un-rounded
    analytic gradient is 2e-3 / 1.8e-2 / 3.6e-2 away (L = 401 / 1025 / 1601) — and that one matches the pure-fp32 oracle."""
    from weclip_b200 import _lib
    torch.set_num_threads(os.cpu_count() or 1)
    model, sd, fg, bg, names, labs, imgs, C = _setup("voc", 1, HW, HW, seed=41)
    h = w = HW // 16
    ids = O.label_ids_from_map(labs[0])
    K = len(ids)
    text = torch.cat([fg[ids], bg], 0)
    taps_o, _ = O.encode_image(imgs, sd)
    x11 = taps_o[-1]
    c_ref, _, _ = O.grad_cam(x11, text, list(range(K)), sd, h, w, True)
    c_f32, _, _ = O.grad_cam(x11, text, list(range(K)), sd, h, w, False)
    args = (x11[:, 0].contiguous().cuda(), 1, h, w, text.cuda(), [0, K + 25], [0, K])
    try:
        c_default = model.engine.gradcam(*args)[0].cpu().numpy()
        _lib.lib().wc_gradcam_set_exact(1)
        c_exact = model.engine.gradcam(*args)[0].cpu().numpy()
    finally:
        _lib.lib().wc_gradcam_set_exact(0)
    assert np.abs(c_default - c_ref).max() < 2e-3
    assert np.abs(c_exact - c_f32).max() < 5e-4

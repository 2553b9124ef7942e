"""msc+flip fusion / scoring kernels (csrc/evalops.cu) against torch's own ops and the oracle's restated
evaluation loop (oracle.msc_flip_image <- test_msc_flip_voc.py:60-96)."""
import tempfile

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import weclip_oracle as O
from weclip_b200 import ops, synth
from weclip_b200.evaluate import MscFlipEvaluator
from weclip_b200.model import WeCLIP

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape,size,sf", [((2, 3, 37, 53), (64, 41), None), ((1, 21, 20, 20), (333, 500), None),
                                           ((1, 3, 96, 128), None, 0.75), ((1, 3, 96, 128), None, 1.25),
                                           ((1, 3, 375, 500), (384, 512), None), ((1, 5, 9, 7), (3, 2), None)])
def test_resize_bilinear_matches_torch(shape, size, sf):
    x = torch.randn(*shape, generator=torch.Generator().manual_seed(1))
    want = F.interpolate(x, size=size, scale_factor=sf, mode="bilinear", align_corners=False)
    got = ops.resize_bilinear(x.cuda(), size=size, scale_factor=sf).cpu()
    assert got.shape == want.shape
    assert (got - want).abs().max().item() < 1e-5


def test_flip_concat_is_exact():
    x = torch.randn(1, 3, 45, 67)
    assert torch.equal(ops.flip_concat(x.cuda()).cpu(), torch.cat([x, x.flip(-1)], 0))


def test_msc_accumulate_matches_torch():
    g = torch.Generator().manual_seed(2)
    C, h, w = 21, 20, 27
    segs = [torch.randn(2, C, h, w, generator=g), torch.randn(2, C, 15, 20, generator=g), torch.randn(2, C, 25, 33, generator=g)]
    acc = torch.empty(1, C, h, w, device="cuda")
    lst = []
    for i, s in enumerate(segs):
        ops.msc_accumulate(s.cuda(), acc, 1.0 / 3, i == 0)
        r = F.interpolate(s, size=(h, w), mode="bilinear", align_corners=False) if i else s
        lst.append((r[0] + r[1].flip(-1)) / 2)
    want = torch.stack(lst).mean(0)
    assert (acc[0].cpu() - want).abs().max().item() < 1e-5
    # batched form: [2B,C,hs,ws] = B plain then B flipped
    B = 3
    sb = torch.randn(2 * B, C, 15, 20, generator=g)
    accb = torch.empty(B, C, h, w, device="cuda")
    ops.msc_accumulate(sb.cuda(), accb, 0.5, True)
    r = F.interpolate(sb, size=(h, w), mode="bilinear", align_corners=False)
    wantb = 0.5 * (r[:B] + r[B:].flip(-1)) / 2
    assert (accb.cpu() - wantb).abs().max().item() < 1e-5


@pytest.mark.parametrize("C", [21, 81, 150])
def test_upsample_argmax_hist_matches_torch_and_fast_hist(C):
    g = torch.Generator().manual_seed(C)
    logits = torch.randn(C, 23, 31, generator=g)
    Ho, Wo = 181, 250
    gt = torch.randint(0, C, (Ho, Wo), generator=g).to(torch.int16)
    gt[torch.rand(Ho, Wo, generator=g) < 0.1] = 255            # ignore label
    hist = torch.zeros(C, C, dtype=torch.int64, device="cuda")
    pred = ops.upsample_argmax_hist(logits.cuda(), (Ho, Wo), gt.cuda(), hist).cpu()
    want = F.interpolate(logits[None], size=(Ho, Wo), mode="bilinear", align_corners=False)[0].argmax(0)
    assert (pred == want).float().mean().item() >= 0.9995     # ties / last-bit interpolation differences only
    # integer work is bit-exact: the histogram of OUR predictions equals _fast_hist of the same arrays
    assert np.array_equal(hist.cpu().numpy(), O.fast_hist(gt.numpy(), pred.numpy(), C))
    hist2 = torch.zeros(C, C, dtype=torch.int64, device="cuda")
    ops.confusion_hist(pred.long().cuda(), gt.cuda(), hist2)
    assert torch.equal(hist, hist2)


def test_msc_flip_image_matches_oracle():
    """One dataset image through the whole evaluation loop, small enough for the CPU oracle (long side -> 160)."""
    sd = synth.make_state_dict(0, 21)
    fg, bg = synth.make_text_features(7, 20, 25)
    rng = np.random.default_rng(4)
    lab = synth.make_label_map(rng, 93, 130, [4, 14])
    root = tempfile.mkdtemp()
    synth.write_label_png(root, "SegmentationClassAug", "e0", lab)
    img = synth.make_images(9, 1, 93, 130)
    model = WeCLIP(num_classes=21, dataset_root_path=root, device="cuda", state_dict=sd, fg_text_features=fg,
                   bg_text_features=bg)
    ev = MscFlipEvaluator(model, 21, (1.0, 0.75, 1.25), resize_long=160)
    gt = torch.from_numpy(lab.astype(np.int16)).cuda()
    seg_pred, msc_pred, cam, msc = ev.process("e0", img.cuda(), gt, want_preds=True)
    seg_o, msc_o, cam_o, msc_segs_o = O.msc_flip_image(img, lab, sd, fg, bg, (1.0, 0.75, 1.25), resize_long=160)
    assert ((msc.cpu() - msc_segs_o[0]).abs().max() / msc_segs_o.abs().max()).item() < 2e-2
    assert (seg_pred.cpu() == seg_o).float().mean().item() >= 0.995
    assert (msc_pred.cpu() == msc_o).float().mean().item() >= 0.995
    assert (cam.cpu() == cam_o).float().mean().item() >= 0.995
    h = ev.hist.cpu().numpy()
    assert np.array_equal(h[0], O.fast_hist(lab, seg_pred.cpu().numpy(), 21))
    assert np.array_equal(h[1], O.fast_hist(lab, msc_pred.cpu().numpy(), 21))
    assert np.array_equal(h[2], O.fast_hist(lab, cam.cpu().numpy(), 21))
    sc = ev.scores()
    assert 0.0 <= sc["msc_seg"]["miou"] <= 1.0


def test_cam_passes_live_equals_reference_call_pattern():
    """cam_passes="reference" (names = name+name with every [x, flip(x)] call, test_msc_flip_voc.py:52,69,84: six CAM / PAR
    passes per dataset image) and cam_passes="live" (only the pass whose labels are read, :71) give the same three
    confusion matrices and predictions."""
    sd = synth.make_state_dict(0, 21)
    fg, bg = synth.make_text_features(7, 20, 25)
    rng = np.random.default_rng(12)
    root = tempfile.mkdtemp()
    B, H, W = 2, 96, 128
    labs, names = [], []
    for i in range(B):
        lab = synth.make_label_map(rng, H, W, synth.draw_label_set(rng, 20))
        synth.write_label_png(root, "SegmentationClassAug", f"c{i}", lab)
        labs.append(lab); names.append(f"c{i}")
    img = synth.make_images(14, B, H, W).cuda()
    gt = torch.from_numpy(np.stack(labs).astype(np.int16)).cuda()
    model = WeCLIP(num_classes=21, dataset_root_path=root, device="cuda", state_dict=sd, fg_text_features=fg,
                   bg_text_features=bg)
    calls = []
    fwd = model.forward
    model.forward = lambda x, names_, mode="train": (calls.append(len(names_)), fwd(x, names_, mode=mode))[1]
    ref = MscFlipEvaluator(model, 21, (1.0, 0.75, 1.25), resize_long=128, cam_passes="reference")
    live = MscFlipEvaluator(model, 21, (1.0, 0.75, 1.25), resize_long=128, cam_passes="live")
    assert ref.cam_passes_per_image == 6 and live.cam_passes_per_image == 1
    r = ref.process_batch(names, img, gt, want_preds=True)
    assert calls == [2 * B] * 3
    del calls[:]
    l_ = live.process_batch(names, img, gt, want_preds=True)
    assert sorted(calls) == [0, 0, B]
    assert torch.equal(r[0], l_[0]) and torch.equal(r[1], l_[1]) and torch.equal(r[3], l_[3])
    assert (r[2] == l_[2]).float().mean().item() >= 0.999        # fp atomics order on the CAM path
    assert torch.equal(ref.hist[:2], live.hist[:2])
    assert int((ref.hist[2] - live.hist[2]).abs().sum()) <= 2e-3 * int(ref.hist[2].sum())


def test_msc_flip_matches_reference_validate_golden(golden_dir):
    """The evaluation driver against the REAL validate() loop of test_msc_flip_voc.py:30-113 (tests/golden/mscflip_voc.npz,
    generated by oracle/make_golden.py::golden_mscflip): two dataset images of different sizes (short sides not multiples
    of 16 after the long-side resize), one image per call as the reference's DataLoader(batch_size=1) does, plus the
    .npy logit dump of :111."""
    import os
    from oracle.make_golden import MSCFLIP_CASE, mscflip_case
    g = np.load(os.path.join(golden_dir, "mscflip_voc.npz"))
    c = MSCFLIP_CASE
    sd = synth.make_state_dict(c["weights_seed"], 21)
    fg, bg = synth.make_text_features(c["text_seed"], 20, 25)
    root, dump = tempfile.mkdtemp(), tempfile.mkdtemp()
    items = mscflip_case()
    for name, _, lab in items:
        synth.write_label_png(root, "SegmentationClassAug", name, lab)
    model = WeCLIP(num_classes=21, dataset_root_path=root, device="cuda", state_dict=sd, fg_text_features=fg,
                   bg_text_features=bg)
    ev = MscFlipEvaluator(model, 21, c["scales"], resize_long=c["resize_long"], logit_dir=dump)
    agree = np.zeros(3); total = 0
    for i, (name, img, lab) in enumerate(items):
        gt = torch.from_numpy(lab.astype(np.int16)).cuda()
        seg_pred, msc_pred, cam, msc = ev.process(name, img[None].cuda(), gt, want_preds=True)
        ref_msc = torch.from_numpy(g[f"msc_segs{i}"])
        assert ((msc.cpu() - ref_msc[0]).abs().max() / ref_msc.abs().max()).item() < 2e-2
        for j, (got, key) in enumerate(((seg_pred, "seg_pred"), (msc_pred, "msc_pred"), (cam, "cam"))):
            agree[j] += (got.cpu().numpy() == g[f"{key}{i}"]).sum()
        total += lab.size
    assert (agree / total >= 0.995).all(), agree / total
    h = ev.hist.cpu().numpy()
    assert h.sum(axis=(1, 2)).tolist() == g["hist"].sum(axis=(1, 2)).tolist()        # same valid-pixel counts, exactly
    assert (np.abs(h - g["hist"]).sum(axis=(1, 2)) <= 2 * 0.005 * g["hist"].sum(axis=(1, 2))).all()
    ev.scores()                                                                       # drains the logit dumper
    for i, (name, _, _) in enumerate(items):
        d = np.load(os.path.join(dump, name + ".npy"), allow_pickle=True).item()       # the CRF stage's read, :140-141
        assert set(d) == {"segs", "msc_segs"} and d["segs"].shape == g[f"segs{i}"].shape
        assert d["msc_segs"].dtype == np.float32 and d["msc_segs"].shape == g[f"msc_segs{i}"].shape
        assert np.abs(d["segs"] - g[f"segs{i}"]).max() / np.abs(g[f"segs{i}"]).max() < 2e-2
        assert np.abs(d["msc_segs"] - g[f"msc_segs{i}"]).max() / np.abs(g[f"msc_segs{i}"]).max() < 2e-2


def test_batched_bucket_equals_image_by_image():
    """A (H,W) bucket of B dataset images run as one [2B,...] forward per scale gives the same three confusion
    matrices as the reference's image-by-image loop (images are independent)."""
    sd = synth.make_state_dict(0, 21)
    fg, bg = synth.make_text_features(7, 20, 25)
    rng = np.random.default_rng(6)
    root = tempfile.mkdtemp()
    B, H, W = 3, 96, 128
    labs, names = [], []
    for i in range(B):
        lab = synth.make_label_map(rng, H, W, synth.draw_label_set(rng, 20))
        synth.write_label_png(root, "SegmentationClassAug", f"q{i}", lab)
        labs.append(lab); names.append(f"q{i}")
    img = synth.make_images(10, B, H, W).cuda()
    gt = torch.from_numpy(np.stack(labs).astype(np.int16)).cuda()
    model = WeCLIP(num_classes=21, dataset_root_path=root, device="cuda", state_dict=sd, fg_text_features=fg,
                   bg_text_features=bg)
    one = MscFlipEvaluator(model, 21, (1.0, 0.75, 1.25), resize_long=128)
    for i in range(B):
        one.process(names[i], img[i:i + 1], gt[i])
    bat = MscFlipEvaluator(model, 21, (1.0, 0.75, 1.25), resize_long=128)
    bat.process_batch(names, img, gt)
    assert bat.n_images == one.n_images == B
    assert torch.equal(one.hist[:2], bat.hist[:2])               # segmentation histograms: bit-identical
    # the CAM path reduces with floating-point atomics (affinity.cu, gradcam.cu): near-tie pixels may differ with the batch shape
    assert int(one.hist[2].sum()) == int(bat.hist[2].sum())
    assert int((one.hist[2] - bat.hist[2]).abs().sum()) <= 2e-3 * int(one.hist[2].sum())
    assert int(bat.hist[0].sum()) == int((gt >= 0).logical_and(gt < 21).sum())


def test_scales_on_separate_streams_equal_sequential():
    """The three scale passes run on their own CUDA streams with the fusion tail on a fourth and no wait on the caller's
    stream between images; segmentation histograms and predictions are bit-identical to the sequential single-stream
    driver, also when several images are in flight before anything is read."""
    sd = synth.make_state_dict(0, 21)
    fg, bg = synth.make_text_features(7, 20, 25)
    rng = np.random.default_rng(16)
    root = tempfile.mkdtemp()
    B, H, W = 4, 96, 128
    labs, names = [], []
    for i in range(B):
        lab = synth.make_label_map(rng, H, W, synth.draw_label_set(rng, 20))
        synth.write_label_png(root, "SegmentationClassAug", f"s{i}", lab)
        labs.append(lab); names.append(f"s{i}")
    img = synth.make_images(11, B, H, W).cuda()
    gt = torch.from_numpy(np.stack(labs).astype(np.int16)).cuda()
    model = WeCLIP(num_classes=21, dataset_root_path=root, device="cuda", state_dict=sd, fg_text_features=fg,
                   bg_text_features=bg)
    seq = MscFlipEvaluator(model, 21, (1.0, 0.75, 1.25), resize_long=128, overlap_scales=False)
    ovl = MscFlipEvaluator(model, 21, (1.0, 0.75, 1.25), resize_long=128, overlap_scales=True)
    assert ovl.overlap_scales and not seq.overlap_scales
    want = [seq.process(names[i], img[i:i + 1], gt[i], want_preds=True) for i in range(B)]
    for rep in range(3):                      # images back to back, nothing read in between
        for i in range(B):
            ovl.process(names[i], img[i:i + 1].clone(), gt[i].clone())
    got = [ovl.process(names[i], img[i:i + 1], gt[i], want_preds=True) for i in range(B)]
    # segmentation outputs (GEMM / attention only) are bit-identical.  The CAM pseudo-label path reduces with floating-point
    # atomics (affinity.cu, gradcam.cu): their order — hence the last bits of a near-tie pixel — depends on what else is
    # running, so labels are compared at the domain tolerance (>= 99.9 % here; the parity bar is 99.5 %).
    for w_, g_ in zip(want, got):
        seg_w, msc_pred_w, cam_w, msc_w = w_
        seg_g, msc_pred_g, cam_g, msc_g = g_
        assert torch.equal(seg_w, seg_g) and torch.equal(msc_pred_w, msc_pred_g) and torch.equal(msc_w, msc_g)
        assert (cam_w == cam_g).float().mean().item() >= 0.999
    hs, ho = seq.hist, ovl.hist
    assert torch.equal(ho[:2], 4 * hs[:2])
    assert int(ho[2].sum()) == 4 * int(hs[2].sum())
    assert int((ho[2] - 4 * hs[2]).abs().sum()) <= 2e-3 * int(ho[2].sum())
    ovl.reset()
    assert int(ovl.hist.sum()) == 0


def _small_dataset(root, seed=21):
    """Six dataset images of three different sizes (short sides not multiples of 16 after the resize) and varying K."""
    rng = np.random.default_rng(seed)
    items = []
    for i, (h0, w0) in enumerate([(93, 130), (120, 88), (93, 130), (101, 135), (120, 88), (93, 130)]):
        lab = synth.make_label_map(rng, h0, w0, synth.draw_label_set(rng, 20))
        synth.write_label_png(root, "SegmentationClassAug", f"g{i}", lab)
        items.append((f"g{i}", synth.make_images(40 + i, 1, h0, w0)[0], lab))
    return items


def test_cuda_graph_buckets_equal_the_eager_driver():
    """use_graphs: one CUDA graph per (image size, label size, classes-per-image) bucket and replay slot holds the whole
    step; replays (incl. re-used buckets, both slots, several images in flight) give the eager driver's histograms —
    bit-identical, CAM pseudo-labels included (every reduction on the path has a fixed order)."""
    sd = synth.make_state_dict(0, 21)
    fg, bg = synth.make_text_features(7, 20, 25)
    root = tempfile.mkdtemp()
    items = _small_dataset(root)
    model = WeCLIP(num_classes=21, dataset_root_path=root, device="cuda", state_dict=sd, fg_text_features=fg,
                   bg_text_features=bg)
    eager = MscFlipEvaluator(model, 21, (1.0, 0.75, 1.25), resize_long=160, overlap_scales=False)
    graphed = MscFlipEvaluator(model, 21, (1.0, 0.75, 1.25), resize_long=160, use_graphs=True)
    assert graphed.use_graphs
    for rep in range(2):
        for name, img, lab in items:
            gt = torch.from_numpy(lab.astype(np.int16)).cuda()
            eager.process(name, img[None].cuda(), gt)
            graphed.process(name, img[None].cuda().clone(), gt.clone())
    he, hg = eager.hist, graphed.hist
    assert graphed.n_images == eager.n_images == 12 and graphed.graph_replays == 12
    assert 3 <= graphed.graph_captures <= 12              # buckets x 2 slots at most; far fewer captures than images
    assert torch.equal(he, hg)                # all three histograms: the path has no floating-point atomics
    before = graphed.graph_captures
    for name, img, lab in items:                          # third pass: every bucket / slot pair that recurs is a pure replay
        graphed.process(name, img[None].cuda(), torch.from_numpy(lab.astype(np.int16)).cuda())
    assert graphed.graph_captures <= before + 3
    assert torch.equal(graphed.hist[:2] * 2, he[:2] * 3)


def test_validate_with_pinned_ring_and_graphs_matches_eager():
    """evaluate.validate(): one image per call from a pinned input ring, CUDA graphs per bucket, final all-reduce."""
    from weclip_b200.evaluate import validate
    sd = synth.make_state_dict(0, 21)
    fg, bg = synth.make_text_features(7, 20, 25)
    root = tempfile.mkdtemp()
    items = _small_dataset(root, seed=22)
    model = WeCLIP(num_classes=21, dataset_root_path=root, device="cuda", state_dict=sd, fg_text_features=fg,
                   bg_text_features=bg)
    dataset = [(n, im.numpy(), lab.astype(np.int64), None) for n, im, lab in items]
    got = validate(model, dataset, test_scales=[1.0, 0.75], resize_long=160, use_graphs=True)
    want = validate(model, dataset, test_scales=[1.0, 0.75], resize_long=160, use_graphs=False)
    for k in ("seg", "msc_seg"):
        assert got[k]["miou"] == want[k]["miou"] and got[k]["pAcc"] == want[k]["pAcc"]
    assert abs(got["cam"]["pAcc"] - want["cam"]["pAcc"]) < 2e-3

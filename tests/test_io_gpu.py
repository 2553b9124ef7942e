"""SURVEY.md §8f-4 on the device: csrc/jpeg.cu (de-quantise + islow IDCT + fancy up-sampling + fixed-point colour conversion
+ fp64-exact normalisation) returns PIL's pixels BIT FOR BIT and the reference loader's normalised CHW tensor exactly;
label sets from a device label map; the dataset feeds evaluate.validate."""
import io
import os
import tempfile

import numpy as np
import pytest
import torch
from PIL import Image

from weclip_b200 import dataio, synth
from tests.test_io_cpu import CASES, make_jpeg, pil_pixels

pytestmark = pytest.mark.gpu


def reference_normalize(arr: np.ndarray) -> np.ndarray:
    """datasets/transforms.py:8-15 + np.transpose (voc.py:236), verbatim arithmetic."""
    proc = np.empty_like(arr, np.float32)
    for c, (m, s) in enumerate(zip([123.675, 116.28, 103.53], [58.395, 57.12, 57.375])):
        proc[..., c] = (arr[..., c] - m) / s
    return np.transpose(proc, (2, 0, 1))


@pytest.mark.parametrize("sub,q,hw", CASES)
def test_decode_jpeg_is_bit_exact_vs_pil(sub, q, hw):
    data = make_jpeg(*hw, sub, q, seed=q)
    chw, rgb = dataio.decode_jpeg(data, "cuda", normalize=True, want_rgb=True)
    want = pil_pixels(data)
    assert np.array_equal(rgb.cpu().numpy(), want)                               # integer work: every pixel
    assert np.array_equal(chw.cpu().numpy(), reference_normalize(want))          # fp64 division rounded once, like numpy


def test_greyscale_and_restart_markers_on_device():
    for kw in (dict(grey=True), dict(restart_marker_rows=1), dict(optimize=True)):
        data = make_jpeg(101, 77, 2, 70, seed=5, **kw)
        rgb = dataio.decode_jpeg(data, "cuda", normalize=False, want_rgb=True)
        assert np.array_equal(rgb.cpu().numpy(), pil_pixels(data))


def test_normalize_and_label_set_kernels():
    rng = np.random.default_rng(0)
    arr = rng.integers(0, 256, (53, 71, 3)).astype(np.uint8)
    got = dataio.normalize_u8(torch.from_numpy(arr).cuda()).cpu().numpy()
    assert np.array_equal(got, reference_normalize(arr))
    lab = synth.make_label_map(rng, 90, 120, [2, 9, 19])                         # has 0 (bg) and a 255 stripe
    ids = dataio.label_ids(torch.from_numpy(lab).cuda())
    ref = (np.unique(lab) - 1).tolist()                                          # clip_tool.py:113-118
    assert ids == [int(i) for i in ref if i not in (254, 255)] == [2, 9, 19]
    assert dataio.label_ids(torch.zeros(8, 8, dtype=torch.uint8, device="cuda")) == []


def test_dataset_feeds_validate_like_the_reference_loader():
    """VOC-layout directory (JPEGImages/*.jpg, SegmentationClassAug/*.png, val.txt): the device-decoding dataset and the
    reference-style host loader (PIL + numpy) give the same items, and evaluate.validate the same scores from both."""
    from weclip_b200.evaluate import validate
    from weclip_b200.model import WeCLIP
    root = tempfile.mkdtemp()
    os.makedirs(os.path.join(root, "JPEGImages"))
    rng = np.random.default_rng(3)
    names = []
    for i, (h, w) in enumerate([(93, 130), (120, 88), (93, 130)]):
        name = f"2007_{i:06d}"
        with open(os.path.join(root, "JPEGImages", name + ".jpg"), "wb") as f:
            f.write(make_jpeg(h, w, 2, 90, seed=i))
        synth.write_label_png(root, "SegmentationClassAug", name, synth.make_label_map(rng, h, w, synth.draw_label_set(rng, 20)))
        names.append(name)
    with open(os.path.join(root, "val.txt"), "w") as f:
        f.write("\n".join(names) + "\n")
    dev_ds = dataio.VOC12SegDataset(root, name_list_dir=root, split="val", device="cuda")
    host_ds = dataio.VOC12SegDataset(root, name_list_dir=root, split="val", device=None)
    assert len(dev_ds) == 3
    for a, b in zip(dev_ds, host_ds):
        assert a[0] == b[0] and np.array_equal(a[1].cpu().numpy(), b[1]) and np.array_equal(a[2], b[2])
    sd = synth.make_state_dict(0, 21)
    fg, bg = synth.make_text_features(7, 20, 25)
    model = WeCLIP(num_classes=21, dataset_root_path=root, device="cuda", state_dict=sd, fg_text_features=fg, bg_text_features=bg)
    s_dev = validate(model, [(n, im.cpu().numpy(), lab, c) for n, im, lab, c in dev_ds], test_scales=[1.0, 0.75], resize_long=160)
    s_host = validate(model, list(host_ds), test_scales=[1.0, 0.75], resize_long=160)
    assert s_dev["seg"]["miou"] == s_host["seg"]["miou"] and s_dev["msc_seg"]["pAcc"] == s_host["msc_seg"]["pAcc"]

"""SURVEY.md §8f-4, host side: the library's JPEG marker parser + sequential Huffman decoder (C, no GPU needed) and the
oracle's numpy restatement of libjpeg's reconstruction, pinned bit for bit against the pixels PIL returns (= what
imageio.imread returns in datasets/voc.py:50)."""
import io

import numpy as np
import pytest
from PIL import Image

from oracle import jpeg_oracle as J
from weclip_b200 import _lib, dataio, synth

CASES = [(0, 95, (64, 80)), (2, 75, (375, 500)), (2, 90, (333, 500)), (1, 85, (121, 67)), (2, 50, (17, 23)), (0, 75, (8, 8)),
         (2, 95, (500, 375)), (1, 60, (48, 100))]


def make_jpeg(h, w, subsampling, quality, seed=0, grey=False, **kw) -> bytes:
    img = synth.make_images(seed + h + w, 1, h, w)[0].numpy()
    arr = np.clip(img.transpose(1, 2, 0) * 50 + 120 + np.random.default_rng(seed).normal(0, 6, (h, w, 3)), 0, 255).astype(np.uint8)
    im = Image.fromarray(arr[..., 0] if grey else arr)
    buf = io.BytesIO()
    if grey:
        im.save(buf, "JPEG", quality=quality, **kw)
    else:
        im.save(buf, "JPEG", quality=quality, subsampling=subsampling, **kw)
    return buf.getvalue()


def pil_pixels(data: bytes) -> np.ndarray:
    return np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))


@pytest.fixture(scope="module", autouse=True)
def _built():
    _lib.build()


@pytest.mark.parametrize("sub,q,hw", CASES)
def test_entropy_decode_plus_oracle_reconstruction_equals_pil(sub, q, hw):
    data = make_jpeg(*hw, sub, q, seed=q)
    info, coef = dataio.jpeg_entropy_decode(data, pinned=False)
    assert (info.height, info.width, info.ncomp) == (hw[0], hw[1], 3)
    assert (info.hs[0], info.vs[0]) == {0: (1, 1), 1: (2, 1), 2: (2, 2)}[sub]
    got = J.reconstruct(info, coef.numpy())
    assert np.array_equal(got, pil_pixels(data))


def test_greyscale_restart_intervals_and_optimised_tables():
    for kw in (dict(grey=True), dict(optimize=True), dict(optimize=True, grey=True), dict(restart_marker_blocks=3),
               dict(restart_marker_rows=1), dict(restart_marker_rows=2, grey=True)):
        data = make_jpeg(70, 90, 2, 80, seed=3, **kw)
        info, coef = dataio.jpeg_entropy_decode(data, pinned=False)
        if "restart_marker_blocks" in kw or "restart_marker_rows" in kw:
            assert info.restart_interval > 0
        assert np.array_equal(J.reconstruct(info, coef.numpy()), pil_pixels(data))


def test_unsupported_streams_fail_loudly():
    data = make_jpeg(40, 40, 2, 80, progressive=True)
    with pytest.raises(_lib.WeclipError, match="progressive"):
        dataio.jpeg_entropy_decode(data, pinned=False)
    with pytest.raises(_lib.WeclipError):
        dataio.jpeg_entropy_decode(b"not a jpeg at all", pinned=False)
    good = make_jpeg(40, 40, 2, 80)
    with pytest.raises(_lib.WeclipError):
        dataio.jpeg_entropy_decode(good[: len(good) // 3], pinned=False)      # truncated scan


def test_logit_dump_format(tmp_path):
    """test_msc_flip_voc.py:111 / :140-141: a pickled dict with 'segs' and 'msc_segs' read back through .item()."""
    segs, msc = np.random.rand(1, 21, 7, 10).astype(np.float32), np.random.rand(1, 21, 7, 10).astype(np.float32)
    p = str(tmp_path / "x.npy")
    dataio.save_logits(p, segs, msc)
    d = np.load(p, allow_pickle=True).item()
    assert set(d) == {"segs", "msc_segs"} and np.array_equal(d["segs"], segs) and np.array_equal(d["msc_segs"], msc)

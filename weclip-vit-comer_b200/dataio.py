"""Input pipeline on the device: JPEG -> normalised CHW tensor, label PNG -> label map / label set,
the per-image ``.npy`` logit dump, and a VOC12SegDataset-shaped dataset that feeds ``evaluate.validate``.

Mirrors datasets/voc.py:25-67,183-271 (``VOC12SegDataset(stage='val', aug=False)`` -> ``(name, image CHW fp32, label HW,
cls_label)``), datasets/transforms.py:8-15 (``normalize_img``) and clip/clip_tool.py:111-118 (label ids from the PNG).
The serial entropy decode of a JPEG runs on a host thread (C, inside libweclip_sm100.so); de-quantisation, inverse DCT,
chroma up-sampling, colour conversion and normalisation run as CUDA kernels (csrc/jpeg.cu) and return exactly PIL's pixels.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional, Tuple

import numpy as np
import torch

from . import _lib

IMAGENET_MEAN = (123.675, 116.28, 103.53)        # datasets/transforms.py:8
IMAGENET_STD = (58.395, 57.12, 57.375)


class JpegInfo(ctypes.Structure):
    _fields_ = [("width", ctypes.c_int), ("height", ctypes.c_int), ("ncomp", ctypes.c_int),
                ("hs", ctypes.c_int * 3), ("vs", ctypes.c_int * 3), ("hmax", ctypes.c_int), ("vmax", ctypes.c_int),
                ("mcus_x", ctypes.c_int), ("mcus_y", ctypes.c_int),
                ("blocks_w", ctypes.c_int * 3), ("blocks_h", ctypes.c_int * 3),
                ("coef_off", ctypes.c_long * 3), ("coef_count", ctypes.c_long),
                ("comp_qt", ctypes.c_int * 3), ("restart_interval", ctypes.c_int),
                ("qt", (ctypes.c_uint16 * 64) * 4)]


def jpeg_entropy_decode(data: bytes, pinned: bool = True, out: torch.Tensor = None) -> Tuple[JpegInfo, torch.Tensor]:
    """HOST half: parse the stream and Huffman-decode it -> (info, int16 coefficients [coef_count], pinned host tensor).
    ``out``: a caller-owned (pinned) int16 buffer to decode into — allocating pinned memory costs more than decoding a VOC-sized
    image, so a loader keeps a small ring of them; the returned tensor is then a view of its first ``coef_count`` elements."""
    L = _lib.lib()
    buf = (ctypes.c_uint8 * len(data)).from_buffer_copy(data)
    info = JpegInfo()
    _lib.check(L.wc_jpeg_parse(buf, ctypes.c_size_t(len(data)), ctypes.byref(info)), "wc_jpeg_parse")
    if out is not None:
        if out.dtype != torch.int16 or out.numel() < int(info.coef_count) or not out.is_contiguous():
            raise ValueError("out must be a contiguous int16 tensor of at least coef_count elements")
        coef = out.view(-1)[: int(info.coef_count)]
    else:
        coef = torch.empty(int(info.coef_count), dtype=torch.int16, pin_memory=pinned and torch.cuda.is_available())
    rc = L.wc_jpeg_entropy_decode(buf, ctypes.c_size_t(len(data)), ctypes.c_void_p(coef.data_ptr()), ctypes.c_long(coef.numel()))
    _lib.check(rc, "wc_jpeg_entropy_decode")
    return info, coef


def decode_jpeg(data, device="cuda", normalize: bool = True, want_rgb: bool = False, mean=IMAGENET_MEAN, std=IMAGENET_STD):
    """JPEG file (path or bytes) -> fp32 [3,H,W] = (RGB - mean) / std on ``device`` (and / or uint8 [H,W,3] = PIL's pixels).
    Returns ``chw``, ``rgb`` or ``(chw, rgb)``."""
    _lib.require_device()
    if isinstance(data, (str, os.PathLike)):
        with open(data, "rb") as f:
            data = f.read()
    info, coef_h = jpeg_entropy_decode(data)
    L = _lib.lib()
    dev = torch.device(device)
    coef = coef_h.to(dev, non_blocking=True)
    H, W = info.height, info.width
    chw = torch.empty(3, H, W, dtype=torch.float32, device=dev) if normalize else None
    rgb = torch.empty(H, W, 3, dtype=torch.uint8, device=dev) if want_rgb else None
    ws = torch.empty(int(L.wc_jpeg_workspace_bytes(ctypes.byref(info))), dtype=torch.uint8, device=dev)
    m, s = (ctypes.c_double * 3)(*mean), (ctypes.c_double * 3)(*std)
    rc = L.wc_jpeg_reconstruct(ctypes.byref(info), _lib.ptr(coef), _lib.ptr(rgb), _lib.ptr(chw), m, s, _lib.ptr(ws),
                               ctypes.c_size_t(ws.numel()), _lib.stream_ptr())
    _lib.check(rc, "wc_jpeg_reconstruct")
    if normalize and want_rgb:
        return chw, rgb
    return chw if normalize else rgb


def normalize_u8(img_hwc: torch.Tensor, mean=IMAGENET_MEAN, std=IMAGENET_STD) -> torch.Tensor:
    """uint8 [H,W,3] on the device -> fp32 [3,H,W] (normalize_img + transpose, datasets/transforms.py:8-15, voc.py:236)."""
    _lib.require_device()
    img_hwc = img_hwc.contiguous()
    H, W, _ = img_hwc.shape
    out = torch.empty(3, H, W, dtype=torch.float32, device=img_hwc.device)
    m, s = (ctypes.c_double * 3)(*mean), (ctypes.c_double * 3)(*std)
    _lib.check(_lib.lib().wc_normalize_u8(_lib.ptr(img_hwc), _lib.ptr(out), H, W, m, s, _lib.stream_ptr()), "wc_normalize_u8")
    return out


def label_ids(label_u8: torch.Tensor) -> List[int]:
    """clip/clip_tool.py:111-118 on a uint8 label map that already lives on the device: np.unique, uint8 '-1' (0 -> 255
    background, 255 -> 254 ignore), both dropped -> sorted foreground class ids.  (Reads 1 KB back to the host.)"""
    _lib.require_device()
    label_u8 = label_u8.contiguous()
    present = torch.empty(256, dtype=torch.int32, device=label_u8.device)
    _lib.check(_lib.lib().wc_label_presence(_lib.ptr(label_u8), ctypes.c_long(label_u8.numel()), _lib.ptr(present),
                                            _lib.stream_ptr()), "wc_label_presence")
    vals = torch.nonzero(present).flatten().tolist()
    ids = [(v - 1) % 256 for v in vals]
    return sorted(i for i in ids if i not in (254, 255))


def read_label_png(path: str) -> np.ndarray:
    """imageio.imread(label_dir) of datasets/voc.py:56-62 (palette PNG -> uint8 class indices)."""
    from PIL import Image
    return np.asarray(Image.open(path))


def save_logits(path: str, segs: np.ndarray, msc_segs: np.ndarray) -> None:
    """test_msc_flip_voc.py:111: np.save(<work_dir>/logit/<name>.npy, {"segs": ..., "msc_segs": ...}) (read back with
    np.load(..., allow_pickle=True).item() by the CRF stage, :140-141)."""
    np.save(path, {"segs": np.asarray(segs, np.float32), "msc_segs": np.asarray(msc_segs, np.float32)})


class VOC12SegDataset:
    """datasets/voc.py:183-271 for stage='val', aug=False: item = (name, image CHW fp32 normalised, label HW uint8,
    cls_label).  ``device`` selects where the image is produced: "cuda" decodes through csrc/jpeg.cu (the image tensor is
    then already resident for ``evaluate.validate``), None / "cpu" returns what the reference returns (PIL + numpy)."""

    def __init__(self, root_dir: str, name_list_dir: Optional[str] = None, split: str = "val", names=None, device="cuda",
                 cls_labels: Optional[dict] = None):
        self.img_dir = os.path.join(root_dir, "JPEGImages")
        self.label_dir = os.path.join(root_dir, "SegmentationClassAug")
        if names is None:
            with open(os.path.join(name_list_dir, split + ".txt")) as f:          # datasets/voc.py:14-16
                names = [x for x in f.read().split("\n") if x]
        self.name_list = list(names)
        self.device = device
        self.cls_labels = cls_labels or {}

    def __len__(self):
        return len(self.name_list)

    def __getitem__(self, idx):
        name = self.name_list[idx]
        label = read_label_png(os.path.join(self.label_dir, name + ".png"))
        path = os.path.join(self.img_dir, name + ".jpg")
        if self.device is not None and torch.device(self.device).type == "cuda":
            image = decode_jpeg(path, self.device)
        else:
            from PIL import Image
            arr = np.asarray(Image.open(path).convert("RGB"))
            proc = np.empty(arr.shape, np.float32)
            for c in range(3):
                proc[..., c] = (arr[..., c] - IMAGENET_MEAN[c]) / IMAGENET_STD[c]
            image = np.transpose(proc, (2, 0, 1))
        return name, image, label, self.cls_labels.get(name, 0)

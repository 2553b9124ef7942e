"""msc+flip evaluation driver — the caller that defines "one dataset image" for the metric.

Mirrors ``validate(model, dataset, test_scales)`` of test_msc_flip_voc.py:32-113 (COCO twin
test_msc_flip_coco.py:33-116) and ``utils/evaluate.py:9-36``: long-side-512 resize, [x, flip(x)] batch per
scale, flip-average, scale-mean, resize to the label size, argmax, three confusion matrices (single-scale
seg, msc seg, CAM pseudo-label).  All arithmetic runs in libweclip_sm100.so (csrc/evalops.cu); the three
C x C int64 histograms stay on the device and are summed over ranks with ONE all-reduce at the end
(images are sharded round-robin over ranks, no data-path collective).
"""
from __future__ import annotations

from typing import Iterable, Optional, Sequence

import numpy as np
import torch

from . import ops


def scores(hist: np.ndarray) -> dict:
    """utils/evaluate.py:24-36 on an accumulated confusion matrix."""
    hist = np.asarray(hist, dtype=np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        acc = np.diag(hist).sum() / hist.sum()
        acc_cls = np.nanmean(np.diag(hist) / hist.sum(axis=1))
        iu = np.diag(hist) / (hist.sum(axis=1) + hist.sum(axis=0) - np.diag(hist))
        valid = hist.sum(axis=1) > 0
        return {"pAcc": acc, "mAcc": acc_cls, "miou": np.nanmean(iu[valid]), "iou": dict(enumerate(iu))}


def shard_indices(n_items: int, rank: int, world: int) -> range:
    """Round-robin image sharding (the reference's own multi-GPU idiom, clip/generate_cams_voc12.py:96-100)."""
    return range(rank, n_items, world)


class MscFlipEvaluator:
    def __init__(self, model, num_classes: int = 21, test_scales: Sequence[float] = (1.0, 0.75, 1.25),
                 resize_long: int = 512, device="cuda"):
        self.model = model
        self.C = num_classes
        self.scales = list(test_scales)
        self.resize_long = resize_long
        self.device = torch.device(device)
        self.hist = torch.zeros(3, num_classes, num_classes, dtype=torch.int64, device=self.device)  # seg, msc, cam
        self.n_images = 0

    def reset(self):
        self.hist.zero_()
        self.n_images = 0

    def process(self, name: str, inputs: torch.Tensor, labels: torch.Tensor, want_preds: bool = False):
        """One dataset image (test_msc_flip_voc.py:50-101).  inputs fp32 [1,3,H0,W0]; labels int16 [Hl,Wl]
        (both already on the device).  Returns (seg_pred, msc_pred, cam, msc_segs) when want_preds."""
        out = self.process_batch([name], inputs, labels.unsqueeze(0), want_preds)
        if want_preds:
            seg_pred, msc_pred, cam, msc = out
            return seg_pred[0], msc_pred[0], (cam[0] if cam is not None else None), msc[0]
        return None

    def process_batch(self, names: Sequence[str], inputs: torch.Tensor, labels: torch.Tensor, want_preds: bool = False):
        """B same-size dataset images per model call (the reference loops image by image; images are independent,
        so a (H,W) bucket is run as one [2B,3,H,W] forward per scale: B plain images followed by their B flipped
        twins).  inputs fp32 [B,3,H0,W0]; labels int16 [B,Hl,Wl]."""
        B = inputs.shape[0]
        names = list(names)      # B names for 2B images: the CAM / PAR path runs for the B plain images only, exactly as the
                                 # reference's one-name call does for its [image, flip] pair (model_attn_aff_voc.py:139)
        _, _, h0, w0 = inputs.shape
        ratio = self.resize_long / max(h0, w0)
        base = ops.resize_bilinear(inputs, size=(int(h0 * ratio), int(w0 * ratio)))          # :60-63
        segs_cat, cam, _ = self.model(ops.flip_concat(base), names, mode="val")             # :68-69
        C, h, w = segs_cat.shape[1:]
        n_scales = 1 + sum(1 for s in self.scales if s != 1.0)
        msc = torch.empty(B, C, h, w, dtype=torch.float32, device=self.device)
        ops.msc_accumulate(segs_cat, msc, 1.0 / n_scales, True)                              # :74
        for s in self.scales:                                                                # :79-88
            if s != 1.0:
                sc, _, _ = self.model(ops.flip_concat(ops.resize_bilinear(base, scale_factor=s)), names, mode="val")
                ops.msc_accumulate(sc, msc, 1.0 / n_scales, False)
        size = labels.shape[-2:]
        labels = labels.contiguous()
        seg_pred = ops.upsample_argmax_hist(segs_cat[:B], size, labels, self.hist[0], want_preds)  # :93-94
        msc_pred = ops.upsample_argmax_hist(msc, size, labels, self.hist[1], want_preds)           # :95-96
        if cam is not None:                                                                  # :71, :107
            ops.confusion_hist(cam[:B], labels, self.hist[2])
        self.n_images += B
        if want_preds:
            return seg_pred, msc_pred, (cam[:B] if cam is not None else None), msc
        return None

    def all_reduce(self):
        """The only collective of the path: sum the three C x C histograms over ranks (NCCL over NVLink)."""
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.hist, op=dist.ReduceOp.SUM)
        return self.hist

    def scores(self) -> dict:
        h = self.hist.cpu().numpy()
        return {"seg": scores(h[0]), "msc_seg": scores(h[1]), "cam": scores(h[2])}


def validate(model, dataset: Iterable, test_scales=None, num_classes: int = 21, resize_long: int = 512,
             rank: int = 0, world: int = 1, device="cuda"):
    """Drop-in for validate() of test_msc_flip_voc.py: dataset yields (name, image CHW fp32, label HW, cls_label).
    Returns the three score dicts after the cross-rank histogram reduction."""
    ev = MscFlipEvaluator(model, num_classes, test_scales or [1.0, 0.75], resize_long, device)
    items = list(dataset) if not hasattr(dataset, "__getitem__") else dataset
    for i in shard_indices(len(items), rank, world):
        name, image, label = items[i][:3]
        image = torch.as_tensor(image, dtype=torch.float32).unsqueeze(0).to(ev.device, non_blocking=True)
        label = torch.as_tensor(np.asarray(label)).to(torch.int16).to(ev.device, non_blocking=True).contiguous()
        ev.process(str(name), image, label)
    ev.all_reduce()
    return ev.scores()

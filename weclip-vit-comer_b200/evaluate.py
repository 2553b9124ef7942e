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


class LogitDumper:
    """Per-image ``.npy`` logit dump of test_msc_flip_voc.py:111 — ``np.save(work_dir/logit/<name>.npy, {"segs": [1,C,h,w],
    "msc_segs": [1,C,h,w]})``, the file the CRF stage reads back with ``np.load(..., allow_pickle=True).item()`` (:140-141).
    Device -> pinned-host copies are asynchronous; files are written when their copy has completed (``drain``), so the
    evaluation loop is never synchronised per image."""

    def __init__(self, logit_dir: str):
        import os
        self.dir = logit_dir
        os.makedirs(logit_dir, exist_ok=True)
        self._pending = []

    def add(self, names, segs: torch.Tensor, msc: torch.Tensor):
        """segs / msc fp32 [B,C,h,w] on the device, valid on the current stream."""
        hs = torch.empty(segs.shape, dtype=torch.float32, pin_memory=True)
        hm = torch.empty(msc.shape, dtype=torch.float32, pin_memory=True)
        hs.copy_(segs, non_blocking=True)
        hm.copy_(msc, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._pending.append((ev, list(names), hs, hm))
        self.drain(block=False)

    def drain(self, block: bool = True):
        import os
        keep = []
        for ev, names, hs, hm in self._pending:
            if not block and not ev.query():
                keep.append((ev, names, hs, hm))
                continue
            ev.synchronize()
            for i, name in enumerate(names):
                np.save(os.path.join(self.dir, str(name) + ".npy"),
                        {"segs": hs[i:i + 1].numpy().copy(), "msc_segs": hm[i:i + 1].numpy().copy()})
        self._pending = keep


def shard_indices(n_items: int, rank: int, world: int) -> range:
    """Round-robin image sharding (the reference's own multi-GPU idiom, clip/generate_cams_voc12.py:96-100)."""
    return range(rank, n_items, world)


class MscFlipEvaluator:
    def __init__(self, model, num_classes: int = 21, test_scales: Sequence[float] = (1.0, 0.75, 1.25),
                 resize_long: int = 512, device="cuda", overlap_scales: bool = True, cam_passes: str = "reference",
                 logit_dir: Optional[str] = None):
        """cam_passes: how often the CAM / affinity / PAR path runs per dataset image.
        "reference" = the reference's call pattern: ``names = name+name`` (test_msc_flip_voc.py:52, coco :53) is passed
        with EVERY ``[x_s, flip(x_s)]`` batch (:69, :84), so pseudo-labels are built for the plain AND the flipped image at
        every scale: 2 x n_scales passes per dataset image (6 for scales 1.0/0.75/1.25).
        "live" = dead-output elimination: the caller only reads ``cam[0]`` of the scale-1.0 call (:71; ``cam_cat`` of :84
        is never read), so ONE pass per dataset image produces the same three histograms."""
        if cam_passes not in ("reference", "live"):
            raise ValueError("cam_passes must be 'reference' or 'live'")
        self.cam_passes = cam_passes
        self.dumper = LogitDumper(logit_dir) if logit_dir else None
        self._names = None
        self.model = model
        self.C = num_classes
        self.scales = list(test_scales)
        self.resize_long = resize_long
        self.device = torch.device(device)
        self._hist = torch.zeros(3, num_classes, num_classes, dtype=torch.int64, device=self.device)  # seg, msc, cam
        self.n_images = 0
        # The scales of one image are independent forward passes: each runs on its own CUDA stream so that one pass's
        # tail waves and bandwidth-bound kernels (LayerNorm, PAR) fill in under another pass's tensor-core kernels.  The
        # fusion / histogram tail runs on a fourth stream and the caller's stream is NOT made to wait for it, so the
        # next image's passes start while this image's longest pass is still running; join() (called by all_reduce /
        # scores, or by the caller) orders the caller's stream after everything enqueued so far.
        self.overlap_scales = bool(overlap_scales) and self.device.type == "cuda"
        self._side_streams = []
        self._acc_stream = None

    def _streams(self, n):
        while len(self._side_streams) < n:
            self._side_streams.append(torch.cuda.Stream(self.device))
        if self._acc_stream is None:
            self._acc_stream = torch.cuda.Stream(self.device)
        return self._side_streams[:n], self._acc_stream

    @property
    def hist(self) -> torch.Tensor:
        """The three C x C confusion matrices (seg, msc seg, CAM), valid on the current stream."""
        self.join()
        return self._hist

    def join(self):
        """Order the current stream after all evaluation work enqueued so far (no host synchronisation)."""
        if self._acc_stream is not None:
            main = torch.cuda.current_stream(self.device)
            main.wait_stream(self._acc_stream)
            for st in self._side_streams:
                main.wait_stream(st)

    def _forward_scales(self, base: torch.Tensor, names, side):
        """model([x_s, flip(x_s)]) for scale 1.0 and every other test scale -> list of (segs, cam), scale 1.0 first.
        side: one stream per scale (already ordered after `base`), or None to run on the current stream."""
        scales = [1.0] + [s for s in self.scales if s != 1.0]

        def fwd(s):
            x = base if s == 1.0 else ops.resize_bilinear(base, scale_factor=s)
            if self.cam_passes == "reference":
                nm = names + names                       # names = name+name (test_msc_flip_voc.py:52): plain + flipped
            else:
                nm = names if s == 1.0 else []           # only cam[0] of the scale-1.0 call is read (:71)
            segs, cam, _ = self.model(ops.flip_concat(x), nm, mode="val")
            return segs, cam

        if side is None:
            return [fwd(s) for s in scales]
        outs = [None] * len(scales)
        for i in sorted(range(len(scales)), key=lambda i: -scales[i]):      # longest pass first
            with torch.cuda.stream(side[i]):
                outs[i] = fwd(scales[i])
        return outs

    @property
    def cam_passes_per_image(self) -> int:
        """CAM / affinity / PAR passes per dataset image under the current ``cam_passes`` mode."""
        n_sc = 1 + sum(1 for s in self.scales if s != 1.0)
        return 2 * n_sc if self.cam_passes == "reference" else 1

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
        names = list(names)      # one name per dataset image; _forward_scales repeats them for the flipped twins
        self._names = names
        _, _, h0, w0 = inputs.shape
        ratio = self.resize_long / max(h0, w0)
        n_sc = 1 + sum(1 for s in self.scales if s != 1.0)
        labels = labels.contiguous()
        base = ops.resize_bilinear(inputs, size=(int(h0 * ratio), int(w0 * ratio)))          # :60-63
        if not self.overlap_scales:
            return self._fuse(self._forward_scales(base, names, None), B, labels, want_preds)
        main = torch.cuda.current_stream(self.device)
        side, acc = self._streams(n_sc)
        ready = torch.cuda.Event()
        ready.record(main)
        for st in side + [acc]:
            st.wait_event(ready)
        outs = self._forward_scales(base, names, side)                                       # :68-69, :79-88
        for st in side:
            acc.wait_stream(st)
            base.record_stream(st)           # allocated on the caller's stream, read on the side streams
        for t in (labels, inputs):
            t.record_stream(acc)
        for segs, cam in outs:               # allocated on a side stream, read on the fusion stream
            segs.record_stream(acc)
            if cam is not None:
                cam.record_stream(acc)
        with torch.cuda.stream(acc):
            res = self._fuse(outs, B, labels, want_preds)
        if want_preds:
            main.wait_stream(acc)
            for t in res:
                if t is not None:
                    t.record_stream(main)
        return res

    def _fuse(self, outs, B, labels, want_preds):
        segs_cat, cam = outs[0]
        C, h, w = segs_cat.shape[1:]
        n_scales = len(outs)
        msc = torch.empty(B, C, h, w, dtype=torch.float32, device=self.device)
        ops.msc_accumulate(segs_cat, msc, 1.0 / n_scales, True)                              # :74
        for sc, _ in outs[1:]:
            ops.msc_accumulate(sc, msc, 1.0 / n_scales, False)
        size = labels.shape[-2:]
        seg_pred = ops.upsample_argmax_hist(segs_cat[:B], size, labels, self._hist[0], want_preds)  # :93-94
        msc_pred = ops.upsample_argmax_hist(msc, size, labels, self._hist[1], want_preds)           # :95-96
        if cam is not None:                                                                  # :71, :107
            ops.confusion_hist(cam[:B], labels, self._hist[2])
        self.n_images += B
        if self.dumper is not None:                                                          # :111
            self.dumper.add(self._names, segs_cat[:B], msc)
        if want_preds:
            return seg_pred, msc_pred, (cam[:B] if cam is not None else None), msc
        return None

    def all_reduce(self, in_place: bool = True):
        """The only collective of the path: sum the three C x C histograms over ranks (NCCL over NVLink).
        in_place=False reduces a copy (a running evaluation keeps accumulating its local histograms)."""
        import torch.distributed as dist
        hist = self.hist if in_place else self.hist.clone()
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(hist, op=dist.ReduceOp.SUM)
        return hist

    def scores(self) -> dict:
        if self.dumper is not None:
            self.dumper.drain()
        h = self.hist.cpu().numpy()
        return {"seg": scores(h[0]), "msc_seg": scores(h[1]), "cam": scores(h[2])}


def validate(model, dataset: Iterable, test_scales=None, num_classes: int = 21, resize_long: int = 512,
             rank: int = 0, world: int = 1, device="cuda", cam_passes: str = "reference",
             logit_dir: Optional[str] = None):
    """Drop-in for validate() of test_msc_flip_voc.py: dataset yields (name, image CHW fp32, label HW, cls_label).
    Returns the three score dicts after the cross-rank histogram reduction."""
    ev = MscFlipEvaluator(model, num_classes, test_scales or [1.0, 0.75], resize_long, device, cam_passes=cam_passes,
                          logit_dir=logit_dir)
    items = list(dataset) if not hasattr(dataset, "__getitem__") else dataset
    for i in shard_indices(len(items), rank, world):
        name, image, label = items[i][:3]
        image = torch.as_tensor(image, dtype=torch.float32).unsqueeze(0).to(ev.device, non_blocking=True)
        label = torch.as_tensor(np.asarray(label)).to(torch.int16).to(ev.device, non_blocking=True).contiguous()
        ev.process(str(name), image, label)
    ev.all_reduce()
    return ev.scores()

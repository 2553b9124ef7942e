"""msc+flip evaluation driver — the caller that defines "one dataset image" for the metric.

Mirrors ``validate(model, dataset, test_scales)`` of test_msc_flip_voc.py:32-113 (COCO twin
test_msc_flip_coco.py:33-116) and ``utils/evaluate.py:9-36``: long-side-512 resize, [x, flip(x)] batch per
scale, flip-average, scale-mean, resize to the label size, argmax, three confusion matrices (single-scale
seg, msc seg, CAM pseudo-label).  All arithmetic runs in libweclip_sm100.so (csrc/evalops.cu); the three
C x C int64 histograms stay on the device and are summed over ranks with ONE all-reduce at the end
(images are sharded round-robin over ranks, no data-path collective).
"""
from __future__ import annotations

import math
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


class PinnedRing:
    """Input staging for the evaluation loop: a few pinned host slots and a copy stream, so that the
    host->device copy of image i+1 runs under the kernels of image i.  put() fills a slot from host memory and issues its
    copies; get() makes the current stream wait for them; done() marks the slot reusable once the work enqueued so far on
    the current stream has finished."""

    def __init__(self, device="cuda", slots: int = 3):
        self.device = torch.device(device)
        self.copy_stream = torch.cuda.Stream(self.device)
        self.slots = [dict(img=None, lab=None, ready=None, free=None) for _ in range(slots)]
        self._next = 0
        self.h2d_bytes = 0

    @staticmethod
    def _fit(buf, shape, dtype):
        n = int(np.prod(shape))
        if buf is None or buf.numel() < n or buf.dtype != dtype:
            buf = torch.empty(n, dtype=dtype, pin_memory=True)
        return buf

    def put(self, image, label):
        """image [3,H,W] float32, label [H,W] integer (numpy arrays or CPU tensors) -> slot handle."""
        i = self._next
        self._next = (i + 1) % len(self.slots)
        s = self.slots[i]
        if s["free"] is not None:
            s["free"].synchronize()                      # the step that read this slot last has completed
        img = torch.as_tensor(np.asarray(image), dtype=torch.float32)
        lab = torch.as_tensor(np.asarray(label)).to(torch.int16)
        s["img"], s["lab"] = self._fit(s["img"], img.shape, torch.float32), self._fit(s["lab"], lab.shape, torch.int16)
        pi, pl = s["img"][:img.numel()].view(img.shape), s["lab"][:lab.numel()].view(lab.shape)
        pi.copy_(img); pl.copy_(lab)
        with torch.cuda.stream(self.copy_stream):
            s["d_img"] = pi.to(self.device, non_blocking=True)
            s["d_lab"] = pl.to(self.device, non_blocking=True)
            s["ready"] = torch.cuda.Event()
            s["ready"].record(self.copy_stream)
        self.h2d_bytes += img.numel() * 4 + lab.numel() * 2
        return i

    def get(self, i):
        s = self.slots[i]
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(s["ready"])
        s["d_img"].record_stream(cur); s["d_lab"].record_stream(cur)
        return s["d_img"], s["d_lab"]

    def done(self, i):
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.slots[i]["free"] = ev


def shard_indices(n_items: int, rank: int, world: int) -> range:
    """Round-robin image sharding (the reference's own multi-GPU idiom, clip/generate_cams_voc12.py:96-100)."""
    return range(rank, n_items, world)


class MscFlipEvaluator:
    def __init__(self, model, num_classes: int = 21, test_scales: Sequence[float] = (1.0, 0.75, 1.25),
                 resize_long: int = 512, device="cuda", overlap_scales: bool = True, cam_passes: str = "reference",
                 logit_dir: Optional[str] = None, use_graphs: bool = False, graph_slots: int = 2):
        """cam_passes: how often the CAM / affinity / PAR path runs per dataset image.
        "reference" = the reference's call pattern: ``names = name+name`` (test_msc_flip_voc.py:52, coco :53) is passed
        with EVERY ``[x_s, flip(x_s)]`` batch (:69, :84), so pseudo-labels are built for the plain AND the flipped image at
        every scale: 2 x n_scales passes per dataset image (6 for scales 1.0/0.75/1.25).
        "live" = dead-output elimination: the caller only reads ``cam[0]`` of the scale-1.0 call (:71; ``cam_cat`` of :84
        is never read), so ONE pass per dataset image produces the same three histograms."""
        if cam_passes not in ("reference", "live"):
            raise ValueError("cam_passes must be 'reference' or 'live'")
        # use_graphs: one CUDA graph per (image size, label size, batch, classes-per-image signature) bucket holds the
        # whole step — long-side resize, the three [x, flip(x)] forwards with their CAM / PAR passes on forked streams,
        # fusion, argmax, histograms (~500 launches) — and every later image of the bucket is ONE graph launch
        # (SURVEY §8f-1).  graph_slots replay slots with their own streams, workspaces and static buffers take turns so
        # that image i+1 starts under the tail of image i (one image per call leaves most of the 148 SMs idle in the narrow
        # kernels; more slots = more images in flight).
        self.use_graphs = bool(use_graphs) and torch.device(device).type == "cuda"
        self.graph_slots = max(1, int(graph_slots))
        self._graphs = {}
        self._slots = []
        self._slot = 0
        self._pool = None
        self.graph_captures = 0
        self.graph_replays = 0
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
        main = torch.cuda.current_stream(self.device) if self.device.type == "cuda" else None
        if self._acc_stream is not None:
            main.wait_stream(self._acc_stream)
            for st in self._side_streams:
                main.wait_stream(st)
        for slot in self._slots:
            main.wait_stream(slot["launch"])

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

    @torch.no_grad()      # evaluation never differentiates: keeps WeCLIP on its fused inference kernels (no autograd tape)
    def process_batch(self, names: Sequence[str], inputs: torch.Tensor, labels: torch.Tensor, want_preds: bool = False):
        """B same-size dataset images per model call (the reference loops image by image; images are independent,
        so a (H,W) bucket is run as one [2B,3,H,W] forward per scale: B plain images followed by their B flipped
        twins).  inputs fp32 [B,3,H0,W0]; labels int16 [B,Hl,Wl]."""
        B = inputs.shape[0]
        names = list(names)      # one name per dataset image; _forward_scales repeats them for the flipped twins
        self._names = names
        if self.use_graphs and not want_preds and self.dumper is None:
            return self._process_graphed(names, inputs, labels.contiguous())
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

    # ---- CUDA-graph path -----------------------------------------------------------------------------------------
    def _scale_list(self):
        return [1.0] + [s for s in self.scales if s != 1.0]

    def _names_for(self, s, names):
        if self.cam_passes == "reference":
            return names + names                         # names = name+name (test_msc_flip_voc.py:52)
        return names if s == 1.0 else []                 # only cam[0] of the scale-1.0 call is read (:71)

    def _host_metas(self, names, B, h0, w0):
        """WeCLIP.prepare() for every scale's [x_s, flip(x_s)] call (host lists only; sizes as F.interpolate computes them)."""
        ratio = self.resize_long / max(h0, w0)
        hb, wb = int(h0 * ratio), int(w0 * ratio)
        metas = []
        for s in self._scale_list():
            hs, ws = (hb, wb) if s == 1.0 else (int(math.floor(hb * s)), int(math.floor(wb * s)))
            metas.append(self.model.prepare(self._names_for(s, names), "val", 2 * B, hs, ws, device_tensors=False))
        return metas

    def _slot_state(self):
        if not self._slots:
            for _ in range(self.graph_slots):
                # graphs of one slot replay in order on its launch stream and may share one memory pool; the two slots
                # run concurrently, so each has its own pool, streams and (through the per-stream arenas) workspaces
                self._slots.append(dict(launch=torch.cuda.Stream(self.device), acc=torch.cuda.Stream(self.device),
                                        side=[torch.cuda.Stream(self.device) for _ in self._scale_list()],
                                        pool=torch.cuda.graph_pool_handle()))
        slot = self._slots[self._slot]
        self._slot = (self._slot + 1) % len(self._slots)
        return slot

    @torch.no_grad()
    def _graph_step(self, slot, ent):
        """The step as it is captured: every launch goes to the slot's streams, forked from / joined to its launch stream."""
        launch, side, acc = slot["launch"], slot["side"], slot["acc"]
        inputs, labels, metas = ent["inputs"], ent["labels"], ent["metas"]
        B = inputs.shape[0]
        _, _, h0, w0 = inputs.shape
        ratio = self.resize_long / max(h0, w0)
        base = ops.resize_bilinear(inputs, size=(int(h0 * ratio), int(w0 * ratio)))          # :60-63
        scales = self._scale_list()
        for st in side + [acc]:
            st.wait_stream(launch)
        outs = [None] * len(scales)
        for i in sorted(range(len(scales)), key=lambda i: -scales[i]):                        # longest pass first
            with torch.cuda.stream(side[i]):
                x = base if scales[i] == 1.0 else ops.resize_bilinear(base, scale_factor=scales[i])
                segs, cam, _ = self.model.run(ops.flip_concat(x), metas[i])
                outs[i] = (segs, cam)
        for st in side:
            acc.wait_stream(st)
        with torch.cuda.stream(acc):
            self._fuse(outs, B, labels, False, count=False)
        launch.wait_stream(acc)

    def _capture(self, key, slot, inputs, labels, metas_h):
        from .model import ForwardMeta
        launch = slot["launch"]
        ent = dict(inputs=torch.empty_like(inputs), labels=torch.empty_like(labels), metas=[])
        for m_h in metas_h:
            m = ForwardMeta()
            m.__dict__.update(m_h.__dict__)
            if m_h.cam_path:
                m.rows = torch.empty(m_h.rows_host.shape, dtype=torch.int64, device=self.device)
                m.lut = torch.empty(m_h.lut_host.shape, dtype=torch.int64, device=self.device)
            ent["metas"].append(m)
        self._refresh(ent, inputs, labels, metas_h)
        # warm-up on the slot's own streams: sizes every workspace the capture will use (per-stream arenas), sets the
        # kernels' shared-memory attributes; its histogram counts are taken back
        torch.cuda.synchronize(self.device)       # nothing else may be adding to the histograms while they are saved / restored
        saved = self._hist.clone()
        self._graph_step(slot, ent)
        launch.synchronize()
        self._hist.copy_(saved)
        g = torch.cuda.CUDAGraph()
        # thread_local: other threads of the process (the NCCL watchdog polls events) must not invalidate the capture
        with torch.cuda.graph(g, pool=slot["pool"], stream=launch, capture_error_mode="thread_local"):
            self._graph_step(slot, ent)
        ent["graph"] = g
        self._graphs[key] = ent
        self.graph_captures += 1
        return ent

    def _refresh(self, ent, inputs, labels, metas_h):
        ent["inputs"].copy_(inputs, non_blocking=True)
        ent["labels"].copy_(labels, non_blocking=True)
        for m, m_h in zip(ent["metas"], metas_h):
            if m_h.cam_path:
                m.rows.copy_(m_h.rows_host.pin_memory(), non_blocking=True)
                m.lut.copy_(m_h.lut_host.pin_memory(), non_blocking=True)

    def _process_graphed(self, names, inputs, labels):
        B = inputs.shape[0]
        _, _, h0, w0 = inputs.shape
        metas_h = self._host_metas(names, B, h0, w0)
        slot = self._slot_state()
        launch = slot["launch"]
        caller = torch.cuda.current_stream(self.device)
        launch.wait_stream(caller)                        # the caller's stream produced inputs / labels
        eng = self.model.engine
        if any(m.cam_path for m in metas_h):
            self.model._text_table(self.device)           # (re)build the cached text table before it becomes part of the key
        key = (id(slot), B, h0, w0, tuple(labels.shape[-2:]), self.cam_passes, tuple(m.signature() for m in metas_h),
               id(eng), eng.generation, getattr(self.model, "_text_key", None), str(self.model.par.affinity_dtype))
        with torch.cuda.stream(launch):
            ent = self._graphs.get(key)
            if ent is None:
                ent = self._capture(key, slot, inputs, labels, metas_h)
            else:
                self._refresh(ent, inputs, labels, metas_h)
            ent["graph"].replay()
        inputs.record_stream(launch)
        labels.record_stream(launch)
        self.graph_replays += 1
        self.n_images += B
        self.model.iter_num += len(metas_h)
        return None

    def _fuse(self, outs, B, labels, want_preds, count=True):
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
        if count:
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
             logit_dir: Optional[str] = None, use_graphs: bool = True):
    """Drop-in for validate() of test_msc_flip_voc.py: dataset yields (name, image CHW fp32, label HW, cls_label).
    One dataset image per model call like the reference's DataLoader(batch_size=1); images are sharded round-robin over
    ranks, staged through a pinned ring (copy of image i+1 under the kernels of image i) and — with use_graphs — replayed
    from one CUDA graph per (size, classes-per-image) bucket.  Returns the three score dicts after the cross-rank
    histogram reduction."""
    ev = MscFlipEvaluator(model, num_classes, test_scales or [1.0, 0.75], resize_long, device, cam_passes=cam_passes,
                          logit_dir=logit_dir, use_graphs=use_graphs and logit_dir is None)
    items = list(dataset) if not hasattr(dataset, "__getitem__") else dataset
    idx = list(shard_indices(len(items), rank, world))
    ring = PinnedRing(ev.device)
    pending = ring.put(items[idx[0]][1], items[idx[0]][2]) if idx else None
    for n, i in enumerate(idx):
        slot = pending
        if n + 1 < len(idx):
            nxt = items[idx[n + 1]]
            pending = ring.put(nxt[1], nxt[2])           # staged while image i computes
        image, label = ring.get(slot)
        ev.process(str(items[i][0]), image.unsqueeze(0), label.contiguous())
        ring.done(slot)
    ev.all_reduce()
    return ev.scores()

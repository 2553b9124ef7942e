"""Full-path bench workloads (registered into bench.py's WORKLOADS table).

``mscflip`` = BASELINE.json configs[2]: test_msc_flip_voc inference, scales 1.0/0.75/1.25 + flip at 512x512,
21 classes, images sharded across ranks.  One step = IMAGES_PER_STEP dataset images through
``evaluate.MscFlipEvaluator.process_batch`` (6 WeCLIP forwards each: encoder, decoder and — with the reference's
``names = name+name`` call pattern, test_msc_flip_voc.py:52 — Grad-CAM, affinity refinement, CAM upsample and PAR at label
resolution for all six images; then fusion + argmax + 3 confusion matrices).  ``value`` / ``e2e`` are measured with that
call pattern (cam_passes="reference": 6 CAM/PAR passes per dataset image); ``value_live_outputs`` is the same step with
the dead passes eliminated (cam_passes="live": 1 pass, identical histograms).
"""
from __future__ import annotations

import ctypes
import os
import tempfile
import time

import numpy as np
import torch

from . import _lib, synth

IMAGES_PER_STEP = 8
KINDS_TENSOR = ("gemm_tc", "attention")
# profiler class -> kernel-name fragments in the ncu traffic summary (tools/ncu_traffic.py -> profiles/*ncu_traffic*.json)
NCU_NAMES = {"gemm_tc": ("gemm_tc",), "attention": ("attn_fwd_kernel", "attn_bwd_colsum_kernel"),
             "softmax_headmean": ("attn_headmean_kernel", "softmax_headmean_kernel"), "par_iter": ("par_iter",),
             "par_affinity": ("par_aff", "par_resize"), "layernorm": ("layernorm_kernel",)}


def ncu_traffic():
    """DRAM bytes (read + write) per launch and per kernel class, from the newest committed ncu traffic summary of ONE
    msc+flip step (same workload, 8 dataset images); None when no summary is in the tree."""
    import glob
    import json
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    import re

    def version(path):      # r01_ncu_traffic_v13.json -> (1, 13): numeric, so that v13 sorts after v8
        return tuple(int(x) for x in re.findall(r"\d+", os.path.basename(path)))
    files = sorted(glob.glob(os.path.join(repo, "profiles", "*ncu_traffic*.json")), key=version)
    if not files:
        return None
    kern = json.load(open(files[-1]))["kernels"]
    out = {"_file": os.path.relpath(files[-1], repo)}
    for cls, frags in NCU_NAMES.items():
        sel = [v for k, v in kern.items() if any(f in k for f in frags)]
        n = sum(v["launches"] for v in sel)
        if n:
            out[cls] = int(sum(v["dram_bytes_total"] for v in sel) / n)
    return out


def prof_summary():
    L = _lib.lib()
    n = L.wc_prof_kinds()
    ms, work, cnt = (ctypes.c_double * n)(), (ctypes.c_double * n)(), (ctypes.c_long * n)()
    L.wc_prof_kind_name.restype = ctypes.c_char_p
    _lib.check(L.wc_prof_summary(ms, work, cnt), "wc_prof_summary")
    return {L.wc_prof_kind_name(i).decode(): (ms[i], work[i], cnt[i]) for i in range(n)}


class MscFlipVoc:
    name = ("cfg3: test_msc_flip_voc val, scales 1.0/0.75/1.25 + flip, 512x512 images and labels, 21 classes, "
            "K per image from the VOC histogram")
    dtype = "f16"
    n_classes, n_fg, n_bg, dataset, subdir = 21, 20, 25, "voc", "SegmentationClassAug"
    scales = (1.0, 0.75, 1.25)
    H = W = 512
    units = IMAGES_PER_STEP
    # per forward pair at 512x512 the attention scores alone are 2*12*1025*1032*4 B = 101 MB per block; with taps, MLP
    # activations and PAR affinities the working set per image is > 126 MB
    l2_note = "working set larger than L2 (>= 101 MB of attention scores per block at 512x512, x3 scales)"

    def _make_data(self, seed, n_img, root):
        rng = np.random.default_rng(seed)
        names, labs = [], []
        for i in range(n_img):
            ids = synth.draw_label_set(rng, self.n_fg)
            lab = synth.make_label_map(rng, self.H, self.W, ids)
            synth.write_label_png(root, self.subdir, f"b{seed}_{i}", lab)
            names.append(f"b{seed}_{i}")
            labs.append(lab)
        imgs = synth.make_images(1000 + seed, n_img, self.H, self.W)
        return names, labs, imgs

    def __init__(self, rank: int, device):
        from .evaluate import MscFlipEvaluator
        from .model import WeCLIP
        self.dev = device
        self.root = tempfile.mkdtemp(prefix="wcbench")
        self.sd = synth.make_state_dict(0, self.n_classes)
        self.fg, self.bg = synth.make_text_features(7, self.n_fg, self.n_bg)
        self.names, labs, imgs = self._make_data(100 + rank, IMAGES_PER_STEP, self.root)
        self.model = WeCLIP(num_classes=self.n_classes, dataset_root_path=self.root, device=device,
                            dataset=self.dataset, state_dict=self.sd, fg_text_features=self.fg,
                            bg_text_features=self.bg)
        overlap = os.environ.get("WC_OVERLAP_SCALES", "1") != "0"
        graphs = os.environ.get("WC_BENCH_GRAPHS", "1") != "0"
        self.graphs = graphs
        self.ev = MscFlipEvaluator(self.model, self.n_classes, self.scales, 512, device, overlap_scales=overlap,
                                   cam_passes="reference", use_graphs=graphs)
        self.ev_live = MscFlipEvaluator(self.model, self.n_classes, self.scales, 512, device, overlap_scales=overlap,
                                        cam_passes="live", use_graphs=graphs)
        # one dataset image per call (the reference's DataLoader(batch_size=1)): the same 512x512 images, and a VOC-like
        # size mix (long side 500 -> short sides 384 / 341 / 340 / 287 ... after the resize, mostly not multiples of 16)
        self.ev_b1 = MscFlipEvaluator(self.model, self.n_classes, self.scales, 512, device, overlap_scales=overlap,
                                      cam_passes="reference", use_graphs=graphs,
                                      graph_slots=int(os.environ.get("WC_GRAPH_SLOTS", "4")))
        voc = synth.make_voc_like_dataset(500 + rank, IMAGES_PER_STEP, self.root, self.subdir, self.n_fg)
        self.voc_items = [(n, im[None].to(device), torch.from_numpy(lab.astype(np.int16)).to(device)) for n, im, lab in voc]
        self.voc_sizes = [tuple(im.shape[-2:]) for _, im, _ in self.voc_items]
        self.cam_passes_per_image = self.ev.cam_passes_per_image if self.dataset == "voc" else 0
        self.imgs_h = imgs.pin_memory()
        self.labs_h = torch.from_numpy(np.stack(labs).astype(np.int16)).pin_memory()
        self.imgs_d = self.imgs_h.to(device)
        self.labs_d = self.labs_h.to(device)
        self.hist_h = torch.empty(3, self.n_classes, self.n_classes, dtype=torch.int64).pin_memory()
        self.h2d = self.imgs_h.numel() * 4 + self.labs_h.numel() * 2
        self.d2h = self.hist_h.numel() * 8
        self.launches = None

    def step_device(self, record=False):
        # the step's images share one (H,W) bucket: one [2B,3,H,W] forward per scale
        self.ev.process_batch(self.names, self.imgs_d, self.labs_d)

    def step_device_live(self):
        """Same step with cam_passes="live": only the CAM / PAR pass whose labels the driver reads."""
        self.ev_live.process_batch(self.names, self.imgs_d, self.labs_d)

    def step_b1(self):
        """The step's 8 images ONE PER CALL, as the reference's batch_size-1 loop drives the model."""
        for i in range(IMAGES_PER_STEP):
            self.ev_b1.process(self.names[i], self.imgs_d[i:i + 1], self.labs_d[i])

    def step_b1_voc_sizes(self):
        for name, img, lab in self.voc_items:
            self.ev_b1.process(name, img, lab)

    def join(self):
        """Order the current stream after every step enqueued so far (the passes run on the evaluator's own streams)."""
        for ev in (getattr(self, "ev", None), getattr(self, "ev_live", None), getattr(self, "ev_b1", None)):
            if ev is not None:              # workloads without an evaluator run on the current stream already
                ev.join()

    def step_e2e(self):
        """Host buffers in, histograms out: H2D of the step's images + labels from pinned memory, the step, the cross-rank
        histogram sum (the path's only collective; a no-op at world size 1) and the D2H read of the result."""
        img = self.imgs_h.to(self.dev, non_blocking=True)
        lab = self.labs_h.to(self.dev, non_blocking=True)
        self.ev.process_batch(self.names, img, lab)
        self.hist_h.copy_(self.ev.all_reduce(in_place=False), non_blocking=True)   # .hist orders this stream after the passes
        torch.cuda.current_stream().synchronize()
        return self.hist_h

    def profile(self, steps: int):
        """Separate pass with the C-side profiler on: per-kernel-class device time and algorithmic work."""
        L = _lib.lib()
        L.wc_prof_reset(); L.wc_prof_enable(1)
        ev = getattr(self, "ev", None)
        overlap = ev.overlap_scales if ev is not None else False
        graphs = ev.use_graphs if ev is not None else False
        if ev is not None:
            ev.overlap_scales = False      # per-kernel times are taken with each kernel alone on the device,
            ev.use_graphs = False          # launched eagerly (the profiler brackets launches, not graph nodes)
        for _ in range(steps):
            self.step_device()
        if ev is not None:
            ev.join()
            ev.overlap_scales, ev.use_graphs = overlap, graphs
        summ = prof_summary()
        L.wc_prof_enable(0); L.wc_prof_reset()
        self._prof = {k: (ms / steps, work / steps, cnt // steps) for k, (ms, work, cnt) in summ.items()}
        self.launches = sum(c for _, _, c in self._prof.values())
        return self._prof

    def roofline(self, peaks):
        prof = self._prof
        total = sum(ms for ms, _, _ in prof.values()) or 1.0
        top = max(prof, key=lambda k: prof[k][0])
        rows = {}
        for k, (ms, work, cnt) in prof.items():
            if cnt == 0:
                continue
            row = {"ms_per_step": round(ms, 3), "share_of_profile_pass": round(ms / total, 4), "launches_per_step": cnt}
            if work > 0 and ms > 0:
                if k in KINDS_TENSOR:
                    a = work / (ms * 1e-3) / 1e12
                    row.update(achieved=round(a, 1), unit="TFLOP/s", frac=round(a / peaks["bf16_tflops_sustained"], 4))
                else:
                    a = work / (ms * 1e-3) / 1e9
                    row.update(achieved=round(a, 1), unit="GB/s", frac=round(a / peaks["hbm_gbs"], 4))
            rows[k] = row
        traffic = ncu_traffic() or {}
        for k, row in rows.items():
            if k in traffic:
                row["dram_bytes_per_launch_ncu"] = traffic[k]
        ms, work, cnt = prof[top]
        tensor = top in KINDS_TENSOR
        peak = peaks["bf16_tflops_sustained"] if tensor else peaks["hbm_gbs"]
        ach = (work / (ms * 1e-3) / (1e12 if tensor else 1e9)) if work > 0 else None
        return {"bound": "tensor" if tensor else "hbm", "kernel": top, "achieved": round(ach, 1) if ach else None,
                "peak": peak, "peak_source": peaks["source"] + (" (sustained cuBLAS bf16)" if tensor else " (copy)"),
                "unit": "TFLOP/s" if tensor else "GB/s", "frac": round(ach / peak, 4) if ach else None,
                "traffic": traffic.get(top), "traffic_source": traffic.get("_file"),
                "kernel_ms_per_launch": round(ms / max(cnt, 1), 4),
                "kernels_note": "per-class device time from a separate SERIALISED profile pass (one stream, events around every "
                                "launch); the timed step overlaps its three scale passes, so the classes sum to more than "
                                "ms_per_step",
                "kernels": rows}

    # ---- CPU baseline / reference arm (oracle port of the evaluation loop; bounded sample) ---------------
    @classmethod
    def cpu(cls):
        wl = cls.__new__(cls)
        wl.root = tempfile.mkdtemp(prefix="wcbench")
        wl.sd = synth.make_state_dict(0, cls.n_classes)
        wl.fg, wl.bg = synth.make_text_features(7, cls.n_fg, cls.n_bg)
        wl.names, wl.labs, wl.imgs_h = wl._make_data(100, 1, wl.root)
        return wl

    def cpu_sample(self, n_img=1):
        from oracle import weclip_oracle as O
        if not hasattr(self, "labs"):
            self.labs = [self.labs_h[i].numpy().astype(np.uint8) for i in range(n_img)]
        t0 = time.perf_counter()
        for i in range(n_img):
            O.msc_flip_image(self.imgs_h[i:i + 1].clone(), self.labs[i], self.sd, self.fg, self.bg, self.scales, 512,
                             self.dataset)
        dt = time.perf_counter() - t0
        return n_img / dt, (f"{n_img} of the step's {IMAGES_PER_STEP} dataset images (6 forwards with names = name+name, i.e. "
                            f"{self.cam_passes_per_image} CAM/PAR passes, + fusion each), oracle port of the reference loop "
                            "(pinned against the real validate(), tests/golden/mscflip_voc.npz), torch CPU fp32")

    def gpu_library_sample(self, n_img=3):
        """The "library-call bar" (SURVEY §8d): the same loop as ATen / cuBLAS / cuDNN library calls on this GPU, in the
        reference's own GPU mixed precision (fp32 in-proj + softmax, fp16 conv1 / out-proj / MLP: clip/model.py:461-476,
        clip/myAtt.py:199,321), with the reference's host round trips (OpenCV boxes, cv2.resize, numpy CAM weights).
        It is the oracle run on cuda: a baseline, never a product path."""
        from oracle import weclip_oracle as O
        dev = self.dev
        sd = {k: v.to(dev) for k, v in self.sd.items()}
        fg, bg = self.fg.to(dev), self.bg.to(dev)
        labs = [self.labs_h[i].numpy().astype(np.uint8) for i in range(min(n_img, IMAGES_PER_STEP))]

        def run(i):
            return O.msc_flip_image(self.imgs_d[i:i + 1], labs[i], sd, fg, bg, self.scales, 512, self.dataset,
                                    mimic_fp16="gpu")
        run(0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(len(labs)):
            run(i)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        return len(labs) / dt, (f"{len(labs)} dataset images after 1 warm-up, oracle loop on cuda: torch {torch.__version__} "
                                "ATen/cuBLAS library calls, reference GPU mixed precision, host OpenCV/numpy round trips")


MscFlipVoc.cam_passes_per_image = 6


class MscFlipCoco(MscFlipVoc):
    cam_passes_per_image = 0
    name = ("cfg4: test_msc_flip_coco val, scales 1.0/0.75/1.25 + flip, 512x512, 81 classes "
            "(val mode returns before CAM/PAR, model_attn_aff_coco.py:131-132)")
    n_classes, n_fg, n_bg, dataset = 81, 80, 23, "coco"
    subdir = os.path.join("SegmentationClass", "train")


class PseudoLabelVoc(MscFlipVoc):
    """BASELINE.json configs[1]: the training-time CAM + PAR pseudo-label step (scripts/dist_clip_voc.py:238-267 calls
    WeCLIP.forward(img, img_name, mode='train')): batch 32 crops of 320x320, 21 classes, one name per crop, so every image
    goes through encoder, decoder, Grad-CAM, affinity refinement (iter_num > 15000: RFM branch), upsample and PAR at crop
    resolution.  Unit = one crop; the step returns (seg, pseudo labels, attn_pred) on the device, e2e copies the labels out."""
    name = "cfg2: WeCLIP VOC CAM+PAR pseudo-label step, batch 32 at 320x320 crop, 21 classes, mode='train' (RFM branch)"
    H = W = 320
    BATCH = 32
    units = BATCH
    cam_passes_per_image = 1          # one name per crop: one CAM / PAR pass per crop
    l2_note = "working set larger than L2 (32 crops: 157 MB of MLP activations per block, 315 MB of PAR affinities)"

    def __init__(self, rank: int, device):
        from .model import WeCLIP
        self.dev = device
        self.root = tempfile.mkdtemp(prefix="wcbench")
        self.sd = synth.make_state_dict(0, self.n_classes)
        self.fg, self.bg = synth.make_text_features(7, self.n_fg, self.n_bg)
        self.names, self.labs, imgs = self._make_data(300 + rank, self.BATCH, self.root)
        self.model = WeCLIP(num_classes=self.n_classes, dataset_root_path=self.root, device=device,
                            dataset=self.dataset, state_dict=self.sd, fg_text_features=self.fg,
                            bg_text_features=self.bg)
        self.imgs_h = imgs.pin_memory()
        self.imgs_d = self.imgs_h.to(device)
        self.labels_h = torch.empty(self.BATCH, self.H, self.W, dtype=torch.int64).pin_memory()
        self.h2d = self.imgs_h.numel() * 4
        self.d2h = self.labels_h.numel() * 8
        self.launches = None

    @torch.no_grad()       # the pseudo-label step itself (forward); the decoder's training backward is not part of cfg-2
    def step_device(self, record=False):
        self.model.iter_num = 20000            # RFM branch of clip_tool.py:152-175
        return self.model(self.imgs_d, self.names, mode="train")

    @torch.no_grad()
    def step_e2e(self):
        img = self.imgs_h.to(self.dev, non_blocking=True)
        self.model.iter_num = 20000
        _, labels, _ = self.model(img, self.names, mode="train")
        self.labels_h.copy_(labels, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self.labels_h

    @classmethod
    def cpu(cls):
        wl = cls.__new__(cls)
        wl.root = tempfile.mkdtemp(prefix="wcbench")
        wl.sd = synth.make_state_dict(0, cls.n_classes)
        wl.fg, wl.bg = synth.make_text_features(7, cls.n_fg, cls.n_bg)
        wl.names, wl.labs, wl.imgs_h = wl._make_data(300, 2, wl.root)
        return wl

    def cpu_sample(self, n_img=2):
        from oracle import weclip_oracle as O
        n_img = min(n_img, 2)
        t0 = time.perf_counter()
        O.weclip_forward(self.imgs_h[:n_img].clone(), [np.asarray(l) for l in self.labs[:n_img]], self.sd, self.fg, self.bg,
                         mode="train", iter_num=20001, dataset=self.dataset)
        dt = time.perf_counter() - t0
        return n_img / dt, (f"{n_img} of the step's {self.BATCH} crops (full forward incl. Grad-CAM, affinity refinement, PAR), "
                            "oracle port of WeCLIP.forward(mode='train'), torch CPU fp32")


WORKLOADS = {"mscflip": MscFlipVoc, "mscflip_coco": MscFlipCoco, "cfg2": PseudoLabelVoc}

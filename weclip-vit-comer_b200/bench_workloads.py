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
                                      graph_slots=int(os.environ.get("WC_GRAPH_SLOTS", "2")))
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

    def step_device_fp32_par(self):
        """Same step as step_device with the PAR affinities stored in fp32 as the reference computes them (PAR.py is fp32 end to
        end; the pipeline default stores them in fp16: 2.3e-3 vs 2e-5 max error, labels >= 99.5 % either way)."""
        par = self.model.par
        keep, par.affinity_dtype = par.affinity_dtype, torch.float32
        try:
            self.ev.process_batch(self.names, self.imgs_d, self.labs_d)
        finally:
            par.affinity_dtype = keep

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
    metric = "CAM+PAR pseudo-label step crops/sec"
    unit = "crops/s"
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


class TrainStepVoc(PseudoLabelVoc):
    """ONE optimisation step of the reference's training loop (scripts/dist_clip_voc.py:238-267) at its own
    configuration (configs/voc_attn_reg.yaml: samples_per_gpu 4, crop 320, AdamW lr 2e-4 / wd 0.01): WeCLIP.forward in train mode
    under autograd (frozen encoder + CAM / refinement / PAR pseudo-labels on the inference kernels, trainable decoder through
    train.DecoderFunction), F.interpolate of the logits, affinity + segmentation loss on the device (losses.py), backward through
    the decoder (tcgen05 dgrad / wgrad GEMMs + train.cu), AdamW step (torch's fused element-wise optimiser: plumbing).
    Unit = one crop."""
    name = ("train: scripts/dist_clip_voc.py optimisation step, batch 4 at 320x320 crop (samples_per_gpu 4), 21 classes: forward "
            "mode='train' (RFM branch) + pseudo-labels + aff / seg loss + decoder backward + AdamW")
    metric = "WeCLIP decoder training step crops/sec"
    unit = "crops/s"
    BATCH = 4
    units = BATCH
    l2_note = ("4 crops: 26 MB of fp16 MLP activations per encoder block, 39 MB of PAR affinities; smaller than L2, as the "
               "reference's own step is; L2 is not flushed between steps")

    def __init__(self, rank: int, device):
        super().__init__(rank, device)
        from .train import TrainStep
        self.graphs = os.environ.get("WC_BENCH_GRAPHS", "1") != "0"
        # optimiser state on the device (capturable) with a tensor learning rate: the whole step can live in a CUDA graph
        self.opt = torch.optim.AdamW(self.model.get_param_groups()[3], lr=torch.tensor(2e-4, device=device), weight_decay=0.01,
                                     betas=(0.9, 0.999), capturable=True)
        self.step = TrainStep(self.model, self.opt, radius=8, aff_weight=0.1, use_graphs=self.graphs)
        self.loss_h = torch.zeros(1).pin_memory()
        self.d2h = 4

    def _train_step(self, img):
        self.model.iter_num = 20000                                            # RFM branch of clip_tool.py:152-175
        return self.step(img, self.names)

    def step_device(self, record=False):
        return self._train_step(self.imgs_d)

    def profile(self, steps: int):
        on, self.step.use_graphs = self.step.use_graphs, False                 # the profiler brackets launches, not graph nodes
        try:
            return super().profile(steps)
        finally:
            self.step.use_graphs = on

    def step_e2e(self):
        loss = self._train_step(self.imgs_h.to(self.dev, non_blocking=True))
        self.loss_h.copy_(loss.reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self.loss_h

    def cpu_sample(self, n_img=2):
        import torch.nn.functional as F
        from oracle import weclip_oracle as O
        n_img = min(n_img, 2)
        img = self.imgs_h[:n_img].clone()
        labs = [np.asarray(l) for l in self.labs[:n_img]]
        keys = [k for k in self.sd if k.startswith("decoder.") or k.startswith("decoder_fts_fuse.")]
        sd = {k: v.clone() for k, v in self.sd.items()}
        leaves = {k: sd[k].float().clone().requires_grad_(True) for k in keys}
        sd.update(leaves)
        opt = torch.optim.AdamW(list(leaves.values()), lr=2e-4, weight_decay=0.01)
        h = w = self.H // 16
        t0 = time.perf_counter()
        with torch.no_grad():        # pseudo-labels (encoder, Grad-CAM re-enables grad itself, refinement, PAR)
            _, cam, _ = O.weclip_forward(img, labs, self.sd, self.fg, self.bg, mode="train", iter_num=20001, dataset=self.dataset)
            taps, _ = O.encode_image(img, self.sd)
        cam = torch.as_tensor(np.asarray(cam)).long()
        fts = O.decoder_fts_fuse([t.detach() for t in taps], sd, h, w)
        seg = O.decoder(fts, sd)
        attn_pred = O.attn_pred_from_fts(fts)
        segs = F.interpolate(seg, size=cam.shape[1:], mode="bilinear", align_corners=False)
        aff_label = O.affinity_label(cam, O.mask_by_radius(h, w, 8), 255)
        pos, neg = aff_label == 1, aff_label == 0
        attn_loss = 0.5 * (1 - attn_pred)[pos].sum() / (pos.sum() + 1) + 0.5 * attn_pred[neg].sum() / (neg.sum() + 1)
        nll = -torch.log_softmax(segs, dim=1).gather(1, cam.clamp(0, self.n_classes - 1).unsqueeze(1)).squeeze(1)
        valid = cam != 255
        bgm, fgm = valid & (cam == 0), valid & (cam != 0)
        loss = 0.5 * (nll[bgm].sum() / bgm.sum().clamp_min(1) + nll[fgm].sum() / fgm.sum().clamp_min(1)) + 0.1 * attn_loss
        opt.zero_grad()
        loss.backward()
        opt.step()
        dt = time.perf_counter() - t0
        return n_img / dt, (f"{n_img} of the step's {self.BATCH} crops: oracle port of the same step (forward incl. pseudo-labels; the "
                            "frozen encoder runs twice because the oracle's forward does not return its taps; decoder forward + "
                            "torch autograd backward + AdamW), torch CPU fp32")


class InputPipelineVoc:
    """the dataset side of the evaluation loop (datasets/voc.py:25-95: PIL.Image.open -> np.asarray ->
    normalize_img -> transpose): baseline JPEGs of VOC-like sizes (4:2:0, quality 90) -> fp32 [3,H,W] on the device.
    `value`: Huffman-decoded coefficient blocks already resident in HBM (the device half: dequantise + islow IDCT, fancy
    up-sampling, fixed-point colour conversion, fp64-exact normalise); `e2e`: compressed bytes in host memory -> host marker
    parser + Huffman decoder (C, one thread) -> pinned coefficients -> H2D -> device half -> one pixel read back.
    Unit = one image."""
    name = "io: VOC12SegDataset JPEG decode + normalise, 16 baseline 4:2:0 JPEGs of VOC-like sizes (long side 500), quality 90"
    metric = "JPEG decode + normalise images/sec"
    cpu_kind = "reference"      # PIL / libjpeg + numpy is the reference's own code path for this row
    cpu_cores = 2               # test_msc_flip_voc.py:36 DataLoader(num_workers=2); our host half uses two threads as well
    host_threads = 2
    units = 16
    dtype = "i16/u8"
    cam_passes_per_image = 0
    l2_note = "16 images: 18 MB of coefficients + 36 MB of fp32 output per step (smaller than L2; not flushed)"

    def __init__(self, rank: int, device):
        from . import dataio
        self.dev, self.dataio = device, dataio
        self.streams = self._make_streams(500 + rank, self.units)
        self.infos, self.coef_h, self.coef_d = [], [], []
        for data in self.streams:
            info, coef = dataio.jpeg_entropy_decode(data)
            self.infos.append(info); self.coef_h.append(coef); self.coef_d.append(coef.to(device))
        L = _lib.lib()
        self.out = [torch.empty(3, i.height, i.width, dtype=torch.float32, device=device) for i in self.infos]
        self.ws = [torch.empty(int(L.wc_jpeg_workspace_bytes(ctypes.byref(i))), dtype=torch.uint8, device=device) for i in self.infos]
        self.mean = (ctypes.c_double * 3)(*dataio.IMAGENET_MEAN)
        self.std = (ctypes.c_double * 3)(*dataio.IMAGENET_STD)
        self.h2d = sum(c.numel() * 2 for c in self.coef_h)
        self.d2h = 4 * self.units
        self.px = sum(i.height * i.width for i in self.infos)
        self.pix_h = torch.zeros(self.units).pin_memory()
        self.launches = 0
        self._ms_dev = None

    @staticmethod
    def _make_streams(seed, n):
        """n synthetic photographs-like images (smooth structure + sensor noise) of VOC-like sizes as baseline 4:2:0 JPEGs."""
        import io
        from PIL import Image
        rng = np.random.default_rng(seed)
        out = []
        for i, (H, W) in enumerate(synth.draw_voc_like_sizes(rng, n)):
            y, x = np.mgrid[0:H, 0:W]
            base = np.stack([127 + 100 * np.sin(x / 37.0 + i), 127 + 100 * np.cos(y / 29.0 - i), (x + y) % 256], -1)
            img = np.clip(base + rng.normal(0, 12, (H, W, 3)), 0, 255).astype(np.uint8)
            buf = io.BytesIO()
            Image.fromarray(img).save(buf, format="JPEG", quality=90, subsampling=2)
            out.append(buf.getvalue())
        return out

    def _reconstruct(self, i, coef):
        L = _lib.lib()
        rc = L.wc_jpeg_reconstruct(ctypes.byref(self.infos[i]), _lib.ptr(coef), ctypes.c_void_p(0), _lib.ptr(self.out[i]), self.mean,
                                   self.std, _lib.ptr(self.ws[i]), ctypes.c_size_t(self.ws[i].numel()), _lib.stream_ptr())
        _lib.check(rc, "wc_jpeg_reconstruct")

    def step_device(self, record=False):
        for i in range(self.units):
            self._reconstruct(i, self.coef_d[i])

    def step_e2e(self):
        if not hasattr(self, "_pool"):
            from concurrent.futures import ThreadPoolExecutor
            self._pool = ThreadPoolExecutor(self.host_threads)
        # host half on `host_threads` threads (the C decoder runs without the GIL), into a pinned ring reused once the step has synced
        futs = [self._pool.submit(self.dataio.jpeg_entropy_decode, data, True, self.coef_h[i]) for i, data in enumerate(self.streams)]
        for i, f in enumerate(futs):
            info, coef = f.result()
            self._reconstruct(i, coef.to(self.dev, non_blocking=True))
            self.pix_h[i:i + 1].copy_(self.out[i][0, 0, :1], non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self.pix_h

    def profile(self, steps: int):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            self.step_device()
        e1.record(); torch.cuda.synchronize()
        self._ms_dev = e0.elapsed_time(e1) / steps
        self.launches = 3 * self.units          # IDCT (all components), up-sample + colour, normalise: per image

    def roofline(self, peaks):
        # algorithmic bytes per image (4:2:0): coefficients 2 B x 1.5 samples / px in, u8 planes 1.5 B / px written and read back by
        # the colour pass, fp32 [3,H,W] out
        by = self.px * (2 * 1.5 + 1.5 + 1.5 + 12)
        a = by / (self._ms_dev * 1e-3) / 1e9
        return {"bound": "hbm", "kernel": "jpeg_idct + jpeg_color (wc_jpeg_reconstruct)", "achieved": round(a, 1),
                "peak": peaks["hbm_gbs"], "peak_source": peaks["source"] + " (copy)", "unit": "GB/s",
                "frac": round(a / peaks["hbm_gbs"], 4), "traffic": None,
                "note": "18 B per pixel (4:2:0): 3 B coefficients + 3 B planes + 12 B fp32 output; images of ~0.19 MP each are "
                        "launch-bound (3 launches of a few microseconds per image)"}

    @classmethod
    def cpu(cls):
        return cls.__new__(cls)

    def cpu_sample(self, n_img=16):
        import io
        from PIL import Image
        if not hasattr(self, "streams"):
            self.streams = self._make_streams(500, self.units)
        mean = np.array([123.675, 116.28, 103.53]); std = np.array([58.395, 57.12, 57.375])
        from concurrent.futures import ThreadPoolExecutor
        reps = 20

        def load(data):
            arr = np.asarray(Image.open(io.BytesIO(data)).convert("RGB")).astype(np.float32)
            return np.ascontiguousarray(np.transpose((arr - mean) / std, (2, 0, 1)))
        with ThreadPoolExecutor(self.cpu_cores) as pool:
            t0 = time.perf_counter()
            for _ in range(reps):
                list(pool.map(load, self.streams))
            dt = time.perf_counter() - t0
        return reps * len(self.streams) / dt, (f"{reps} x {len(self.streams)} images: PIL (libjpeg-turbo) decode + the reference's numpy "
                                              f"normalise / transpose (datasets/voc.py), {self.cpu_cores} host threads as the "
                                              "reference's DataLoader(num_workers=2) — the reference's own code path")


WORKLOADS = {"mscflip": MscFlipVoc, "mscflip_coco": MscFlipCoco, "cfg2": PseudoLabelVoc, "train": TrainStepVoc,
             "io": InputPipelineVoc}

"""WeCLIP model wrapper — CUDA (sm_100a) drop-in for WeCLIP_model/model_attn_aff_voc.py:60-175 and its COCO twin
model_attn_aff_coco.py:56-170.

Same ``forward(img, img_names, mode) -> (seg, cam_labels, attn_pred)`` contract, same attributes callers touch
(``iter_num``, ``par``, ``cam_bg_thres``, ``root_path``, ``fg_text_features``, ``bg_text_features``) and the same
``state_dict`` key names (``encoder.visual.*``, ``decoder_fts_fuse.*``, ``decoder.*``, ``par.kernel``).

Out of scope by SURVEY.md §2 (init-time only): the CLIP text tower / tokenizer.  ``fg_text_features`` /
``bg_text_features`` are therefore inputs (ctor kwargs or attributes), not computed here.

Trainable decoder: ``decoder_fts_fuse`` and ``decoder`` are sub-modules holding live ``nn.Parameter``s
(``get_param_groups()[3]``, ``state_dict()``); under grad mode ``seg`` and ``attn_pred`` come from
``train.DecoderFunction`` (tcgen05 forward + hand-written backward, Dropout2d(0.1) while ``self.training``) and carry an
autograd graph, so scripts/dist_clip_voc.py:246-267 (loss, backward, optimizer.step) runs against this module.  Under
``torch.no_grad()`` the fused inference kernels run (wc_decoder_fwd).  The CLIP encoder is frozen (the reference never
steps it: its parameter groups 0-2 are empty, model_attn_aff_voc.py:94-103).
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from .engine import Engine
from .par import PAR, argmax_lut
from .train import DecoderFunction, DecoderTransformer, SegFormerHead, decoder_parameters, make_dropout_mask

_DATASETS = {
    # window, box threshold, RFM switch iteration, label sub-directory   (clip_tool.py:155,179 / :277,301;
    # model_attn_aff_voc.py:86,146 / model_attn_aff_coco.py:78,135,141)
    "voc": dict(window=6, box_thr=0.4, switch=15000, subdir="SegmentationClassAug"),
    "coco": dict(window=10, box_thr=0.7, switch=40000, subdir=os.path.join("SegmentationClass", "train")),
}


class LabelReader:
    """clip_tool.py:111-118: the image-level label set is read from the ground-truth PNG inside forward().
    The decoded (ids, height, width) are memoised per path + mtime: msc+flip evaluation opens the same PNG
    2 x n_scales times per dataset image."""

    def __init__(self):
        self._cache: Dict[str, Tuple[float, List[int], int, int]] = {}

    def __call__(self, path: str):
        from PIL import Image
        mtime = os.path.getmtime(path)          # FileNotFoundError like the reference's Image.open
        hit = self._cache.get(path)
        if hit is not None and hit[0] == mtime:
            return hit[1], hit[2], hit[3]
        lab = np.asarray(Image.open(path))
        ids = (np.unique(lab) - 1).tolist()     # uint8 wrap: 0 -> 255 (bg), 255 -> 254 (ignore)
        ids = [int(i) for i in ids if i not in (254, 255)]
        if len(self._cache) > 4096:
            self._cache.clear()
        self._cache[path] = (mtime, ids, lab.shape[0], lab.shape[1])
        return ids, lab.shape[0], lab.shape[1]


def load_clip_checkpoint(path: str) -> dict:
    """clip/clip.py:127-140 (clip.load): the configured checkpoint (configs/*.yaml clip_pretrain_path: ViT-B-16.pt) is a
    TorchScript archive, so ``torch.jit.load`` is tried first and its ``state_dict()`` taken; a plain pickled state dict
    (optionally wrapped as ``{"state_dict": ...}``) is the fallback, exactly as in the reference."""
    try:
        return dict(torch.jit.load(path, map_location="cpu").eval().state_dict())
    except RuntimeError:
        obj = torch.load(path, map_location="cpu", weights_only=False)
    if isinstance(obj, dict) and "state_dict" in obj and isinstance(obj["state_dict"], dict):
        obj = obj["state_dict"]
    if hasattr(obj, "state_dict") and not isinstance(obj, dict):
        obj = obj.state_dict()
    return dict(obj)


def _to_device_async(t: torch.Tensor, device) -> torch.Tensor:
    """Small host tensor -> device without a stream synchronisation (pinned staging, asynchronous copy)."""
    if torch.device(device).type != "cuda":
        return t.to(device)
    return t.pin_memory().to(device, non_blocking=True)


class ForwardMeta:
    """What forward() derives from the image names before any kernel runs (see WeCLIP.prepare)."""
    mode = "val"
    cam_path = False
    require_seg_trans = True
    nl = 0
    n_cam = 0
    ids_all = size = text_off = k_off = ch_off = rows = lut = rows_host = lut_host = None

    def signature(self):
        """Everything that shapes the launches (grid sizes, by-value tables): equal signatures can share a CUDA graph."""
        return (self.mode, self.cam_path, self.require_seg_trans, self.nl, self.n_cam,
                tuple(self.k_off) if self.k_off else None, self.size)


class WeCLIP(nn.Module):
    def __init__(self, num_classes=None, clip_model=None, embedding_dim=256, in_channels=512, dataset_root_path=None,
                 device="cuda", dataset: str = "voc", fg_text_features: Optional[torch.Tensor] = None,
                 bg_text_features: Optional[torch.Tensor] = None, state_dict: Optional[dict] = None):
        """``clip_model``: path to a state dict with the reference's CLIP keys (``visual.*``) — or pass a full
        WeCLIP ``state_dict`` directly.  ``dataset`` selects the VOC / COCO constants."""
        super().__init__()
        assert embedding_dim == 256, "the decoder kernels are instantiated for embedding_dim=256"
        self.num_classes = num_classes
        self.embedding_dim = embedding_dim
        self.in_channels = in_channels
        self.dataset = dataset
        self.cfg = _DATASETS[dataset]
        self.root_path = os.path.join(dataset_root_path or "", self.cfg["subdir"].split(os.sep)[0])
        self._label_subdir = os.sep.join(self.cfg["subdir"].split(os.sep)[1:])
        self.cam_bg_thres = 1
        self.iter_num = 0
        self.require_all_fts = True
        self.fg_text_features = fg_text_features
        self.bg_text_features = bg_text_features
        self.par = PAR(num_iter=20, dilations=[1, 2, 4, 8, 12, 24], affinity_dtype=torch.float16)
        self.label_reader = LabelReader()
        self.keep_intermediates = False      # debugging / test hook: keep raw CAMs, box masks, refined CAMs in self.last
        self.last: dict = {}
        self._device = torch.device(device)
        self._sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()      # frozen encoder weights (reference key names)
        self.engine: Optional[Engine] = None
        self.decoder_fts_fuse = SegFormerHead()
        self.decoder = DecoderTransformer(output_dim=num_classes or 21)
        self._dec_versions = None
        if state_dict is None and clip_model is not None:
            clip_sd = load_clip_checkpoint(clip_model)
            state_dict = {("encoder." + k.replace("module.", "")): v for k, v in clip_sd.items()
                          if k.replace("module.", "").startswith(("visual.", "logit_scale"))}
        if state_dict is not None:
            self.load_state_dict(state_dict, strict=False)

    # ---- state dict with the reference's key names -------------------------------------------------------
    def _decoder_state(self) -> "OrderedDict[str, torch.Tensor]":
        sd = OrderedDict()
        for prefix, mod in (("decoder_fts_fuse.", self.decoder_fts_fuse), ("decoder.", self.decoder)):
            for k, v in mod.state_dict().items():
                sd[prefix + k] = v
        return sd

    def state_dict(self, *args, **kwargs):
        sd = OrderedDict(self._sd)
        sd.update({k: v.detach() for k, v in self._decoder_state().items()})
        sd["par.kernel"] = self.par.kernel
        return sd

    def load_state_dict(self, state_dict, strict: bool = False):
        """test_msc_flip_voc.py:196 loads with strict=False; encoder keys may come from the CLIP checkpoint."""
        enc = {k: v for k, v in state_dict.items() if k.startswith(("encoder.visual.", "encoder.logit_scale"))}
        dec = {k: v for k, v in state_dict.items() if k.startswith(("decoder_fts_fuse.", "decoder."))}
        self._sd.update({k: v.detach().float().cpu() for k, v in enc.items()})
        if "decoder.linear_pred.weight" in dec and dec["decoder.linear_pred.weight"].shape[0] != self.decoder.linear_pred.out_channels:
            self.decoder = DecoderTransformer(output_dim=int(dec["decoder.linear_pred.weight"].shape[0]))
        own = self._decoder_state()
        loaded = set()
        with torch.no_grad():
            for k, v in dec.items():
                if k in own and own[k].shape == v.shape:
                    own[k].copy_(v.detach().to(own[k].dtype))
                    loaded.add(k)
        self._dec_loaded = getattr(self, "_dec_loaded", set()) | loaded
        need = ("encoder.visual.conv1.weight",)
        have_dec = {"decoder.linear_pred.weight", "decoder_fts_fuse.linear_fuse.weight"} <= self._dec_loaded
        if all(k in self._sd for k in need) and have_dec and torch.cuda.is_available():
            self.decoder_fts_fuse.to(self._device)
            self.decoder.to(self._device)
            full = OrderedDict(self._sd)
            full.update(self._decoder_state())
            self.engine = Engine(full, self.num_classes, self._device)
            self.num_classes = self.engine.num_classes
            self._dec_versions = self._param_versions()
        missing = [k for k in ("encoder.visual.conv1.weight",) if k not in self._sd] + \
                  [k for k in ("decoder.linear_pred.weight", "decoder_fts_fuse.linear_fuse.weight") if k not in self._dec_loaded]
        unexpected = [k for k in state_dict if k not in enc and k not in loaded and k != "par.kernel"]
        if strict and (missing or unexpected):
            raise RuntimeError(f"missing {missing}, unexpected {unexpected}")
        return nn.modules.module._IncompatibleKeys(missing, unexpected)

    def get_param_groups(self):
        """model_attn_aff_voc.py:94-103: [[], [], [], decoder + fuse parameters] — the live nn.Parameters."""
        groups = [[], [], [], []]
        groups[3] += list(self.decoder.parameters())
        groups[3] += list(self.decoder_fts_fuse.parameters())
        return groups

    def _param_versions(self):
        return tuple((p.data_ptr(), p._version) for p in decoder_parameters(self.decoder_fts_fuse, self.decoder))

    def _sync_decoder_weights(self):
        """The fused inference kernels read fp16 device copies of the decoder weights: refresh them (in place, same
        addresses, so captured CUDA graphs stay valid) after an optimiser step or a manual edit of the parameters."""
        ver = self._param_versions()
        if ver != self._dec_versions:
            self.engine.refresh_decoder(self._decoder_state())
            self._dec_versions = ver

    # ---- forward ---------------------------------------------------------------------------------------------
    def _label_path(self, name) -> str:
        return os.path.join(self.root_path, self._label_subdir, str(name) + ".png")

    @torch.no_grad()
    def _text_table(self, device):
        """[fg; bg] text features as one fp32 device table (rebuilt only when the attributes change)."""
        fg, bg = self.fg_text_features, self.bg_text_features
        key = (fg.data_ptr(), bg.data_ptr(), fg._version, bg._version, tuple(fg.shape), tuple(bg.shape), str(device))
        if getattr(self, "_text_key", None) != key:
            self._text_tbl = torch.cat([fg.to(device, torch.float32), bg.to(device, torch.float32)], dim=0).contiguous()
            self._text_key = key
            if torch.device(device).type == "cuda":
                torch.cuda.current_stream(device).synchronize()   # one-time: other streams may read the table next
        return self._text_tbl, fg.shape[0], bg.shape[0]

    def prepare(self, img_names, mode: str, N: int, H: int, W: int, device_tensors: bool = True) -> "ForwardMeta":
        """Host half of forward(): read the label sets (clip_tool.py:111-118), build the per-image text-row / class-offset
        tables and the valid_key LUT.  The two index vectors travel to the device from pinned memory (no host blocking);
        ``device_tensors=False`` leaves them on the host (CUDA-graph replay copies them into the graph's own buffers)."""
        cfg = self.cfg
        if isinstance(img_names, str):
            img_names = [img_names]
        m = ForwardMeta()
        m.mode = mode
        m.cam_path = not (self.dataset == "coco" and mode == "val") and len(img_names) > 0   # model_attn_aff_coco.py:131-132
        m.require_seg_trans = self.iter_num > cfg["switch"] or mode == "val"
        m.nl = cfg["window"] if m.require_seg_trans else 8                 # clip_tool.py:155 / :173
        m.n_cam = len(img_names)
        if m.n_cam > N:
            raise IndexError("index %d is out of bounds for dimension 0 with size %d" % (N, N))   # img[i], reference :141
        if not m.cam_path:
            return m
        fg, bg = self.fg_text_features, self.bg_text_features
        n_fg, n_bg = fg.shape[0], bg.shape[0]
        ids_all, sizes, text_off, k_off, rows = [], [], [0], [0], []
        for name in img_names:
            ids, oh, ow = self.label_reader(self._label_path(name))
            ids_all.append(ids)
            sizes.append((H, W) if mode == "train" else (oh, ow))     # clip_tool.py:194-197
            rows += list(ids) + list(range(n_fg, n_fg + n_bg))
            text_off.append(text_off[-1] + len(ids) + n_bg)
            k_off.append(k_off[-1] + len(ids))
        if any(len(i) == 0 for i in ids_all):
            raise RuntimeError("image without foreground label (torch.stack([]) in the reference, clip_tool.py:212)")
        if any(i < 0 or i >= n_fg for ids in ids_all for i in ids):
            raise IndexError("label id outside the foreground text features")
        if len(set(sizes)) != 1:
            raise RuntimeError("stack expects each tensor to be equal size (model_attn_aff_voc.py:173)")
        m.ids_all, m.size, m.text_off, m.k_off = ids_all, sizes[0], text_off, k_off
        m.ch_off = [k_off[i] + i for i in range(m.n_cam + 1)]
        m.rows_host = torch.tensor(rows, dtype=torch.int64)
        m.lut_host = torch.tensor([v for ids in ids_all for v in [0] + [i + 1 for i in ids]], dtype=torch.int64)
        if device_tensors:
            dev = self.engine.device if self.engine is not None else self._device
            m.rows, m.lut = _to_device_async(m.rows_host, dev), _to_device_async(m.lut_host, dev)
        return m

    def run(self, img, m: "ForwardMeta"):
        """Device half of forward(): every launch goes to the current stream; nothing here blocks the host or reads a
        host value back, so the whole call can be captured into a CUDA graph (evaluate.MscFlipEvaluator, use_graphs)."""
        eng, cfg = self.engine, self.cfg
        N, _, H, W = img.shape
        h, w = H // 16, W // 16
        n_cam, nl = m.n_cam, m.nl
        taps, x_last, attn = eng.encoder(img, n_attn=(nl - 1) if m.cam_path else 0, attn_slots=nl if m.cam_path else 0,
                                         attn_imgs=n_cam if m.cam_path else None)
        params = decoder_parameters(self.decoder_fts_fuse, self.decoder)
        if torch.is_grad_enabled() and any(p.requires_grad for p in params):
            # training step (scripts/dist_clip_voc.py:246-267): seg / attn_pred carry the autograd graph of the decoder
            mask = make_dropout_mask(N, self.decoder_fts_fuse.dropout.p, self.training and self.decoder_fts_fuse.training,
                                     eng.device)
            seg, attn_pred_g = DecoderFunction.apply(taps, (N, h, w), mask, *params)
            attn_pred = attn_pred_g.detach()          # the CAM / PAR branch is detached (clip_tool.py:169-175)
        else:
            self._sync_decoder_weights()
            seg, attn_pred, _ = eng.decoder(taps, N, h, w)
            attn_pred_g = attn_pred
        if not m.cam_path:
            return seg, None, attn_pred_g
        # the text rows are gathered on the device from a cached [fg; bg] table with the index vector of prepare()
        table, _, _ = self._text_table(eng.device)
        text = table.index_select(0, m.rows)
        sums = eng.last_attn_sums
        cams, _, _ = eng.gradcam(x_last, n_cam, h, w, text, m.text_off, m.k_off, attn_last_out=attn[nl - 1],
                                 attn_last_sum=sums[nl - 1])
        refined, bmasks = eng.affinity_refine(attn, nl, attn_pred, n_cam, h, w, m.require_seg_trans, cams, m.k_off,
                                              cfg["box_thr"], want_masks=self.keep_intermediates, attn_sums=sums)
        if self.keep_intermediates:
            self.last = dict(cams=cams, box_masks=bmasks, refined=refined, k_off=m.k_off, ids=m.ids_all,
                             rfm_sel=eng.last_rfm_sel)
        Hc, Wc = m.size
        par_in = eng.cam_upsample_bg(refined, m.k_off, n_cam, h, w, Hc, Wc, float(self.cam_bg_thres))
        par_out = self.par.refine_ragged(img[:n_cam], par_in, m.ch_off)
        labels = argmax_lut(par_out, m.lut, m.ch_off)
        return seg, labels, attn_pred_g

    def forward(self, img, img_names="2007_000032", mode="train"):
        _lib.require_device()
        if self.engine is None:
            raise _lib.WeclipError("WeCLIP weights are not loaded (encoder + decoder state dict required)")
        img = img.to(self.engine.device, torch.float32)
        N, _, H, W = img.shape
        self.iter_num += 1
        # the CAM / PAR path runs once per NAME (model_attn_aff_voc.py:139 `for i, img_name in enumerate(img_names)`); the
        # reference's msc+flip driver passes names = name+name (test_msc_flip_voc.py:52), i.e. one name per image of its
        # [image, flipped image] batch.  Fewer names than images is allowed (only the leading images get pseudo-labels):
        # evaluate.MscFlipEvaluator(cam_passes="live") uses that to skip passes whose labels the driver never reads
        return self.run(img, self.prepare(img_names, mode, N, H, W))

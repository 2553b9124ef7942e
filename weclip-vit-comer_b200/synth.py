"""Seeded synthetic weights / images / labels (there is no network for datasets or checkpoints).

Everything here is deterministic CPU torch/numpy so that the GPU box regenerates bit-identical
inputs from a seed; tests/golden/ therefore only stores *outputs* of the reference.

State-dict key names follow the reference: ``encoder.visual.*`` (clip/model.py:246-262),
``decoder_fts_fuse.*`` (WeCLIP_model/segformer_head.py:49-66), ``decoder.*``
(WeCLIP_model/Decoder/TransDecoder.py:104-110), ``par.kernel`` (WeCLIP_model/PAR.py:33).

All weight values are exactly representable in bf16 *and* fp16 (8-bit mantissa, |w| >= 2^-14 or 0), so
the reference's fp16 ``convert_weights`` (clip/model.py:461-476), the fp32 oracle and the bf16 CUDA path
all see the same numbers.
"""
from __future__ import annotations

import math
import os
from collections import OrderedDict

import numpy as np
import torch

VIT_WIDTH = 768
VIT_HEADS = 12
VIT_LAYERS = 12
VIT_PATCH = 16
VIT_GRID = 14
EMBED_DIM = 512
DEC_WIDTH = 256
DEC_HEADS = 8
DEC_LAYERS = 3
N_TAPS = 11

# K (classes per image) histogram of VOC train_aug, SURVEY.md §8d [probed]
VOC_K_HIST = {1: 0.596, 2: 0.289, 3: 0.091, 4: 0.024}


def bf16_exact(t: torch.Tensor) -> torch.Tensor:
    """Round to bf16 and flush magnitudes below fp16's normal range, return as fp32."""
    t = t.to(torch.bfloat16).to(torch.float32)
    return torch.where(t.abs() < 2.0 ** -14, torch.zeros_like(t), t)


def _randn(g, shape, std):
    return bf16_exact(torch.randn(shape, generator=g, dtype=torch.float32) * std)


def _ln(g, width, prefix, sd):
    sd[prefix + ".weight"] = bf16_exact(1.0 + 0.1 * torch.randn(width, generator=g))
    sd[prefix + ".bias"] = _randn(g, (width,), 0.05)


def _resblock(g, sd, prefix, width, qk_std):
    w_in = torch.randn(3 * width, width, generator=g, dtype=torch.float32)
    w_in[: 2 * width] *= qk_std
    w_in[2 * width:] *= width ** -0.5
    sd[prefix + ".attn.in_proj_weight"] = bf16_exact(w_in)
    sd[prefix + ".attn.in_proj_bias"] = _randn(g, (3 * width,), 0.02)
    sd[prefix + ".attn.out_proj.weight"] = _randn(g, (width, width), 0.7 * width ** -0.5)
    sd[prefix + ".attn.out_proj.bias"] = _randn(g, (width,), 0.02)
    _ln(g, width, prefix + ".ln_1", sd)
    sd[prefix + ".mlp.c_fc.weight"] = _randn(g, (4 * width, width), width ** -0.5)
    sd[prefix + ".mlp.c_fc.bias"] = _randn(g, (4 * width,), 0.02)
    sd[prefix + ".mlp.c_proj.weight"] = _randn(g, (width, 4 * width), 0.7 * (4 * width) ** -0.5)
    sd[prefix + ".mlp.c_proj.bias"] = _randn(g, (width,), 0.02)
    _ln(g, width, prefix + ".ln_2", sd)


def make_state_dict(seed: int = 0, num_classes: int = 21) -> "OrderedDict[str, torch.Tensor]":
    """Random-init ViT-B/16 visual tower + WeCLIP decoder, fp32 CPU tensors with bf16-exact values."""
    g = torch.Generator().manual_seed(seed)
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    W = VIT_WIDTH
    v = "encoder.visual"
    sd[v + ".class_embedding"] = _randn(g, (W,), W ** -0.5)
    sd[v + ".positional_embedding"] = _randn(g, (VIT_GRID * VIT_GRID + 1, W), 0.3)
    sd[v + ".proj"] = _randn(g, (W, EMBED_DIM), W ** -0.5)
    sd[v + ".conv1.weight"] = _randn(g, (W, 3, VIT_PATCH, VIT_PATCH), 0.03)
    _ln(g, W, v + ".ln_pre", sd)
    for i in range(VIT_LAYERS):
        _resblock(g, sd, f"{v}.transformer.resblocks.{i}", W, qk_std=0.05)
    _ln(g, W, v + ".ln_post", sd)
    sd["encoder.logit_scale"] = torch.tensor(math.log(1 / 0.07), dtype=torch.float32)
    D = DEC_WIDTH
    for i in range(N_TAPS):
        p = f"decoder_fts_fuse.linears_modulelist.{i}"
        sd[p + ".proj.weight"] = _randn(g, (D, W), W ** -0.5)
        sd[p + ".proj.bias"] = _randn(g, (D,), 0.02)
        sd[p + ".proj_2.weight"] = _randn(g, (D, D), D ** -0.5)
        sd[p + ".proj_2.bias"] = _randn(g, (D,), 0.02)
    sd["decoder_fts_fuse.linear_fuse.weight"] = _randn(g, (D, D * N_TAPS, 1, 1), 0.25 * (D * N_TAPS) ** -0.5)
    sd["decoder_fts_fuse.linear_fuse.bias"] = _randn(g, (D,), 0.02)
    for i in range(DEC_LAYERS):
        _resblock(g, sd, f"decoder.transformer.resblocks.{i}", D, qk_std=0.08)
    sd["decoder.linear_pred.weight"] = _randn(g, (num_classes, D, 1, 1), D ** -0.5)
    sd["decoder.linear_pred.bias"] = _randn(g, (num_classes,), 0.02)
    return sd


def make_text_features(seed: int, n_fg: int, n_bg: int):
    """Unit-norm Gaussian stand-ins for zeroshot_classifier output (model_attn_aff_voc.py:34-46,81-82).

    The text tower is init-time only and out of scope; the build takes these as inputs.
    """
    g = torch.Generator().manual_seed(seed)
    fg = torch.randn(n_fg, EMBED_DIM, generator=g)
    bg = torch.randn(n_bg, EMBED_DIM, generator=g)
    fg = fg / fg.norm(dim=-1, keepdim=True)
    bg = bg / bg.norm(dim=-1, keepdim=True)
    return fg.contiguous(), bg.contiguous()


def make_images(seed: int, n: int, H: int, W: int, smooth: bool = True) -> torch.Tensor:
    """[n,3,H,W] fp32 in the ImageNet-normalised range; ``smooth`` blurs so PAR affinities are not degenerate."""
    g = torch.Generator().manual_seed(seed)
    if not smooth:
        return torch.randn(n, 3, H, W, generator=g)
    # low-resolution noise upsampled bicubically + a little pixel noise: smooth regions with edges
    lo = torch.randn(n, 3, max(2, H // 24 + 1), max(2, W // 24 + 1), generator=g)
    img = torch.nn.functional.interpolate(lo, size=(H, W), mode="bicubic", align_corners=False)
    img = 1.3 * torch.tanh(1.5 * img) + 0.05 * torch.randn(n, 3, H, W, generator=g)
    return img.contiguous()


def draw_label_set(rng: np.random.Generator, n_classes_fg: int = 20, k_hist=None) -> list:
    k_hist = k_hist or VOC_K_HIST
    ks, ps = zip(*sorted(k_hist.items()))
    k = int(rng.choice(ks, p=np.array(ps) / sum(ps)))
    return sorted(int(c) for c in rng.choice(n_classes_fg, size=k, replace=False))


def make_label_map(rng: np.random.Generator, H: int, W: int, class_ids) -> np.ndarray:
    """uint8 label map: 0 = background, c+1 = foreground class c, a 255 'ignore' border stripe."""
    lab = np.zeros((H, W), dtype=np.uint8)
    for c in class_ids:
        h = int(rng.integers(max(2, H // 6), max(3, H // 2)))
        w = int(rng.integers(max(2, W // 6), max(3, W // 2)))
        y = int(rng.integers(0, max(1, H - h)))
        x = int(rng.integers(0, max(1, W - w)))
        lab[y:y + h, x:x + w] = c + 1
    for c in class_ids:  # every drawn class must survive occlusion by later rectangles
        if not (lab == c + 1).any():
            lab[int(rng.integers(0, H)), int(rng.integers(0, W))] = c + 1
    lab[0, : max(1, W // 8)] = 255
    return lab


def write_label_png(root: str, subdir: str, name: str, lab: np.ndarray) -> str:
    """Write where the reference looks: <root>/SegmentationClassAug/<name>.png (model_attn_aff_voc.py:86,140)."""
    from PIL import Image
    d = os.path.join(root, subdir)
    os.makedirs(d, exist_ok=True)
    path = os.path.join(d, str(name) + ".png")
    Image.fromarray(lab).save(path)
    return path


# (H0, W0) of VOC-style JPEGs (long side 500) with rough frequencies: after the evaluation's long-side-512 resize
# (test_msc_flip_voc.py:60-63, short side int(h * 512/500)) most short sides are NOT multiples of 16 (341, 340, 342, 287 ...)
VOC_LIKE_SIZES = [((375, 500), 0.42), ((500, 375), 0.13), ((333, 500), 0.14), ((500, 333), 0.05), ((334, 500), 0.04),
                  ((332, 500), 0.04), ((374, 500), 0.04), ((281, 500), 0.03), ((500, 334), 0.02), ((500, 500), 0.02),
                  ((400, 500), 0.02), ((356, 500), 0.02), ((480, 360), 0.02), ((335, 500), 0.01)]


def draw_voc_like_sizes(rng: np.random.Generator, n: int):
    sizes, ps = zip(*VOC_LIKE_SIZES)
    idx = rng.choice(len(sizes), size=n, p=np.array(ps) / sum(ps))
    return [sizes[i] for i in idx]


def make_voc_like_dataset(seed: int, n: int, root: str, subdir: str = "SegmentationClassAug", n_fg: int = 20, prefix: str = "v"):
    """n dataset items (name, image [3,H0,W0] fp32, label [H0,W0] uint8) with VOC-like sizes; label PNGs written under root."""
    rng = np.random.default_rng(seed)
    items = []
    for i, (h0, w0) in enumerate(draw_voc_like_sizes(rng, n)):
        lab = make_label_map(rng, h0, w0, draw_label_set(rng, n_fg))
        name = f"{prefix}{seed}_{i}"
        write_label_png(root, subdir, name, lab)
        items.append((name, make_images(7000 + seed * 131 + i, 1, h0, w0)[0], lab))
    return items

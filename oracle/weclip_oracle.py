"""ORACLE — CPU restatement of the reference's WeCLIP hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs
may import this file; the product path (``weclip-vit-comer_b200/``) must never import it.

Plain PyTorch (fp32) / numpy, functional style over a state dict with the reference's key names.  Tensors may live on
``cuda`` as well: then every op is the ATen / cuBLAS library call the reference itself makes on a GPU (with the
reference's host round trips for OpenCV / numpy), and ``gpu_mixed=True`` reproduces its hard-wired GPU mixed precision
(clip/model.py:461-476: fp16 conv1 / out-proj / MLP weights and activations, fp32 in-proj + softmax, clip/myAtt.py:199,321)
— that is bench.py's ``gpu_library_baseline`` leg, the "library-call bar" of SURVEY.md §8(d).
Every function cites the reference file:line it restates (paths relative to /root/reference).

Parity status: PINNED.  The reference has no tests or golden vectors of its own, so
``oracle/make_golden.py`` imports the *real* reference modules from /root/reference in the build container
(with the module stubs listed in SURVEY.md §8c), runs them on seeded synthetic inputs, and commits the
outputs under ``tests/golden/``; ``tests/test_oracle_golden.py`` checks this restatement against them.

Third-party arithmetic on the path that is not under /root/reference: OpenCV
(``cv2.resize`` INTER_LINEAR, ``cv2.findContours``/``boundingRect``; requirements.txt:9 pins 4.6.0.66,
4.13.0 is installed here).  Both are restated below from their published algorithm and pinned against
the installed cv2 through the golden vectors.
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor

PAR_DILATIONS = (1, 2, 4, 8, 12, 24)
PAR_NUM_ITER = 20


# --------------------------------------------------------------------------------------------------
# A.1 encoder
# --------------------------------------------------------------------------------------------------
def upsample_pos_emb(emb: Tensor, new_size: Tuple[int, int]) -> Tensor:
    """clip/model.py:11-27 — bilinear (align_corners=False) resize of the 14x14 grid, result rounded to fp16."""
    first, grid = emb[:1], emb[1:]
    n, d = grid.shape
    s = int(math.isqrt(n))
    g = grid.permute(1, 0).reshape(1, d, s, s)
    g = F.interpolate(g, size=new_size, mode="bilinear", align_corners=False)
    g = g.reshape(d, -1).permute(1, 0)
    return torch.cat([first, g], 0).half().float()


def layer_norm(x: Tensor, w: Tensor, b: Tensor) -> Tensor:
    """clip/model.py:177-183 — fp32 LayerNorm, eps 1e-5."""
    return F.layer_norm(x.float(), (x.shape[-1],), w, b, 1e-5)


def quick_gelu(x: Tensor) -> Tensor:
    """clip/model.py:186-188."""
    return x * torch.sigmoid(1.702 * x)


_HALF_CACHE: Dict[int, Tensor] = {}


def _h(t: Tensor) -> Tensor:
    """fp16 copy of a weight (convert_weights, clip/model.py:461-476), cached per tensor."""
    key = id(t)
    hit = _HALF_CACHE.get(key)
    if hit is None or hit.device != t.device or hit.shape != t.shape:
        hit = t.half()
        _HALF_CACHE[key] = hit
    return hit


def mha(x: Tensor, sd: Dict[str, Tensor], p: str, heads: int, mimic_fp16: bool = True):
    """clip/myAtt.py:199-201 (fp32 packed in-proj), :257-268 (head split), :53-64 (softmax(q/sqrt(d) k^T) v),
    :320-321 (fp16 out-proj), :323-326 (mean over heads).  x: [L,N,E] -> ([L,N,E], [N,L,L])."""
    L, N, E = x.shape
    hd = E // heads
    qkv = F.linear(x.float(), sd[p + ".in_proj_weight"], sd[p + ".in_proj_bias"])
    q, k, v = qkv.chunk(3, dim=-1)
    q = q.contiguous().view(L, N * heads, hd).transpose(0, 1) / math.sqrt(hd)
    k = k.contiguous().view(L, N * heads, hd).transpose(0, 1)
    v = v.contiguous().view(L, N * heads, hd).transpose(0, 1)
    attn = torch.softmax(torch.bmm(q, k.transpose(-2, -1)), dim=-1)
    o = torch.bmm(attn, v).transpose(0, 1).contiguous().view(L, N, E)
    wo, bo = sd[p + ".out_proj.weight"], sd[p + ".out_proj.bias"]
    if mimic_fp16 == "gpu":
        o = F.linear(o.half(), _h(wo), _h(bo))                        # stays fp16 on the GPU path (myAtt.py:321)
    elif mimic_fp16:
        o = F.linear(o.half(), wo.half(), bo.half()).float()
    else:
        o = F.linear(o, wo, bo)
    return o, attn.view(N, heads, L, L).sum(dim=1) / heads


def resblock(x: Tensor, sd: Dict[str, Tensor], p: str, heads: int, mimic_fp16: bool = True, ln1_out=None):
    """clip/model.py:210-214.  ``ln1_out`` lets the Grad-CAM oracle substitute the hooked ln_1 output."""
    a = layer_norm(x, sd[p + ".ln_1.weight"], sd[p + ".ln_1.bias"]) if ln1_out is None else ln1_out
    if mimic_fp16 == "gpu":
        # the reference's GPU dtype flow: fp16 residual stream, LayerNorm in fp32 cast back to fp16 (model.py:177-183),
        # attention input up-cast to fp32 (myAtt.py:199), fp16 out-proj and MLP
        x = x.half()
        o, w = mha(a.half() if ln1_out is None else a, sd, p + ".attn", heads, "gpu")
        x = x + o.to(x.dtype)
        h = layer_norm(x, sd[p + ".ln_2.weight"], sd[p + ".ln_2.bias"]).half()
        h = F.linear(h, _h(sd[p + ".mlp.c_fc.weight"]), _h(sd[p + ".mlp.c_fc.bias"]))
        h = F.linear(quick_gelu(h), _h(sd[p + ".mlp.c_proj.weight"]), _h(sd[p + ".mlp.c_proj.bias"]))
        return x + h, w
    o, w = mha(a, sd, p + ".attn", heads, mimic_fp16)
    x = x + o
    h = layer_norm(x, sd[p + ".ln_2.weight"], sd[p + ".ln_2.bias"])
    h = F.linear(h, sd[p + ".mlp.c_fc.weight"], sd[p + ".mlp.c_fc.bias"])
    h = F.linear(quick_gelu(h), sd[p + ".mlp.c_proj.weight"], sd[p + ".mlp.c_proj.bias"])
    return x + h, w


def encode_image(img: Tensor, sd: Dict[str, Tensor], mimic_fp16: bool = True):
    """clip/model.py:264-287 + :225-243 (11 of 12 blocks, no ln_post/proj) via clip_tool.py:31-41.
    img [N,3,H,W] -> taps: 11 x [L,N,768]; attn: 11 x [N,L,L]."""
    v = "encoder.visual"
    N, _, H, W = img.shape
    h, w = H // 16, W // 16
    pos = upsample_pos_emb(sd[v + ".positional_embedding"], (h, w))
    if mimic_fp16 == "gpu":
        x = F.conv2d(img.half(), _h(sd[v + ".conv1.weight"]), stride=16).float()
    else:
        x = F.conv2d(img.float(), sd[v + ".conv1.weight"], stride=16)
    x = x.reshape(N, x.shape[1], -1).permute(0, 2, 1)
    cls = sd[v + ".class_embedding"].reshape(1, 1, -1).expand(N, 1, -1)
    x = torch.cat([cls, x], dim=1) + pos
    x = layer_norm(x, sd[v + ".ln_pre.weight"], sd[v + ".ln_pre.bias"])
    x = x.permute(1, 0, 2)
    taps, attns = [], []
    for i in range(11):
        x, a = resblock(x, sd, f"{v}.transformer.resblocks.{i}", 12, mimic_fp16)
        taps.append(x)
        attns.append(a.float())
    return taps, attns


# --------------------------------------------------------------------------------------------------
# A.2 decoder
# --------------------------------------------------------------------------------------------------
def decoder_fts_fuse(taps: Sequence[Tensor], sd: Dict[str, Tensor], h: int, w: int) -> Tensor:
    """WeCLIP_model/segformer_head.py:69-80 (+ MLP :13-28) on the patch tokens of the 11 taps
    (model_attn_aff_voc.py:122-125).  Dropout2d is identity in eval.  -> [N,256,h,w]."""
    zs = []
    for i, t in enumerate(taps):
        p = f"decoder_fts_fuse.linears_modulelist.{i}"
        tok = t[1:].permute(1, 0, 2).float()  # [N,P,768]
        z = F.relu(F.linear(tok, sd[p + ".proj.weight"], sd[p + ".proj.bias"]))
        zs.append(F.linear(z, sd[p + ".proj_2.weight"], sd[p + ".proj_2.bias"]))
    cat = torch.cat(zs, dim=-1)  # [N,P,2816] (channel concat in layer order)
    wf = sd["decoder_fts_fuse.linear_fuse.weight"].flatten(1)
    fts = F.linear(cat, wf, sd["decoder_fts_fuse.linear_fuse.bias"])  # [N,P,256]
    N = fts.shape[0]
    return fts.permute(0, 2, 1).reshape(N, -1, h, w)


def decoder(fts: Tensor, sd: Dict[str, Tensor], mimic_fp16: bool = True) -> Tensor:
    """WeCLIP_model/Decoder/TransDecoder.py:113-125 — 3 blocks (256 wide, 8 heads) + 1x1 conv -> [N,C,h,w]."""
    N, C, h, w = fts.shape
    x = fts.reshape(N, C, h * w).permute(2, 0, 1)
    for i in range(3):      # the decoder's own weights stay fp32 on the GPU too (only its out-proj is forced to fp16)
        x, _ = resblock(x, sd, f"decoder.transformer.resblocks.{i}", 8, bool(mimic_fp16))
    x = x.permute(1, 2, 0).reshape(N, C, h, w)
    return F.conv2d(x, sd["decoder.linear_pred.weight"], sd["decoder.linear_pred.bias"])


def attn_pred_from_fts(fts: Tensor) -> Tensor:
    """model_attn_aff_voc.py:134-137 — sigmoid(F^T F) over flattened patches -> [N,P,P]."""
    N, C, h, w = fts.shape
    f = fts.reshape(N, C, h * w)
    return torch.sigmoid(f.transpose(2, 1).bmm(f))


# --------------------------------------------------------------------------------------------------
# A.3 Grad-CAM
# --------------------------------------------------------------------------------------------------
def scale_cam_image(img: np.ndarray) -> np.ndarray:
    """pytorch_grad_cam/utils/image.py:51-61 without the resize: (x-min)/(1e-7+max(x-min))."""
    img = img - np.min(img)
    return (img / (1e-7 + np.max(img))).astype(np.float32)


def forward_last_layer(x11: Tensor, text: Tensor, sd: Dict[str, Tensor], mimic_fp16: bool = True, ln1_out=None):
    """clip/model.py:407-429 — block 12, ln_post, mean over patch tokens, proj, cosine softmax.
    x11 [L,1,768], text [T,512] -> (p [1,T], attn [1,L,L])."""
    v = "encoder.visual"
    x, attn = resblock(x11, sd, f"{v}.transformer.resblocks.11", 12, mimic_fp16, ln1_out=ln1_out)
    x = x.permute(1, 0, 2)
    x = layer_norm(x, sd[v + ".ln_post.weight"], sd[v + ".ln_post.bias"])
    g = torch.mean(x[:, 1:, :], dim=1) @ sd[v + ".proj"]
    attn = attn.float()
    g = g / g.norm(dim=1, keepdim=True)
    t = text / text.norm(dim=1, keepdim=True)
    logits = sd["encoder.logit_scale"].exp() * g @ t.t()
    return logits.softmax(dim=-1), attn


def grad_cam(x11: Tensor, text: Tensor, targets: Sequence[int], sd: Dict[str, Tensor], h: int, w: int,
             mimic_fp16: bool = True):
    """pytorch_grad_cam/base_cam.py:62-108,116-154 + grad_cam.py:16-23 + activations_and_gradients.py:19-37
    with reshape_transform (model_attn_aff_voc.py:23-30) and ClipOutputTarget (clip_tool.py:20-26).

    Hook tensor A = ln_1 output of block 12; G_k = d p_k / d A; w = mean_hw(G) over patch tokens;
    cam = relu(sum_c w_c A_c); min-max applied twice (compute_cam_per_layer, aggregate_multi_layers).
    Returns (cams [K,h,w] float32 numpy, attn_last [1,L,L], probs [1,T])."""
    p = "encoder.visual.transformer.resblocks.11"
    cams = []
    probs = attn_last = None
    for k in targets:
        with torch.enable_grad():       # the reference runs Grad-CAM with autograd on whatever the caller's grad mode is
            A = layer_norm(x11, sd[p + ".ln_1.weight"], sd[p + ".ln_1.bias"]).detach().requires_grad_(True)
            probs, attn_last = forward_last_layer(x11.detach(), text, sd, mimic_fp16, ln1_out=A)
            probs[:, k].sum().backward()
        # .cpu(): the reference's own device -> host round trips (activations_and_gradients.py:24,35)
        act = A.detach()[1:, 0, :].float().reshape(h, w, -1).permute(2, 0, 1).cpu().numpy()       # [768,h,w]
        grd = A.grad[1:, 0, :].float().reshape(h, w, -1).permute(2, 0, 1).cpu().numpy()
        wts = grd.mean(axis=(1, 2))
        cam = np.maximum((wts[:, None, None] * act).sum(axis=0), 0).astype(np.float32)
        cam = scale_cam_image(cam)
        cam = scale_cam_image(np.maximum(cam, 0))
        cams.append(cam)
    return np.stack(cams, 0), attn_last.detach(), probs.detach()


# --------------------------------------------------------------------------------------------------
# A.4 affinity refinement
# --------------------------------------------------------------------------------------------------
def rfm_attention(attn_stack: Tensor, seg_attn: Tensor, require_seg_trans: bool, window: int = 6, layer_mask=None,
                  info: dict = None) -> Tensor:
    """clip/clip_tool.py:152-175 (VOC window -6; COCO -10 at :277; no-RFM branch: mean of last 8).
    attn_stack [12,L,L] (11 encoder maps + block 12), seg_attn [1,P,P] -> [P,P].
    ``layer_mask`` (test hook, like ``box_masks``) overrides the kept-layer decision ``diff <= mean(diff)``; ``info``
    receives that decision as the reference takes it (fp32 ``torch.sum`` over P^2 terms) and the same sums in fp64:
    at P >= 1024 the sums are ~1e6 and their differences between layers ~1, i.e. a few fp32 ulps — the reference's own
    decision then depends on torch's summation order (tests/parity.py::rfm_checks)."""
    aw = attn_stack[:, 1:, 1:]
    if require_seg_trans:
        aw = aw[-window:]
        diff = torch.sum((seg_attn - aw).flatten(1), dim=1)
        th = torch.mean(diff)
        mask = torch.zeros_like(diff)
        mask[diff <= th] = 1
        if info is not None:
            d64 = torch.sum((seg_attn.double() - aw.double()).flatten(1), dim=1)
            info.update(mask=mask.clone(), diff32=diff.clone(), diff64=d64, margin64=d64 - d64.mean(), P=int(aw.shape[-1]))
        if layer_mask is not None:
            mask = torch.as_tensor(layer_mask, dtype=diff.dtype, device=diff.device).reshape(-1)
        mask = mask.reshape(-1, 1, 1).expand_as(aw)
        out = torch.sum(mask * aw, dim=0) / (torch.sum(mask, dim=0) + 1e-5)
        return out * seg_attn.squeeze(0)
    return torch.mean(aw[-8:], dim=0)


def compute_trans_mat(a: Tensor) -> Tensor:
    """clip/clip_tool.py:64-80 — 3x (column-normalise, row-normalise), symmetrise, square."""
    t = a / torch.sum(a, dim=0, keepdim=True)
    t = t / torch.sum(t, dim=1, keepdim=True)
    for _ in range(2):
        t = t / torch.sum(t, dim=0, keepdim=True)
        t = t / torch.sum(t, dim=1, keepdim=True)
    t = (t + t.transpose(1, 0)) / 2
    return torch.matmul(t, t)


def box_mask(cam: np.ndarray, threshold: float) -> np.ndarray:
    """clip/utils.py:115-142 (scoremap2bbox, multi_contour_eval=True) + clip_tool.py:179-183, restated without
    OpenCV: uint8 truncation, strict '>' against int(thr*max), one box per 8-connected component
    (cv2.findContours(RETR_TREE) also returns hole contours, whose boxes lie inside their parent's),
    x1=min(x+w, W-1), y1=min(y+h, H-1), exclusive-upper slice fill.  -> float32 mask [h,w]."""
    from scipy import ndimage
    H, W = cam.shape
    u8 = (cam * 255).astype(np.uint8)
    binary = u8 > int(threshold * np.max(u8))
    lab, n = ndimage.label(binary, structure=np.ones((3, 3), dtype=bool))
    mask = np.zeros((H, W), dtype=np.float32)
    for sl in ndimage.find_objects(lab):
        y0, y1 = sl[0].start, min(sl[0].stop, H - 1)
        x0, x1 = sl[1].start, min(sl[1].stop, W - 1)
        mask[y0:y1, x0:x1] = 1
    return mask


def refine_cams(cams: np.ndarray, trans_mat: Tensor, threshold: float, masks=None) -> List[Tensor]:
    """clip/clip_tool.py:179-191 — cam' = (T * mask[None,:]) @ cam, per class.  ``masks`` (test hook) overrides the
    box masks so that a test can hold the discontinuous threshold decision fixed."""
    out = []
    for k, cam in enumerate(cams):
        h, w = cam.shape
        bm = box_mask(cam, threshold) if masks is None else np.asarray(masks[k], dtype=np.float32)
        m = torch.from_numpy(bm).view(1, h * w).to(trans_mat.device)            # .cuda() as clip_tool.py:180,188
        out.append(torch.matmul(trans_mat * m, torch.from_numpy(cam).view(-1, 1).to(trans_mat.device)).reshape(h, w))
    return out


# --------------------------------------------------------------------------------------------------
# A.5 upsample + background
# --------------------------------------------------------------------------------------------------
def cv2_resize_linear(src: np.ndarray, dsize: Tuple[int, int]) -> np.ndarray:
    """cv2.resize(src, (w,h)) INTER_LINEAR for a float32 single-channel image (OpenCV imgproc/resize.cpp:
    half-pixel centres, src index clamped with zero weight at the border, no antialias)."""
    dw, dh = dsize
    sh, sw = src.shape
    src = src.astype(np.float32)

    def taps(dn, sn):
        scale = float(sn) / dn
        f = (np.arange(dn, dtype=np.float64) + 0.5) * scale - 0.5
        i0 = np.floor(f).astype(np.int64)
        a = (f - i0).astype(np.float32)
        lo = i0 < 0
        i0[lo], a[lo] = 0, 0
        hi = i0 >= sn - 1
        i0[hi], a[hi] = sn - 1, 0
        return i0, np.minimum(i0 + 1, sn - 1), a

    x0, x1, ax = taps(dw, sw)
    y0, y1, ay = taps(dh, sh)
    rows = src[:, x0] * (1 - ax)[None, :] + src[:, x1] * ax[None, :]
    return (rows[y0] * (1 - ay)[:, None] + rows[y1] * ay[:, None]).astype(np.float32)


def generate_cam_label(cam_refined: Sequence[Tensor], size_wh: Tuple[int, int]) -> Tensor:
    """clip/clip_tool.py:202-216 — per class min-max then cv2.resize to (w,h) -> [K,Hc,Wc] float32."""
    out = [cv2_resize_linear(scale_cam_image(c.cpu().numpy().astype(np.float32)), size_wh) for c in cam_refined]
    return torch.from_numpy(np.stack(out, 0))


def cams_with_bg(cams: Tensor, cam_bg_thres: float = 1.0) -> Tensor:
    """model_attn_aff_voc.py:162-163 — bg = (1-max_k)^thres, concat in front."""
    bg = torch.pow(1 - torch.max(cams, dim=0, keepdim=True)[0], cam_bg_thres)
    return torch.cat([bg, cams], dim=0)


# --------------------------------------------------------------------------------------------------
# A.6 PAR
# --------------------------------------------------------------------------------------------------
def par_offsets(dilations=PAR_DILATIONS):
    """WeCLIP_model/PAR.py:10-24,39-49 — (dy,dx) per neighbour, dilation-major, 3x3 raster order w/o centre."""
    base = [(-1, -1), (-1, 0), (-1, 1), (0, -1), (0, 1), (1, -1), (1, 0), (1, 1)]
    return [(dy * d, dx * d) for d in dilations for dy, dx in base]


def par_neighbors(x: Tensor, dilations=PAR_DILATIONS) -> Tensor:
    """WeCLIP_model/PAR.py:39-49 — replicate-padded dilated neighbours [B,C,H,W] -> [B,C,48,H,W] (by index clamp)."""
    B, C, H, W = x.shape
    ys = torch.arange(H, device=x.device).view(H, 1)
    xs = torch.arange(W, device=x.device).view(1, W)
    out = []
    for dy, dx in par_offsets(dilations):
        yy = (ys + dy).clamp(0, H - 1).expand(H, W)
        xx = (xs + dx).clamp(0, W - 1).expand(H, W)
        out.append(x[:, :, yy, xx])
    return torch.stack(out, dim=2)


def par_pos_aff(dilations=PAR_DILATIONS, w1: float = 0.3) -> Tensor:
    """WeCLIP_model/PAR.py:51-62,78,83 — the data-independent positional term, softmaxed: [48]."""
    ker = torch.ones(8)
    for i in (0, 2, 5, 7):
        ker[i] = np.sqrt(2)
    pos = torch.cat([ker * d for d in dilations])
    std = torch.std(pos)
    return F.softmax(-(pos / (std + 1e-8) / w1) ** 2, dim=0)


def par_affinity(imgs: Tensor, dilations=PAR_DILATIONS, w1: float = 0.3, w2: float = 0.01) -> Tensor:
    """WeCLIP_model/PAR.py:70-86 — [B,3,H,W] -> aff [B,48,H,W] (rows sum to 1+w2)."""
    nb = par_neighbors(imgs, dilations)
    ab = torch.abs(nb - imgs.unsqueeze(2))
    sd = torch.std(nb, dim=2, keepdim=True)
    aff = -(ab / (sd + 1e-8) / w1) ** 2
    aff = aff.mean(dim=1)
    return F.softmax(aff, dim=1) + w2 * par_pos_aff(dilations, w1).view(1, -1, 1, 1).to(imgs.device)


def par_forward(imgs: Tensor, masks: Tensor, dilations=PAR_DILATIONS, num_iter: int = PAR_NUM_ITER) -> Tensor:
    """WeCLIP_model/PAR.py:64-92."""
    imgs = F.interpolate(imgs, size=masks.shape[-2:], mode="bilinear", align_corners=True)
    aff = par_affinity(imgs, dilations).unsqueeze(1)
    for _ in range(num_iter):
        masks = (par_neighbors(masks, dilations) * aff).sum(2)
    return masks


def refine_labels(img: Tensor, cams: Tensor, valid_key: Tensor) -> Tensor:
    """model_attn_aff_voc.py:49-57 — PAR, argmax over K+1, valid_key LUT -> [Hc,Wc] int64."""
    ref = par_forward(img.unsqueeze(0).float(), cams.unsqueeze(0).float())
    return valid_key[ref.argmax(dim=1)].squeeze(0)


# --------------------------------------------------------------------------------------------------
# a1/a7 orchestration
# --------------------------------------------------------------------------------------------------
def label_ids_from_map(lab: np.ndarray) -> List[int]:
    """clip/clip_tool.py:111-118 — np.unique, uint8 '-1' (0->255 bg, 255->254 ignore), both dropped."""
    ids = (np.unique(lab) - 1).tolist()
    return [int(i) for i in ids if i not in (254, 255)]


def weclip_forward(img: Tensor, label_maps: Sequence[np.ndarray], sd: Dict[str, Tensor], fg_text: Tensor,
                   bg_text: Tensor, mode: str = "val", iter_num: int = 1, dataset: str = "voc",
                   mimic_fp16: bool = True, return_intermediates: bool = False, box_masks=None, rfm_masks=None):
    """WeCLIP_model/model_attn_aff_voc.py:107-175 (COCO twin model_attn_aff_coco.py:100-170) with
    clip/clip_tool.py:106-197 inlined.  ``label_maps[i]`` is the decoded label PNG of image i; like the reference's
    ``for i, img_name in enumerate(img_names)`` (:139) the CAM / PAR path runs for the first len(label_maps) images only
    (the msc+flip driver passes ONE name for its [image, flipped image] batch).
    Returns (seg [N,C,h,w], cam_labels [N,Hc,Wc] int64 | None, attn_pred [N,P,P])."""
    N, _, H, W = img.shape
    h, w = H // 16, W // 16
    taps, attns = encode_image(img, sd, mimic_fp16)
    fts = decoder_fts_fuse(taps, sd, h, w)
    seg = decoder(fts, sd, mimic_fp16)
    attn_pred = attn_pred_from_fts(fts)
    inter = {"taps": taps, "attns": attns, "fts": fts}
    if dataset == "coco" and mode == "val":           # model_attn_aff_coco.py:131-132
        return (seg, None, attn_pred, inter) if return_intermediates else (seg, None, attn_pred)
    window, thr, switch = (6, 0.4, 15000) if dataset == "voc" else (10, 0.7, 40000)
    require_seg_trans = iter_num > switch or mode == "val"
    attn_stack = torch.stack(attns, dim=0).permute(1, 0, 2, 3)   # [N,11,L,L]
    labels_out = []
    inter.update(cams=[], cams_refined=[], cams_up=[], trans_mat=[], keys=[], par_in=[])
    for i in range(len(label_maps)):
        ids = label_ids_from_map(label_maps[i])
        ori_h, ori_w = label_maps[i].shape[:2]
        text = torch.cat([fg_text[ids], bg_text], dim=0).to(img.device)
        x11 = taps[-1][:, i:i + 1, :].float()
        cams, attn_last, _ = grad_cam(x11, text, list(range(len(ids))), sd, h, w, mimic_fp16)
        stack = torch.cat([attn_stack[i], attn_last], dim=0)
        rinfo = {} if return_intermediates else None
        aff = rfm_attention(stack, attn_pred[i:i + 1], require_seg_trans, window,
                            None if rfm_masks is None else rfm_masks[i], rinfo)
        if return_intermediates:
            inter.setdefault("rfm", []).append(rinfo)
        tm = compute_trans_mat(aff)
        refined = refine_cams(cams, tm, thr, None if box_masks is None else box_masks[i])
        size_wh = (W, H) if mode == "train" else (ori_w, ori_h)
        up = generate_cam_label(refined, size_wh).to(img.device)              # .cuda(), model_attn_aff_voc.py:160
        full = cams_with_bg(up)
        valid_key = torch.from_numpy(np.pad(np.asarray(ids) + 1, (1, 0), mode="constant")).to(img.device)
        labels_out.append(refine_labels(img[i], full, valid_key))
        inter["cams"].append(cams); inter["cams_refined"].append(torch.stack(refined)); inter["cams_up"].append(up)
        inter["trans_mat"].append(tm); inter["keys"].append(ids); inter["par_in"].append(full)
        if return_intermediates:
            inter.setdefault("box_masks", []).append(np.stack([box_mask(c, thr) for c in cams]))
    out = (seg, torch.stack(labels_out, dim=0), attn_pred)
    return out + (inter,) if return_intermediates else out


# --------------------------------------------------------------------------------------------------
# training-time label prep and losses on the path's outputs
# --------------------------------------------------------------------------------------------------
def mask_by_radius(h: int = 20, w: int = 20, radius: int = 8) -> np.ndarray:
    """scripts/dist_clip_voc.py:116-133 (closed form: the reference's symmetric double write marks exactly the
    pairs whose row and column offsets are both <= radius)."""
    ys, xs = np.divmod(np.arange(h * w), w)
    return ((np.abs(ys[:, None] - ys[None, :]) <= radius) & (np.abs(xs[:, None] - xs[None, :]) <= radius)).astype(
        np.float64)


def affinity_label(cam_label: Tensor, mask=None, ignore_index: int = 255) -> Tensor:
    """utils/camutils.py:226-247: nearest down-sample by 16, pairwise label equality, ignore outside the radius mask
    and on rows / columns whose own label is ignore_index."""
    b, h, w = cam_label.shape
    small = F.interpolate(cam_label.unsqueeze(1).float(), size=[h // 16, w // 16], mode="nearest").reshape(b, -1)
    aff = (small[:, None, :] == small[:, :, None]).long()
    if mask is not None:
        aff[:, torch.as_tensor(np.asarray(mask) == 0)] = ignore_index
    ign = small == ignore_index
    aff[ign[:, None, :].expand_as(aff)] = ignore_index
    aff[ign[:, :, None].expand_as(aff)] = ignore_index
    return aff


def aff_loss(inputs: Tensor, targets: Tensor):
    """utils/losses.py:11-22."""
    pos, neg = targets == 1, targets == 0
    pos_count, neg_count = pos.sum() + 1, neg.sum() + 1
    pos_loss = (1 - inputs)[pos].double().sum() / pos_count
    neg_loss = inputs[neg].double().sum() / neg_count
    return (0.5 * pos_loss + 0.5 * neg_loss).float(), pos_count, neg_count


def seg_loss(pred: Tensor, label: Tensor, ignore_index: int = 255) -> Tensor:
    """utils/losses.py:24-33: mean NLL over label==0 pixels and over foreground pixels, averaged."""
    logp = torch.log_softmax(pred.double(), dim=1)
    lab = label.long()
    valid = lab != ignore_index
    nll = -logp.gather(1, lab.clamp(0, pred.shape[1] - 1).unsqueeze(1)).squeeze(1)
    bg, fg = valid & (lab == 0), valid & (lab != 0)
    return ((nll[bg].sum() / bg.sum() + nll[fg].sum() / fg.sum()) * 0.5).float()


# --------------------------------------------------------------------------------------------------
# a15/a16 msc+flip driver and scores
# --------------------------------------------------------------------------------------------------
def fast_hist(gt: np.ndarray, pred: np.ndarray, num_classes: int) -> np.ndarray:
    """utils/evaluate.py:9-15."""
    gt, pred = gt.flatten(), pred.flatten()
    m = (gt >= 0) & (gt < num_classes)
    return np.bincount(num_classes * gt[m].astype(int) + pred[m], minlength=num_classes ** 2).reshape(
        num_classes, num_classes)


def scores(hist: np.ndarray) -> dict:
    """utils/evaluate.py:18-36."""
    with np.errstate(divide="ignore", invalid="ignore"):
        acc = np.diag(hist).sum() / hist.sum()
        acc_cls = np.nanmean(np.diag(hist) / hist.sum(axis=1))
        iu = np.diag(hist) / (hist.sum(axis=1) + hist.sum(axis=0) - np.diag(hist))
        valid = hist.sum(axis=1) > 0
        return {"pAcc": acc, "mAcc": acc_cls, "miou": np.nanmean(iu[valid]), "iou": dict(enumerate(iu))}


def msc_flip_image(img: Tensor, label_map: np.ndarray, sd, fg_text, bg_text, scales=(1.0, 0.75, 1.25),
                   resize_long: int = 512, dataset: str = "voc", mimic_fp16: bool = True):
    """test_msc_flip_voc.py:50-96 for one dataset image.  img [1,3,H0,W0]; label_map [Hl,Wl] uint8.
    ``names = name+name`` (:52): every model call gets TWO names for its [x, flip(x)] batch, so the CAM / affinity / PAR
    path runs for both images at every scale (2 x n_scales passes); only ``cam[0]`` of the scale-1.0 call is kept (:71).
    Returns (seg_pred [Hl,Wl], msc_pred [Hl,Wl], cam_label [Hl,Wl] | None, msc_segs [1,C,h,w])."""
    _, _, h0, w0 = img.shape
    ratio = resize_long / max(h0, w0)
    inputs = F.interpolate(img, size=(int(h0 * ratio), int(w0 * ratio)), mode="bilinear", align_corners=False)
    cat = torch.cat([inputs, inputs.flip(-1)], dim=0)
    names = [label_map, label_map]                                                  # names = name+name (:52)
    segs_cat, cam, _ = weclip_forward(cat, names, sd, fg_text, bg_text, "val", 1, dataset, mimic_fp16)
    segs = segs_cat[0:1]
    segs_list = [(segs_cat[0] + segs_cat[1].flip(-1)) / 2]
    h, w = segs_cat.shape[-2:]
    for s in scales:
        if s != 1.0:
            _in = F.interpolate(inputs, scale_factor=s, mode="bilinear", align_corners=False)
            cat = torch.cat([_in, _in.flip(-1)], dim=0)
            sc, _, _ = weclip_forward(cat, names, sd, fg_text, bg_text, "val", 1, dataset, mimic_fp16)     # cam_cat unused (:84)
            sc = F.interpolate(sc, size=(h, w), mode="bilinear", align_corners=False)
            segs_list.append((sc[0] + sc[1].flip(-1)) / 2)
    msc = torch.mean(torch.stack(segs_list, dim=0), dim=0).unsqueeze(0)
    size = label_map.shape[:2]
    seg_pred = torch.argmax(F.interpolate(segs, size=size, mode="bilinear", align_corners=False), dim=1)[0]
    msc_pred = torch.argmax(F.interpolate(msc, size=size, mode="bilinear", align_corners=False), dim=1)[0]
    return seg_pred, msc_pred, (cam[0] if cam is not None else None), msc

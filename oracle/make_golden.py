"""Generate tests/golden/*.npz by running the REAL reference from /root/reference (build container only).

    python oracle/make_golden.py

TEST INFRASTRUCTURE.  Inputs come from ``weclip_b200.synth`` seeds (regenerated bit-identically by the tests);
only the reference's *outputs* (plus the small micro-case inputs) are stored.  Library versions that the
outputs depend on are recorded in every file (torch / cv2 / numpy are the arithmetic).
"""
from __future__ import annotations

import os
import sys
import tempfile

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)

from weclip_b200 import synth  # noqa: E402
from oracle import ref_harness  # noqa: E402

OUT = os.path.join(REPO, "tests", "golden")
WECLIP_CASE = dict(weights_seed=0, text_seed=7, label_seed=3, image_seed=1, H=96, W=112, label_hw=(70, 90),
                   names=("syn0", "syn1"))


def _versions():
    import cv2
    return dict(torch=torch.__version__, numpy=np.__version__, cv2=cv2.__version__)


def _save(name, **arrs):
    arrs["versions"] = np.array(repr(_versions()))
    np.savez_compressed(os.path.join(OUT, name), **arrs)
    print("wrote", name, {k: getattr(v, "shape", None) for k, v in arrs.items()})


def golden_par(ref):
    PAR = ref["PAR"]
    par = PAR(num_iter=20, dilations=[1, 2, 4, 8, 12, 24])
    g = torch.Generator().manual_seed(11)
    cases = {"same": (40, 56, 40, 56, 3), "resize": (33, 47, 52, 38, 2), "tiny": (5, 7, 5, 7, 4)}
    for tag, (hi, wi, h, w, kc) in cases.items():
        imgs = synth.make_images(21 + kc, 1, hi, wi)
        masks = torch.rand(1, kc, h, w, generator=g)
        with torch.no_grad():
            out = par(imgs, masks)
            # affinity tensor, recomputed exactly as PAR.forward does (PAR.py:67-86)
            im = torch.nn.functional.interpolate(imgs, size=(h, w), mode="bilinear", align_corners=True)
            nb = par.get_dilated_neighbors(im)
            ab = torch.abs(nb - im.unsqueeze(2).repeat(1, 1, nb.shape[2], 1, 1))
            sd = torch.std(nb, dim=2, keepdim=True)
            aff = (-(ab / (sd + 1e-8) / 0.3) ** 2).mean(dim=1, keepdim=True)
            pos = par.pos.repeat(1, 1, 1, h, w)
            pos_aff = -(pos / (torch.std(pos, dim=2, keepdim=True) + 1e-8) / 0.3) ** 2
            aff = torch.softmax(aff, dim=2) + 0.01 * torch.softmax(pos_aff, dim=2)
        _save(f"par_{tag}.npz", imgs=imgs.numpy(), masks=masks.numpy(), out=out.numpy(), aff=aff[:, 0].numpy())


def golden_small_ops(ref):
    g = torch.Generator().manual_seed(5)
    a = torch.rand(30, 30, generator=g) + 0.01
    _save("transmat.npz", a=a.numpy(), out=ref["compute_trans_mat"](a).numpy())

    rng = np.random.default_rng(9)
    cams, masks, thrs = [], [], []
    for i in range(24):
        h, w = (12, 14) if i % 2 else (20, 20)
        lo = rng.random((h // 2 + 1, w // 2 + 1)).astype(np.float32)
        cam = np.kron(lo, np.ones((2, 2), np.float32))[:h, :w] * rng.random((h, w)).astype(np.float32) ** 0.3
        if i == 0:
            cam[:] = 0
        if i == 1:
            cam[:] = 1
        if i == 2:
            cam[:] = 0; cam[-1, :] = 1        # component confined to the last row
        thr = 0.4 if i % 3 else 0.7
        box, cnt = ref["scoremap2bbox"](scoremap=cam, threshold=thr, multi_contour_eval=True)
        m = np.zeros((h, w), np.float32)
        for j in range(cnt):
            x0, y0, x1, y1 = box[j]
            m[y0:y1, x0:x1] = 1
        pad = np.zeros((20, 20), np.float32); pad[:h, :w] = cam; cams.append(pad)
        pm = np.zeros((20, 20), np.float32); pm[:h, :w] = m; masks.append(pm)
        thrs.append((h, w, thr))
    _save("bbox.npz", cams=np.stack(cams), masks=np.stack(masks), meta=np.array(thrs, dtype=np.float64))

    refined = [torch.rand(6, 7, generator=g) * 3 - 1 for _ in range(3)]
    outs = {}
    for tag, (w, h) in {"up": (112, 96), "odd": (90, 70), "down": (5, 4)}.items():
        outs[tag] = ref["generate_cam_label"](refined, [1, 2, 3], w, h)["refined_cam"].numpy()
    _save("camlabel.npz", refined=torch.stack(refined).numpy(), **outs)

    gts = [rng.integers(0, 21, (17, 19)).astype(np.int16) for _ in range(3)]
    gts[0][:3] = 255
    preds = [rng.integers(0, 21, (17, 19)).astype(np.int16) for _ in range(3)]
    hist, sc = ref["evaluate"].scores(gts, preds, np.zeros((21, 21)))
    _save("scores.npz", gts=np.stack(gts), preds=np.stack(preds), hist=hist,
          summary=np.array([sc["pAcc"], sc["mAcc"], sc["miou"]]), iou=np.array([sc["iou"][i] for i in range(21)]))


def golden_weclip(dataset, num_classes, n_fg, n_bg, subdir):
    c = WECLIP_CASE
    sd = synth.make_state_dict(c["weights_seed"], num_classes)
    fg, bg = synth.make_text_features(c["text_seed"], n_fg, n_bg)
    root = tempfile.mkdtemp()
    rng = np.random.default_rng(c["label_seed"])
    labs = []
    for n in c["names"]:
        ids = synth.draw_label_set(rng, n_fg)
        lab = synth.make_label_map(rng, *c["label_hw"], ids)
        synth.write_label_png(root, subdir, n, lab)
        labs.append(lab)
    img = synth.make_images(c["image_seed"], 2, c["H"], c["W"])
    model = ref_harness.build_reference_model(sd, fg, bg, root, dataset, num_classes)
    switch = 15000 if dataset == "voc" else 40000
    runs = {"val": ("val", 0), "train": ("train", 0), "train_rfm": ("train", switch + 5)}
    for tag, (mode, it) in runs.items():
        model.iter_num = it
        seg, cam, ap = model(img, list(c["names"]), mode=mode)
        arrs = dict(seg=seg.detach().numpy(), attn_pred=ap.detach().numpy(), labels=np.stack(labs))
        if cam is not None:
            arrs["cam"] = cam.numpy().astype(np.int16)
        _save(f"weclip_{dataset}_{tag}.npz", **arrs)
    if dataset == "voc":
        # encoder taps / attention and raw Grad-CAMs for image 0 (the pieces the kernels are diffed against)
        from clip.clip_tool import generate_clip_fts, ClipOutputTarget
        from oracle.weclip_oracle import label_ids_from_map
        with torch.no_grad():
            fts_all, attn_list = generate_clip_fts(img, model.encoder, require_all_fts=True)
        ids = label_ids_from_map(labs[0])
        text = torch.cat([fg[ids], bg], dim=0)
        cams = []
        for k in range(len(ids)):
            cam_k, probs, attn_last = model.grad_cam(
                input_tensor=[fts_all[-1][:, 0:1, :], text, c["H"], c["W"]], targets=[ClipOutputTarget(k)],
                target_size=None)
            cams.append(cam_k[0])
        _save("weclip_voc_parts.npz", tap0=fts_all[0].numpy(), tap10=fts_all[-1].numpy(),
              attn0=attn_list[0].numpy(), attn10=attn_list[-1].numpy(), cams0=np.stack(cams),
              probs0=probs.detach().numpy(), attn_last0=attn_last.detach().numpy())


MSCFLIP_CASE = dict(weights_seed=0, text_seed=7, label_seed=21, image_seed=31, resize_long=160, scales=(1.0, 0.75, 1.25),
                    sizes=((93, 130), (120, 88)), names=("m0", "m1"))


def mscflip_case(n_fg=20):
    """Seeded inputs of the msc+flip golden case (regenerated bit-identically by the tests): per dataset image
    (name, image [3,H0,W0] fp32, label map [H0,W0] uint8)."""
    c = MSCFLIP_CASE
    rng = np.random.default_rng(c["label_seed"])
    items = []
    for i, (h0, w0) in enumerate(c["sizes"]):
        lab = synth.make_label_map(rng, h0, w0, synth.draw_label_set(rng, n_fg))
        img = synth.make_images(c["image_seed"] + i, 1, h0, w0)[0]
        items.append((c["names"][i], img, lab))
    return items


def golden_mscflip():
    """The REAL validate() loop of test_msc_flip_voc.py:30-113 (names = name+name, DataLoader batch_size=1, .npy logit
    dump) on two small dataset images of different sizes, followed by main()'s evaluate.scores calls (:202-204)."""
    c = MSCFLIP_CASE
    sd = synth.make_state_dict(c["weights_seed"], 21)
    fg, bg = synth.make_text_features(c["text_seed"], 20, 25)
    root = tempfile.mkdtemp()
    items = mscflip_case()
    for name, _, lab in items:
        synth.write_label_png(root, "SegmentationClassAug", name, lab)
    model = ref_harness.build_reference_model(sd, fg, bg, root, "voc", 21)
    validate, args = ref_harness.reference_validate("voc")
    args.resize_long = c["resize_long"]
    dataset = [(name, img.numpy(), lab.astype(np.int64), np.zeros(20, np.float32)) for name, img, lab in items]
    calls = []
    fwd = model.forward
    model.forward = lambda img, names, mode="train": (calls.append(len(names)), fwd(img, names, mode=mode))[1]
    gts, preds, msc_preds, cams, ph, mh, ch = validate(model=model, dataset=dataset, test_scales=list(c["scales"]))
    ev = ref_harness.reference_modules()["evaluate"]
    ph, _ = ev.scores(gts, preds, ph)
    mh, _ = ev.scores(gts, msc_preds, mh)
    ch, _ = ev.scores(gts, cams, ch)
    out = dict(hist=np.stack([ph, mh, ch]).astype(np.int64), names_per_call=np.array(calls))
    for i, (name, _, _) in enumerate(items):
        d = np.load(os.path.join(args.work_dir, "logit", name + ".npy"), allow_pickle=True).item()   # :111
        out[f"segs{i}"], out[f"msc_segs{i}"] = d["segs"], d["msc_segs"]
        out[f"seg_pred{i}"], out[f"msc_pred{i}"], out[f"cam{i}"] = preds[i], msc_preds[i], cams[i]
    _save("mscflip_voc.npz", **out)


def training_case(seed=17, B=3, H=96, W=112, C=21):
    """Seeded inputs of the label-prep / loss golden case (regenerated bit-identically by the tests)."""
    g = torch.Generator().manual_seed(seed)
    cam = torch.randint(0, 4, (B, H // 8, W // 8), generator=g).repeat_interleave(8, 1).repeat_interleave(8, 2)
    cam = cam * 5                                             # classes {0, 5, 10, 15}
    cam[:, :9, :] = 255                                       # ignored stripe (covers whole 16-px patch rows)
    cam[1, 40:70, 50:90] = 255
    P = (H // 16) * (W // 16)
    attn = torch.rand(B, P, P, generator=g)
    pred = torch.randn(B, C, H, W, generator=g) * 3
    return cam, attn, pred


def golden_training():
    fns = ref_harness.reference_training_fns()
    cam, attn, pred = training_case()
    B, H, W = cam.shape
    h, w = H // 16, W // 16
    out = {}
    for r in (2, 8):
        mask = fns["get_mask_by_radius"](h=h, w=w, radius=r)
        aff = fns["cams_to_affinity_label"](cam.clone(), mask=mask, ignore_index=255)
        loss, pc, nc = fns["get_aff_loss"](attn, aff)
        out[f"mask_r{r}"] = mask.astype(np.uint8)
        out[f"aff_r{r}"] = aff.numpy().astype(np.int16)
        out[f"affloss_r{r}"] = np.array([float(loss), float(pc), float(nc)])
    aff = fns["cams_to_affinity_label"](cam.clone(), mask=None, ignore_index=255)
    out["aff_nomask"] = aff.numpy().astype(np.int16)
    x = attn.clone().requires_grad_(True)
    loss, _, _ = fns["get_aff_loss"](x, torch.from_numpy(out["aff_r2"].astype(np.int64)))
    loss.backward()
    out["affgrad_r2"] = x.grad.numpy()
    p = pred.clone().requires_grad_(True)
    sl = fns["get_seg_loss"](p, cam.long(), ignore_index=255)
    sl.backward()
    out["segloss"] = np.array([float(sl)])
    out["seggrad_sample"] = p.grad[:, :, ::7, ::5].contiguous().numpy()
    _save("training.npz", **out)


def main():
    assert ref_harness.available(), "/root/reference is required (build container only)"
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)
    ref = ref_harness.reference_modules()
    golden_par(ref)
    golden_small_ops(ref)
    golden_training()
    golden_mscflip()
    golden_weclip("voc", 21, 20, 25, "SegmentationClassAug")
    golden_weclip("coco", 81, 80, 23, os.path.join("SegmentationClass", "train"))


if __name__ == "__main__":
    main()

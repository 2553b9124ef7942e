"""Developer script (GPU box): diagnose full-forward disagreement for a (dataset, H, W, N, mode, it, seeds) case."""
import sys, os, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from weclip_b200 import synth
from weclip_b200.model import WeCLIP
from oracle import weclip_oracle as O
from tools.dev_check import stat


def run(dataset, H, W, N, mode, it, wseed, tseed, iseed, lseed, lab_hw):
    ncls, n_fg, n_bg, sub = (21, 20, 25, "SegmentationClassAug") if dataset == "voc" else \
        (81, 80, 23, os.path.join("SegmentationClass", "train"))
    sd = synth.make_state_dict(wseed, ncls)
    fg, bg = synth.make_text_features(tseed, n_fg, n_bg)
    img = synth.make_images(iseed, N, H, W)
    root = tempfile.mkdtemp()
    rng = np.random.default_rng(lseed)
    labs, names = [], []
    for i in range(N):
        lab = synth.make_label_map(rng, lab_hw[0], lab_hw[1], synth.draw_label_set(rng, n_fg))
        synth.write_label_png(root, sub, f"d{i}", lab)
        labs.append(lab); names.append(f"d{i}")
    model = WeCLIP(num_classes=ncls, dataset_root_path=root, device="cuda", dataset=dataset, state_dict=sd,
                   fg_text_features=fg, bg_text_features=bg)
    model.iter_num = it
    seg, cam, ap = model(img.cuda(), names, mode=mode)
    out = O.weclip_forward(img, labs, sd, fg, bg, mode=mode, iter_num=it + 1, dataset=dataset, return_intermediates=True)
    seg_o, cam_o, ap_o, inter = out
    print(f"== {dataset} {H}x{W} N={N} {mode} it={it} keys={inter.get('keys')}")
    stat("seg", seg, seg_o); stat("attn_pred", ap, ap_o)
    print("seg argmax agree", (seg.cpu().argmax(1) == seg_o.argmax(1)).float().mean().item())
    if cam is None:
        return
    print("cam agreement", [(cam[i].cpu() == cam_o[i]).float().mean().item() for i in range(N)])
    eng = model.engine
    h, w = H // 16, W // 16
    window, thr, switch = (6, 0.4, 15000) if dataset == "voc" else (10, 0.7, 40000)
    rfm = (it + 1) > switch or mode == "val"
    nl = window if rfm else 8
    ids_all = inter["keys"]
    k_off, t_off, rows = [0], [0], []
    for ids in ids_all:
        k_off.append(k_off[-1] + len(ids)); t_off.append(t_off[-1] + len(ids) + n_bg); rows += [fg[ids], bg]
    text = torch.cat(rows).cuda()
    taps, x_last, attn = eng.encoder(img.cuda(), n_attn=nl - 1, attn_slots=nl)
    cams, _, _ = eng.gradcam(x_last, N, h, w, text, t_off, k_off, attn_last_out=attn[nl - 1])
    cams_o = torch.from_numpy(np.concatenate(inter["cams"], 0))
    stat("cams", cams, cams_o)
    seg2, ap2, _ = eng.decoder(taps, N, h, w)
    ref, masks = eng.affinity_refine(attn, nl, ap2, N, h, w, rfm, cams, k_off, thr, want_masks=True)
    m_o = torch.from_numpy(np.stack([O.box_mask(c, thr) for c in cams_o.numpy()]))
    neq = (masks.cpu() != m_o).flatten(1).sum(1)
    print("box-mask cells differing per instance:", neq.tolist())
    ref_o = torch.cat(inter["cams_refined"], 0)
    stat("refined", ref, ref_o)
    # same, but with the oracle's raw cams (isolates the box discontinuity)
    ref2, _ = eng.affinity_refine(attn, nl, ap2, N, h, w, rfm, cams_o.cuda().contiguous(), k_off, thr)
    stat("refined (oracle cams)", ref2, ref_o)


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "coco"):
        for mode, it in (("val", 0), ("train", 0), ("train", 40005)):
            run("coco", 96, 112, 2, mode, it, 0, 7, 1, 3, (70, 90))
    if which in ("all", "voc320"):
        run("voc", 320, 320, 1, "val", 0, 5, 11, 640, 320 * 7 + 1, (329, 314))

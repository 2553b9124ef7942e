"""Diagnose full-size parity: per-image pseudo-label agreement vs the oracle, box-mask equality, RFM layer-selection margins
and Grad-CAM error (run on the GPU box)."""
import os
import sys
import tempfile

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import weclip_oracle as O  # noqa: E402
from weclip_b200 import synth  # noqa: E402
from weclip_b200.model import WeCLIP  # noqa: E402


def run(N, H, W, seed, it, mode, pair=False):
    torch.set_num_threads(os.cpu_count() or 1)
    sd = synth.make_state_dict(0, 21)
    fg, bg = synth.make_text_features(7, 20, 25)
    root = tempfile.mkdtemp()
    rng = np.random.default_rng(seed)
    names, labs = [], []
    for i in range(N):
        lab = synth.make_label_map(rng, H, W, synth.draw_label_set(rng, 20))
        synth.write_label_png(root, "SegmentationClassAug", f"f{seed}_{i}", lab)
        names.append(f"f{seed}_{i}"); labs.append(lab)
    imgs = synth.make_images(500 + seed, N, H, W)
    if pair:
        imgs = torch.cat([imgs, imgs.flip(-1)]); names = names + names; labs = labs + labs
    model = WeCLIP(num_classes=21, dataset_root_path=root, device="cuda", state_dict=sd, fg_text_features=fg,
                   bg_text_features=bg)
    model.keep_intermediates = True
    model.iter_num = it
    seg, cam, ap = model(imgs.cuda(), names, mode=mode)
    seg_o, cam_o, ap_o, inter = O.weclip_forward(imgs, labs, sd, fg, bg, mode=mode, iter_num=it + 1,
                                                 return_intermediates=True)
    last = model.last
    k_off = last["k_off"]
    ours_m = last["box_masks"].cpu().numpy()
    ours_c = last["cams"].cpu().numpy()
    ours_r = last["refined"].cpu().numpy()
    window, switch = 6, 15000
    rfm = it + 1 > switch or mode == "val"
    attn_stack = torch.stack(inter["attns"], 0).permute(1, 0, 2, 3)
    print(f"== N={len(names)} {H}x{W} it={it} mode={mode}: total agreement {(cam.cpu() == cam_o).float().mean():.4f}")
    for i in range(len(names)):
        ag = (cam[i].cpu() == cam_o[i]).float().mean().item()
        m_o = inter["box_masks"][i]
        m_g = ours_m[k_off[i]:k_off[i + 1]].reshape(m_o.shape)
        c_o = inter["cams"][i]
        c_g = ours_c[k_off[i]:k_off[i + 1]].reshape(c_o.shape)
        r_o = inter["cams_refined"][i].numpy()
        r_g = ours_r[k_off[i]:k_off[i + 1]].reshape(r_o.shape)
        msg = f"img {i}: agree {ag:.4f} K={m_o.shape[0]} box_equal={np.array_equal(m_o, m_g)} cam_err={np.abs(c_o - c_g).max():.2e} " \
              f"refined_err={np.abs(r_o - r_g).max() / max(np.abs(r_o).max(), 1e-30):.2e}"
        if rfm:
            # oracle-side RFM margins (clip_tool.py:158-165): d_l vs mean
            stack = torch.cat([attn_stack[i], torch.zeros(1, *attn_stack[i].shape[1:])], 0)   # block-12 map not kept: margins of 11
            aw = attn_stack[i][:, 1:, 1:][-(window - 1):]
            d = torch.sum((ap_o[i:i + 1] - aw).flatten(1), dim=1)
            msg += f" rfm_d(first {window - 1} of window)={[round(float(x), 2) for x in d]}"
        if ag < 0.995:
            msg += "   <<<<"
        print(msg)


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "cfg2"):
        run(32, 320, 320, 48, 20000, "train")
    if which in ("all", "cfg3"):
        for (h, w) in ((512, 512), (384, 384), (640, 640)):
            run(1, h, w, 41, 0, "val", pair=True)


def gradcam_isolation(H, W, seed=41):
    """Where does the Grad-CAM error at large L come from?  (a) our encoder output vs the oracle's, (b) our Grad-CAM stage
    fed with the ORACLE's F_10 (isolates block 12 + head + analytic backward), (c) the full path."""
    torch.set_num_threads(os.cpu_count() or 1)
    sd = synth.make_state_dict(0, 21)
    fg, bg = synth.make_text_features(7, 20, 25)
    root = tempfile.mkdtemp()
    rng = np.random.default_rng(seed)
    lab = synth.make_label_map(rng, H, W, synth.draw_label_set(rng, 20))
    synth.write_label_png(root, "SegmentationClassAug", "g0", lab)
    img = synth.make_images(500 + seed, 1, H, W)
    model = WeCLIP(num_classes=21, dataset_root_path=root, device="cuda", state_dict=sd, fg_text_features=fg,
                   bg_text_features=bg)
    eng = model.engine
    h, w = H // 16, W // 16
    L = h * w + 1
    ids = O.label_ids_from_map(lab)
    text = torch.cat([fg[ids], bg], 0)
    taps_o, attns_o = O.encode_image(img, sd)
    x11_o = taps_o[-1]                                            # [L,1,768]
    cams_o, attn_last_o, probs_o = O.grad_cam(x11_o, text, list(range(len(ids))), sd, h, w)
    taps, x_last, attn = eng.encoder(img.cuda(), n_attn=0)
    e_tap = ((x_last.cpu() - x11_o[:, 0]).abs().max() / x11_o.abs().max()).item()
    t_off, k_off = [0, len(ids) + 25], [0, len(ids)]
    cams_a, al_a, probs_a = eng.gradcam(x_last, 1, h, w, text.cuda(), t_off, k_off, want_probs=True)
    cams_b, al_b, probs_b = eng.gradcam(x11_o[:, 0].contiguous().cuda(), 1, h, w, text.cuda(), t_off, k_off, want_probs=True)
    print(f"== Grad-CAM isolation {H}x{W} (L={L}, K={len(ids)}): encoder F_10 err {e_tap:.2e}")
    print(f"   full path      : cam err {np.abs(cams_a.cpu().numpy() - cams_o).max():.2e}  probs err "
          f"{(probs_a.cpu() - probs_o[0]).abs().max():.2e}  attn_last err {(al_a.cpu() - attn_last_o).abs().max():.2e}")
    print(f"   oracle F_10 in : cam err {np.abs(cams_b.cpu().numpy() - cams_o).max():.2e}  probs err "
          f"{(probs_b.cpu() - probs_o[0]).abs().max():.2e}  attn_last err {(al_b.cpu() - attn_last_o).abs().max():.2e}")
    # how sensitive is the ORACLE itself?  perturb F_10 by the size of our encoder error and re-run it
    noise = torch.randn_like(x11_o) * (x_last.cpu() - x11_o[:, 0]).std()
    cams_p, _, _ = O.grad_cam(x11_o + noise, text, list(range(len(ids))), sd, h, w)
    print(f"   oracle with F_10 + gaussian noise of our encoder error's std: cam change {np.abs(cams_p - cams_o).max():.2e}")
    cams_h, _, _ = O.grad_cam(x11_o.half().float(), text, list(range(len(ids))), sd, h, w)
    print(f"   oracle with F_10 rounded to fp16: cam change {np.abs(cams_h - cams_o).max():.2e}")


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "gc":
    for hw in ((320, 320), (512, 512), (640, 640)):
        gradcam_isolation(*hw)

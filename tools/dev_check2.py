"""Developer script (GPU box): Grad-CAM / affinity / upsample / full forward vs the CPU oracle."""
import sys, os, time, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from weclip_b200 import synth, engine
from weclip_b200.model import WeCLIP
from oracle import weclip_oracle as O
from tools.dev_check import stat

def main():
    H, W, N = (int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (96, 112, 2)
    mode = sys.argv[4] if len(sys.argv) > 4 else "val"
    it = int(sys.argv[5]) if len(sys.argv) > 5 else 1
    sd = synth.make_state_dict(0, 21)
    fg, bg = synth.make_text_features(7, 20, 25)
    img = synth.make_images(1, N, H, W)
    h, w = H // 16, W // 16
    L = h * w + 1
    root = tempfile.mkdtemp()
    rng = np.random.default_rng(3)
    names, labs = [], []
    for i in range(N):
        ids = synth.draw_label_set(rng)
        lab = synth.make_label_map(rng, H - 7, W + 5, ids)
        synth.write_label_png(root, "SegmentationClassAug", f"s{i}", lab)
        names.append(f"s{i}"); labs.append(lab)
    t0 = time.time()
    seg_o, cam_o, ap_o, inter = O.weclip_forward(img, labs, sd, fg, bg, mode=mode, iter_num=it, return_intermediates=True)
    print("oracle forward s", time.time() - t0, "keys", inter["keys"])
    eng = engine.Engine(sd, 21, "cuda")
    ids_all = inter["keys"]
    k_off = [0]; text_off = [0]; rows = []
    for ids in ids_all:
        k_off.append(k_off[-1] + len(ids)); text_off.append(text_off[-1] + len(ids) + 25); rows += [fg[ids], bg]
    text = torch.cat(rows).cuda()
    # --- Grad-CAM isolated: oracle x11
    x11 = inter["taps"][10].permute(1, 0, 2).reshape(N * L, 768).contiguous().cuda()
    cams, attn_last, probs = eng.gradcam(x11, N, h, w, text, text_off, k_off, want_probs=True)
    torch.cuda.synchronize()
    cams_o = torch.from_numpy(np.concatenate(inter["cams"], 0))
    stat("gradcam cams (oracle x11)", cams, cams_o)
    for i in range(N):
        p_o, a_o = O.forward_last_layer(inter["taps"][10][:, i:i + 1], torch.cat([fg[ids_all[i]], bg]), sd)
        stat(f"probs img{i}", probs[text_off[i]:text_off[i + 1]], p_o[0])
        stat(f"attn_last img{i}", attn_last[i], a_o[0])
    # --- affinity isolated
    window, thr, switch = 6, 0.4, 15000
    rfm = it > switch or mode == "val"
    nl = window if rfm else 8
    attn_stack = torch.stack(inter["attns"], 0)          # [11,N,L,L]
    a_last = torch.stack([O.forward_last_layer(inter["taps"][10][:, i:i + 1], torch.cat([fg[ids_all[i]], bg]), sd)[1][0]
                          for i in range(N)])
    full = torch.cat([attn_stack, a_last[None]], 0)[-nl:].contiguous().cuda()
    refined, masks = eng.affinity_refine(full, nl, ap_o.cuda().contiguous(), N, h, w, rfm, cams_o.cuda().contiguous(), k_off, thr, want_masks=True)
    ref_o = torch.cat(inter["cams_refined"], 0)
    stat("refined (oracle inputs)", refined, ref_o)
    m_o = torch.from_numpy(np.stack([O.box_mask(c, thr) for c in cams_o.numpy()]))
    print("box masks equal:", bool((masks.cpu() == m_o).all()))
    # --- upsample isolated
    Hc, Wc = (H, W) if mode == "train" else labs[0].shape
    par_in = eng.cam_upsample_bg(ref_o.cuda().contiguous(), k_off, N, h, w, Hc, Wc)
    stat("par_in (oracle refined)", par_in, torch.cat(inter["par_in"], 0))
    # --- full model
    model = WeCLIP(num_classes=21, dataset_root_path=root, device="cuda", state_dict=sd, fg_text_features=fg, bg_text_features=bg)
    model.iter_num = it - 1
    seg, cam, ap = model(img.cuda(), names, mode=mode)
    torch.cuda.synchronize()
    stat("seg", seg, seg_o); stat("attn_pred", ap, ap_o)
    print("cam label agreement:", (cam.cpu() == cam_o).float().mean().item(), "per image", [(cam[i].cpu() == cam_o[i]).float().mean().item() for i in range(N)])
    # stage-wise end-to-end
    taps, x_last, attn = eng.encoder(img.cuda(), n_attn=nl - 1, attn_slots=nl)
    cams2, _, _ = eng.gradcam(x_last, N, h, w, text, text_off, k_off, attn_last_out=attn[nl - 1])
    stat("gradcam cams (e2e)", cams2, cams_o)
    seg2, ap2, _ = eng.decoder(taps, N, h, w)
    ref2, masks2 = eng.affinity_refine(attn, nl, ap2, N, h, w, rfm, cams2, k_off, thr, want_masks=True)
    print("box masks equal (e2e):", (masks2.cpu() == m_o).float().mean().item())
    stat("refined (e2e)", ref2, ref_o)
    par2 = eng.cam_upsample_bg(ref2, k_off, N, h, w, Hc, Wc)
    stat("par_in (e2e)", par2, torch.cat(inter["par_in"], 0))
    t0 = time.time()
    for _ in range(5):
        model(img.cuda(), names, mode=mode)
    torch.cuda.synchronize()
    print("forward ms", (time.time() - t0) / 5 * 1e3)

if __name__ == "__main__":
    main()

"""Locate numerical error inside the analytic Grad-CAM backward: every intermediate buffer of wc_gradcam (test hook
wc_gradcam_debug_offsets) against a float64 restatement of the same formulas (csrc/gradcam.cu header).  GPU box only."""
import ctypes
import math
import os
import sys
import tempfile

import numpy as np
import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import weclip_oracle as O  # noqa: E402
from weclip_b200 import _lib, synth  # noqa: E402
from weclip_b200.model import WeCLIP  # noqa: E402


def emulate(x11, text, sd, L, k_t, dt=torch.float64):
    """float64 restatement of block 12 forward + the analytic backward.  x11 [L,768]."""
    p, v = "encoder.visual.transformer.resblocks.11", "encoder.visual"
    P = L - 1
    g = lambda k: sd[k].to(dt)
    Wi, bi, Wo, bo = g(p + ".attn.in_proj_weight"), g(p + ".attn.in_proj_bias"), g(p + ".attn.out_proj.weight"), g(p + ".attn.out_proj.bias")
    Wfc, bfc, Wpr, bpr = g(p + ".mlp.c_fc.weight"), g(p + ".mlp.c_fc.bias"), g(p + ".mlp.c_proj.weight"), g(p + ".mlp.c_proj.bias")
    x = x11.to(dt)
    a = F.layer_norm(x, (768,), g(p + ".ln_1.weight"), g(p + ".ln_1.bias"), 1e-5)
    qkv = a @ Wi.t() + bi
    heads = lambda t: t.reshape(L, 12, 64).permute(1, 0, 2)
    q, k, vv = (heads(t) for t in qkv.split(768, dim=1))
    q = q / 8
    Pm = torch.softmax(q @ k.transpose(1, 2), -1)
    o = Pm @ vv
    x1 = x + (o.permute(1, 0, 2).reshape(L, 768) @ Wo.t() + bo)
    ln2 = F.layer_norm(x1, (768,), g(p + ".ln_2.weight"), g(p + ".ln_2.bias"), 1e-5)
    upre = ln2 @ Wfc.t() + bfc
    x2 = x1 + ((upre * torch.sigmoid(1.702 * upre)) @ Wpr.t() + bpr)
    y = F.layer_norm(x2, (768,), g(v + ".ln_post.weight"), g(v + ".ln_post.bias"), 1e-5)
    gv_ = y[1:].mean(0) @ g(v + ".proj")
    gn = gv_.norm(); gh = gv_ / gn
    t = text.to(dt); tn = t.norm(dim=1)
    ls = g("encoder.logit_scale").exp()
    pr = torch.softmax(ls * (t @ gh) / tn, 0)
    dl = pr[k_t] * ((torch.arange(len(pr)) == k_t).to(dt) - pr)
    s = ls * ((dl / tn) @ t)
    dgh = (s - gh * (s @ gh)) / gn
    dm = (g(v + ".proj") @ dgh) / P
    gs = 2.0 ** (3 - math.ceil(math.log2(dm.abs().max().item())))

    def ln_bwd(xin, gamma, dy):
        mu = xin.mean(1, keepdim=True); xc = xin - mu
        rstd = torch.rsqrt((xc * xc).mean(1, keepdim=True) + 1e-5)
        xh = xc * rstd
        gg = gamma * dy
        return rstd * (gg - gg.mean(1, keepdim=True) - xh * (gg * xh).mean(1, keepdim=True))
    dy = (dm * gs).expand(L, 768).clone(); dy[0] = 0
    dx2 = ln_bwd(x2, g(v + ".ln_post.weight"), dy)
    sg = torch.sigmoid(1.702 * upre)
    dh = (dx2 @ Wpr) * (sg * (1 + 1.702 * upre * (1 - sg)))
    da2 = dh @ Wfc
    dx1 = dx2 + ln_bwd(x1, g(p + ".ln_2.weight"), da2)
    dO = dx1 @ Wo
    dOh = dO.reshape(L, 12, 64).permute(1, 0, 2)
    delta = (dOh * o).sum(-1)
    dP = dOh @ vv.transpose(1, 2)
    dS = Pm * (dP - delta.unsqueeze(-1))
    colsum = dS[:, 1:, :].sum(1)
    ds0 = dS[:, :, 0]
    gq = (colsum.unsqueeze(-1) * k).sum(1) / 8
    gk = (-ds0.unsqueeze(-1) * q).sum(1)
    gvv = ((1 - Pm[:, :, 0]).unsqueeze(-1) * dOh).sum(1)
    dqkv = torch.cat([gq.reshape(-1), gk.reshape(-1), gvv.reshape(-1)])
    wv = (dqkv @ Wi) / P / gs
    return dict(a32=a, x1=x1, x2=x2, dm=dm, gs=gs, dx2=dx2, dh=dh, da2=da2, dx1=dx1, dO=dO, colsum=colsum, ds0=ds0,
                p0=Pm[:, :, 0], dqkv=dqkv, wvec=wv, delta=delta, probs=pr)


def main(H, W, seed=41):
    torch.set_num_threads(os.cpu_count() or 1)
    sd = synth.make_state_dict(0, 21)
    fg, bg = synth.make_text_features(7, 20, 25)
    rng = np.random.default_rng(seed)
    lab = synth.make_label_map(rng, H, W, synth.draw_label_set(rng, 20))
    root = tempfile.mkdtemp()
    synth.write_label_png(root, "SegmentationClassAug", "g0", lab)
    img = synth.make_images(500 + seed, 1, H, W)
    model = WeCLIP(num_classes=21, dataset_root_path=root, device="cuda", state_dict=sd, fg_text_features=fg, bg_text_features=bg)
    eng = model.engine
    h, w = H // 16, W // 16
    L = h * w + 1
    ids = O.label_ids_from_map(lab)
    K = len(ids)
    text = torch.cat([fg[ids], bg], 0)
    taps_o, _ = O.encode_image(img, sd)
    x11 = taps_o[-1][:, 0].contiguous()
    c_ref, _, _ = O.grad_cam(taps_o[-1], text, list(range(K)), sd, h, w, True)
    c_f32, _, _ = O.grad_cam(taps_o[-1], text, list(range(K)), sd, h, w, False)
    print(f"== {H}x{W}: oracle reference-exact (fp16 out-proj incl. its backward) vs oracle pure fp32: {np.abs(c_ref - c_f32).max():.2e}")
    for exact in (1, 0):
        _lib.lib().wc_gradcam_set_exact(exact)
        cams, _, probs = eng.gradcam(x11.cuda(), 1, h, w, text.cuda(), [0, K + 25], [0, K], want_probs=True)
        c = cams.cpu().numpy()
        print(f"   wc_gradcam exact={exact}: cam err vs reference-exact oracle {np.abs(c - c_ref).max():.2e}, vs fp32 oracle "
              f"{np.abs(c - c_f32).max():.2e}")
    _lib.lib().wc_gradcam_set_exact(1)
    cams, _, probs = eng.gradcam(x11.cuda(), 1, h, w, text.cuda(), [0, K + 25], [0, K], want_probs=True)
    torch.cuda.synchronize()
    ws = eng._ws[("gc", torch.cuda.current_stream().cuda_stream)]
    offs = (ctypes.c_size_t * 23)()
    chunk = _lib.lib().wc_gradcam_debug_offsets(1, h, w, K, K, offs)
    names = ["a32", "x1", "x2", "lse", "dm", "dqkv", "wvec", "gscale", "o", "upre", "dx2", "da2", "colsum", "ds0", "p0",
             "dx2b", "dh", "dx1b", "dO", "dOT", "q", "k", "v"]
    off = dict(zip(names, offs))

    def buf(name, dtype, shape):
        n = int(np.prod(shape)) * (2 if dtype == torch.float16 else 4)
        return ws[off[name]:off[name] + n].view(dtype).view(*shape).float().cpu().double()
    print(f"== {H}x{W} L={L} K={K} chunk={chunk}")
    for kk in range(K):
        e = emulate(x11, text, sd, L, kk)
        gs_g = buf("gscale", torch.float32, (K,))[kk].item()
        rel = lambda got, want: ((got - want).norm() / want.norm()).item()
        R = slice(kk * L, (kk + 1) * L)
        rows = [("gscale", gs_g / e["gs"] - 1.0)]
        rows.append(("dm*gs", rel(buf("dm", torch.float32, (K, 768))[kk] * gs_g, e["dm"] * e["gs"])))
        rows.append(("dx2", rel(buf("dx2", torch.float32, (chunk * L, 768))[R], e["dx2"])))
        rows.append(("dh", rel(buf("dh", torch.float16, (chunk * L, 3072))[R], e["dh"])))
        rows.append(("da2", rel(buf("da2", torch.float32, (chunk * L, 768))[R], e["da2"])))
        rows.append(("dx1b", rel(buf("dx1b", torch.float16, (chunk * L, 768))[R], e["dx1"])))
        rows.append(("dO", rel(buf("dO", torch.float16, (chunk * L, 768))[R], e["dO"])))
        H12 = slice(kk * 12, (kk + 1) * 12)
        rows.append(("p0", rel(buf("p0", torch.float32, (chunk * 12, L))[H12], e["p0"])))
        rows.append(("ds0", rel(buf("ds0", torch.float32, (chunk * 12, L))[H12], e["ds0"])))
        cs_g, cs_e = buf("colsum", torch.float32, (chunk * 12, L))[H12], e["colsum"]
        rows.append(("colsum", rel(cs_g, cs_e)))
        rows.append(("colsum rowsum(got)/|colsum|_1", (cs_g.sum(1).abs() / cs_g.abs().sum(1)).max().item()))
        rows.append(("colsum rowsum(exact)/|colsum|_1", (cs_e.sum(1).abs() / cs_e.abs().sum(1)).max().item()))
        dq_g, dq_e = buf("dqkv", torch.float32, (K, 2304))[kk], e["dqkv"]
        for nm, sl in (("dq", slice(0, 768)), ("dk", slice(768, 1536)), ("dv", slice(1536, 2304))):
            rows.append((f"dqkv.{nm} (|exact| {dq_e[sl].norm():.3e})", rel(dq_g[sl], dq_e[sl])))
        rows.append(("wvec", rel(buf("wvec", torch.float32, (K, 768))[kk], e["wvec"])))
        print(f"-- class {kk}: " + "; ".join(f"{n} {v:.2e}" for n, v in rows))


if __name__ == "__main__":
    for hw in ((320, 320), (512, 512), (640, 640)):
        main(*hw)

"""Per-parameter gradient error of the trainable decoder vs the oracle's autograd (GPU box)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from oracle import weclip_oracle as O
import test_train_gpu as T

N, H, W = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (2, 96, 112)
torch.set_num_threads(os.cpu_count() or 1)
model, sd, img, names = T._setup(N, H, W, 11 + N)
model.iter_num = 20000
model.decoder_fts_fuse.dropout.p = 0.0
h, w = H // 16, W // 16
radius = 8 if min(h, w) > 8 else 2
params = {"decoder_fts_fuse." + k: v for k, v in model.decoder_fts_fuse.named_parameters()}
params.update({"decoder." + k: v for k, v in model.decoder.named_parameters()})
loss, segs, cam, attn_pred = T._ours_step(model, img.cuda(), names, radius)
loss.backward()
for mimic in (True, False):
    sd_o = {k: v.clone() for k, v in sd.items()}
    leaves = {k: sd_o[k].clone().requires_grad_(True) for k in params}
    sd_o.update(leaves)
    taps, _ = O.encode_image(img, sd_o)
    taps = [t.detach() for t in taps]
    fts = O.decoder_fts_fuse(taps, sd_o, h, w)
    seg = O.decoder(fts, sd_o, mimic)
    import torch.nn.functional as F
    ap = O.attn_pred_from_fts(fts)
    segs_o = F.interpolate(seg, size=cam.shape[1:], mode="bilinear", align_corners=False)
    camc = cam.cpu()
    aff_label = O.affinity_label(camc, O.mask_by_radius(h, w, radius), 255)
    pos, neg = aff_label == 1, aff_label == 0
    attn_loss = 0.5 * (1 - ap)[pos].sum() / (pos.sum() + 1) + 0.5 * ap[neg].sum() / (neg.sum() + 1)
    logp = torch.log_softmax(segs_o, dim=1)
    nll = -logp.gather(1, camc.clamp(0, 20).unsqueeze(1)).squeeze(1)
    valid = camc != 255
    bgm, fgm = valid & (camc == 0), valid & (camc != 0)
    lo = 0.5 * (nll[bgm].sum() / bgm.sum() + nll[fgm].sum() / fgm.sum()) + 0.1 * attn_loss
    lo.backward()
    print(f"== oracle mimic_fp16={mimic}: loss ours {float(loss.detach()):.6f} oracle {float(lo.detach()):.6f}")
    rows = []
    for k, p in params.items():
        g, go = p.grad.detach().cpu().reshape(-1), leaves[k].grad.reshape(-1)
        rows.append((((g - go).abs().max() / go.abs().max().clamp_min(1e-20)).item(), k, go.abs().max().item()))
    rows.sort(reverse=True)
    for e, k, m in rows[:12]:
        print(f"   {e:.3e}  {k}  (|g|max {m:.3e})")
    print("   median err", sorted(r[0] for r in rows)[len(rows) // 2])

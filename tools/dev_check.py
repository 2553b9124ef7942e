"""Developer script (GPU box): error statistics of the CUDA path vs the CPU oracle, stage by stage."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from weclip_b200 import synth, engine
from oracle import weclip_oracle as O

def stat(name, got, want):
    got, want = got.float().cpu(), want.float().cpu()
    d = (got - want).abs()
    print(f"{name:28s} max|d|={d.max().item():.3e} mean|d|={d.mean().item():.3e} max|want|={want.abs().max().item():.3e} rel={(d.norm()/want.norm()).item():.3e}", flush=True)

def main():
    H, W, N = (int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (96, 112, 2)
    sd = synth.make_state_dict(0, 21)
    img = synth.make_images(1, N, H, W)
    h, w = H // 16, W // 16
    L = h * w + 1
    eng = engine.Engine(sd, 21, "cuda")
    t0 = time.time(); taps_o, attn_o = O.encode_image(img, sd); print("oracle enc s", time.time() - t0)
    taps, x_last, attn = eng.encoder(img.cuda(), n_attn=11)
    torch.cuda.synchronize()
    for i in (0, 5, 10):
        stat(f"tap{i}", taps[i].view(N, L, 768), taps_o[i].permute(1, 0, 2))
    stat("x_last", x_last.view(N, L, 768), taps_o[10].permute(1, 0, 2))
    for i in (0, 5, 10):
        stat(f"attn{i}", attn[i], attn_o[i])
    fts_o = O.decoder_fts_fuse(taps_o, sd, h, w)
    seg_o = O.decoder(fts_o, sd)
    ap_o = O.attn_pred_from_fts(fts_o)
    seg, ap, fts = eng.decoder(taps, N, h, w, want_fts=True)
    torch.cuda.synchronize()
    stat("fts", fts.view(N, h * w, 256), fts_o.flatten(2).permute(0, 2, 1))
    stat("attn_pred", ap, ap_o)
    stat("seg", seg, seg_o)
    print("seg argmax agree", (seg.argmax(1).cpu() == seg_o.argmax(1)).float().mean().item())
    # decoder in isolation: feed the oracle's taps (bf16) so that encoder error does not enter
    taps_in = torch.stack([t.permute(1, 0, 2).reshape(N * L, 768) for t in taps_o]).cuda().half().contiguous()
    seg2, ap2, fts2 = eng.decoder(taps_in, N, h, w, want_fts=True)
    stat("fts (oracle taps)", fts2.view(N, h * w, 256), fts_o.flatten(2).permute(0, 2, 1))
    stat("attn_pred (oracle taps)", ap2, ap_o)
    stat("seg (oracle taps)", seg2, seg_o)

if __name__ == "__main__":
    main()

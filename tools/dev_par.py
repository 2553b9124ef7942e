"""Developer check (GPU box): PAR error against the oracle for fp32 / fp16 affinity storage on a few shapes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import weclip_oracle as O
from weclip_b200 import synth
from weclip_b200.par import PAR

DIL = [1, 2, 4, 8, 12, 24]
for H, W, Kc, B in [(64, 80, 3, 1), (97, 131, 5, 2), (130, 66, 4, 3), (320, 320, 3, 2), (33, 65, 2, 1)]:
    imgs = synth.make_images(H + Kc + 1, B, H, W)
    masks = torch.rand(B, Kc, H, W, generator=torch.Generator().manual_seed(H + W))
    want = O.par_forward(imgs, masks)
    for dt in (torch.float32, torch.float16):
        got = PAR(DIL, 20, affinity_dtype=dt).cuda()(imgs.cuda(), masks.cuda()).cpu()
        err = (got - want).abs()
        agree = (got.argmax(1) == want.argmax(1)).float().mean().item()
        print(f"{H}x{W} Kc={Kc} B={B} {dt}: max {err.max().item():.3e} mean {err.mean().item():.3e} agree {agree:.5f}", flush=True)

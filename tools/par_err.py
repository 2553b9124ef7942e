import sys, torch
sys.path.insert(0, "/root/repo")
from oracle import weclip_oracle as O
from weclip_b200 import synth
from weclip_b200.par import PAR
DIL = [1, 2, 4, 8, 12, 24]
for H, W, Kc, B in [(64, 80, 3, 1), (97, 131, 5, 2), (320, 320, 3, 2), (130, 66, 4, 3)]:
    imgs = synth.make_images(H + Kc + 1, B, H, W)
    masks = torch.rand(B, Kc, H, W, generator=torch.Generator().manual_seed(H + W))
    want = O.par_forward(imgs, masks)
    got = PAR(DIL, 20, affinity_dtype=torch.float16).cuda()(imgs.cuda(), masks.cuda()).cpu()
    e = (got - want).abs()
    print(H, W, Kc, B, "max %.2e mean %.2e agree %.5f" % (e.max().item(), e.mean().item(), (got.argmax(1) == want.argmax(1)).float().mean().item()))

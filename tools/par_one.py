"""One PAR forward (for ncu): python tools/par_one.py H Kc f16|f32 [B]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from weclip_b200 import synth
from weclip_b200.par import PAR
H, Kc = int(sys.argv[1]), int(sys.argv[2])
dt = torch.float16 if sys.argv[3] == "f16" else torch.float32
B = int(sys.argv[4]) if len(sys.argv) > 4 else max(1, (4 << 20) // (H * H))
imgs = synth.make_images(H, 1, H, H).cuda().expand(B, 3, H, H).contiguous()
masks = torch.rand(B * Kc, H, H, device="cuda")
par = PAR([1, 2, 4, 8, 12, 24], 20, affinity_dtype=dt).cuda()
out = par.refine_ragged(imgs, masks, [b * Kc for b in range(B + 1)])
torch.cuda.synchronize()
print(float(out.sum()))

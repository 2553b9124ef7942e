"""Profiling target (GPU box): a few fused-attention launches at one encoder shape.  usage: attn_one.py L [hd]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from weclip_b200 import ops

L = int(sys.argv[1]) if len(sys.argv) > 1 else 1025
hd = int(sys.argv[2]) if len(sys.argv) > 2 else 64
N, heads = 16, 12
g = torch.Generator(device="cuda").manual_seed(L)
q = (torch.randn(N, heads, L, hd, device="cuda", generator=g) * 0.25).half()
k = torch.randn(N, heads, L, hd, device="cuda", generator=g).half()
v = torch.randn(N, heads, L, hd, device="cuda", generator=g).half()
for _ in range(3):
    ops.attention_fwd(q, k, v, want_avg=False)
torch.cuda.synchronize()
print("done")

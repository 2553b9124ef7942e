import sys, os
sys.path.insert(0, "/root/repo/tools"); sys.path.insert(0, "/root/repo")
import torch
from weclip_b200 import ops
def run(L, N, heads, avg, iters=10):
    g = torch.Generator(device="cuda").manual_seed(L)
    q = (torch.randn(N, heads, L, 64, device="cuda", generator=g) * 0.25).half()
    k = torch.randn(N, heads, L, 64, device="cuda", generator=g).half()
    v = torch.randn(N, heads, L, 64, device="cuda", generator=g).half()
    for _ in range(2): ops.attention_fwd(q, k, v, want_avg=avg)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): ops.attention_fwd(q, k, v, want_avg=avg)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3
for L in (1601, 577):
    for heads in (3, 6, 12, 24):
        a = run(L, 16, heads, False); b = run(L, 16, heads, True)
        print(f"L={L} heads={heads}: fwd {a:.1f} us, fwd+headmean {b:.1f} us, headmean {b-a:.1f} us", flush=True)

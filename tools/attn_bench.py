"""Developer micro-bench (GPU box): fused attention forward (+ head-mean) at the encoder's shapes.
usage: attn_bench.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from weclip_b200 import _lib, ops

def run(L, N=16, heads=12, avg=False, iters=10):
    g = torch.Generator(device="cuda").manual_seed(L)
    q = (torch.randn(N, heads, L, 64, device="cuda", generator=g) * 0.25).half()
    k = torch.randn(N, heads, L, 64, device="cuda", generator=g).half()
    v = torch.randn(N, heads, L, 64, device="cuda", generator=g).half()
    for _ in range(2):
        ops.attention_fwd(q, k, v, want_avg=avg)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        ops.attention_fwd(q, k, v, want_avg=avg)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    fl = 4.0 * L * L * 64 * heads * N
    print(f"L={L} N={N} heads={heads} avg={avg}: {ms * 1e3:8.1f} us  {fl / ms / 1e9:7.1f} TF/s", flush=True)

if __name__ == "__main__":
    for L in (577, 1025, 1601):
        run(L)
        run(L, avg=True)


"""Developer micro-bench (GPU box): wc_gemm_f16 on given shapes.  usage: gemm_bench.py M,N,K,batch,out16[,act[,inplace]] [...]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from weclip_b200 import ops

def run(M, N, K, b, out16, act=0, inplace=0, iters=20):
    A = torch.randn(b, M, K, device="cuda").half()
    B = torch.randn(b, N, K, device="cuda").half()
    Np = (N + 7) // 8 * 8
    actn = ["none", "relu", "quickgelu", "sigmoid"][act]
    bias = torch.randn(N, device="cuda") if act else None
    out = torch.empty(b, M, Np, device="cuda", dtype=torch.float16 if out16 else torch.float32)[:, :, :N]
    resid = None
    if inplace:          # the out-projection's epilogue: x += A B^T + bias in place (fp32)
        bias = torch.randn(N, device="cuda")
        resid = out if inplace == 1 else torch.zeros_like(out)      # 2: residual read from a second buffer (load + add + store)
        out.zero_()
    for _ in range(3):
        ops.gemm_f16(A, B, out=out, bias=bias, resid=resid, act=actn)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        ops.gemm_f16(A, B, out=out, bias=bias, resid=resid, act=actn)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    fl = 2.0 * M * N * K * b
    by = b * (M * K * 2 + N * K * 2 + M * N * (2 if out16 else 4))
    print(f"M={M} N={N} K={K} b={b} out16={out16} act={actn} inplace={inplace} BN={os.environ.get('WC_GEMM_FORCE_BN', 'auto')}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TF/s  {by / ms / 1e6:.0f} GB/s", flush=True)

if __name__ == "__main__":
    for spec in sys.argv[1:]:
        v = [int(x) for x in spec.split(",")]
        run(*v)

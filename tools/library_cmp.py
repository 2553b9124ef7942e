"""The "library-call bar" per kernel (SURVEY.md §8d last row): our kernels next to the library calls the reference makes
for the same op on the same B200 — cuBLAS (F.linear fp16), SDPA (what a fused library attention gives; the reference itself
materialises softmax(QK^T) in fp32, clip/myAtt.py:53-64), torch LayerNorm, and the reference's own PAR op sequence
(pad + dilated conv2d + cat + mul + sum, WeCLIP_model/PAR.py:39-49,88-90).  CUDA events, median of 10 after 3 warm-ups.
    python tools/library_cmp.py > profiles/r02_library_cmp.txt
"""
import os
import statistics
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from weclip_b200 import ops, synth  # noqa: E402
from weclip_b200.par import PAR  # noqa: E402


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return statistics.median(ts)


def gemms():
    print("# GEMM  out = A @ W^T + bias (+ epilogue as in the step): ours (wc_gemm_f16, tcgen05) vs cuBLAS fp16 F.linear (+ separate")
    print("# elementwise kernels for the activation / residual, as the reference runs them)")
    print("# M, N, K, epilogue: ours us / TF/s | library us / TF/s | ratio (library time / our time)")
    for M in (25616, 16400, 9232):
        for (N, K, epi) in ((2304, 768, "bias"), (768, 768, "bias+residual(fp32)"), (3072, 768, "bias+QuickGELU"), (768, 3072, "bias+residual(fp32)")):
            A = torch.randn(M, K, device="cuda").half()
            W = (torch.randn(N, K, device="cuda") * K ** -0.5).half()
            b = torch.randn(N, device="cuda")
            x = torch.randn(M, N, device="cuda")
            if "residual" in epi:
                out = torch.empty(M, N, device="cuda")
                ours = lambda: ops.gemm_f16(A, W, bias=b, resid=x, out=out)
                lib = lambda: x + F.linear(A, W, b.half()).float()
            elif "QuickGELU" in epi:
                ours = lambda: ops.gemm_f16(A, W, bias=b, act="quickgelu", out_dtype=torch.float16)
                def lib():
                    u = F.linear(A, W, b.half())
                    return u * torch.sigmoid(1.702 * u)
            else:
                ours = lambda: ops.gemm_f16(A, W, bias=b, out_dtype=torch.float16)
                lib = lambda: F.linear(A, W, b.half())
            t_o, t_l = timed(ours), timed(lib)
            fl = 2.0 * M * N * K
            print(f"{M:6d} {N:5d} {K:5d} {epi:22s}: {1e3 * t_o:8.1f} us {fl / t_o / 1e9:7.1f} | {1e3 * t_l:8.1f} us {fl / t_l / 1e9:7.1f} | {t_l / t_o:5.2f}x")


def attention():
    print("# attention, 16 images x 12 heads x head_dim 64: ours (fused tcgen05, + row log-sum-exp) vs F.scaled_dot_product_attention")
    print("# (library fused kernel) vs the reference's materialised fp32 softmax(q k^T) v (clip/myAtt.py:53-64)")
    for L in (577, 1025, 1601):
        g = torch.Generator(device="cuda").manual_seed(L)
        q = (torch.randn(16, 12, L, 64, device="cuda", generator=g) * 0.25).half()
        k = torch.randn(16, 12, L, 64, device="cuda", generator=g).half()
        v = torch.randn(16, 12, L, 64, device="cuda", generator=g).half()
        t_o = timed(lambda: ops.attention_fwd(q, k, v, want_avg=False))
        t_s = timed(lambda: F.scaled_dot_product_attention(q, k, v, scale=1.0))
        qf, kf, vf = q.float(), k.float(), v.float()
        t_r = timed(lambda: torch.softmax(qf @ kf.transpose(-1, -2), dim=-1) @ vf, reps=5)
        t_m = timed(lambda: ops.attention_fwd(q, k, v, want_avg=True))

        def ref_maps():
            p = torch.softmax(qf @ kf.transpose(-1, -2), dim=-1)
            return p @ vf, p.sum(1) / 12
        t_rm = timed(ref_maps, reps=5)
        fl = 4.0 * L * L * 64 * 12 * 16
        print(f"L={L:5d}: ours {1e3 * t_o:7.1f} us ({fl / t_o / 1e9:6.1f} TF/s) | SDPA {1e3 * t_s:7.1f} us ({t_s / t_o:4.2f}x) | "
              f"reference-style fp32 {1e3 * t_r:8.1f} us ({t_r / t_o:5.1f}x) || with the head-mean map: ours {1e3 * t_m:7.1f} us | "
              f"reference-style {1e3 * t_rm:8.1f} us ({t_rm / t_m:5.1f}x)")


def layernorm():
    print("# LayerNorm fp32 [M,768] -> fp16: ours vs torch.nn.functional.layer_norm(...).half()")
    for M in (25616, 16400, 9232):
        x = torch.randn(M, 768, device="cuda")
        w, b = torch.randn(768, device="cuda"), torch.randn(768, device="cuda")
        from weclip_b200 import _lib
        import ctypes
        out = torch.empty(M, 768, dtype=torch.float16, device="cuda")

        def ours():
            _lib.check(_lib.lib().wc_layernorm(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(out), None, ctypes.c_long(M), 768,
                                               _lib.stream_ptr()), "ln")
        t_o, t_l = timed(ours), timed(lambda: F.layer_norm(x, (768,), w, b, 1e-5).half())
        by = M * 768 * 6.0
        print(f"M={M:6d}: ours {1e3 * t_o:6.1f} us ({by / t_o / 1e6:6.0f} GB/s) | torch {1e3 * t_l:6.1f} us ({t_l / t_o:4.2f}x)")


def par():
    print("# PAR.forward, 20 iterations, 6 dilations, Kc = 3: ours (fp16 / fp32 affinity) vs the reference's op sequence on the GPU")
    from oracle import weclip_oracle as O     # baseline leg only: the reference's PAR as plain torch ops on cuda
    for (B, H) in ((8, 512), (32, 320)):
        imgs = synth.make_images(H, 1, H, H).cuda().expand(B, 3, H, H).contiguous()
        masks = torch.rand(B, 3, H, H, device="cuda")
        t16 = timed(lambda: PAR([1, 2, 4, 8, 12, 24], 20, affinity_dtype=torch.float16).cuda()(imgs, masks), reps=5)
        t32 = timed(lambda: PAR([1, 2, 4, 8, 12, 24], 20, affinity_dtype=torch.float32).cuda()(imgs, masks), reps=5)
        t_r = timed(lambda: [O.par_forward(imgs[i:i + 1], masks[i:i + 1]) for i in range(B)], reps=3, warm=1)
        print(f"B={B:2d} {H}x{H}: ours fp16-aff {t16:7.3f} ms | ours fp32-aff {t32:7.3f} ms | reference ops (per image, like the model) "
              f"{t_r:8.2f} ms ({t_r / t16:5.1f}x / {t_r / t32:5.1f}x)")


if __name__ == "__main__":
    print(f"# {torch.cuda.get_device_name(0)}, torch {torch.__version__}")
    gemms(); attention(); layernorm(); par()

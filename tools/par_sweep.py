"""BASELINE.json configs[4]: PAR + attention-affinity refinement micro-bench sweep against the HBM roofline.

  PAR        H = W in {128 .. 1024}, Kc in {2, 3, 6}, 20 iterations, fp16 and fp32 affinity storage:
               full dilation set 1/2/4/8/12/24 (shared-memory tile kernels) and each single dilation (generic kernels);
               algorithmic bytes  B*H*W*[12 + NN s_a + n_iter (NN s_a + 8 Kc)],  NN = 8 n_dil
  refinement wc_affinity_refine on P = h*w in {64 .. 4096} patches (RFM on a window of 6 maps, K = 2 classes):
               algorithmic bytes  4 P^2 (n_win + 2) for the layer selection + 7 passes x 4 P^2 for the factored Sinkhorn / T_s
Batches are sized so that one launch covers >= 4 M pixels (every SM busy) and the working set exceeds L2 unless --l2.
    python tools/par_sweep.py > profiles/r02_par_sweep.txt
"""
import json
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from weclip_b200 import ops, synth  # noqa: E402
from weclip_b200.par import PAR  # noqa: E402

DIL = [1, 2, 4, 8, 12, 24]


def peak():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    return json.load(open(p))["hbm_gbs"] if os.path.exists(p) else 6650.0


def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return statistics.median(ts)


def par_case(H, Kc, dil, dt, pk):
    s_a = 2 if dt == torch.float16 else 4
    B = max(1, min(64, (4 << 20) // (H * H)))
    imgs = synth.make_images(H, 1, H, H).cuda().expand(B, 3, H, H).contiguous()
    masks = torch.rand(B * Kc, H, H, device="cuda")
    ch_off = [b * Kc for b in range(B + 1)]
    par = PAR(dil, 20, affinity_dtype=dt).cuda()
    ms = timed(lambda: par.refine_ragged(imgs, masks, ch_off))
    raw = ops.par_affinity(imgs, dil, dtype=dt, raw=True)
    ms_it = timed(lambda: ops.par_propagate(raw, masks, ch_off, dil, 20)) / 20
    nn_ = 8 * len(dil)
    px = B * H * H
    alg = px * (12 + nn_ * s_a + 20 * (nn_ * s_a + 8 * Kc))
    alg_it = px * (nn_ * s_a + 8 * Kc)
    return dict(H=H, B=B, Kc=Kc, dil="all" if len(dil) == 6 else str(dil[0]), aff="f16" if s_a == 2 else "f32",
                ms=round(ms, 3), gbs=round(alg / ms / 1e6, 1), frac=round(alg / ms / 1e6 / pk, 3),
                iter_us=round(1e3 * ms_it, 1), iter_gbs=round(alg_it / ms_it / 1e6, 1), iter_frac=round(alg_it / ms_it / 1e6 / pk, 3),
                aff_mb=round(px * nn_ * s_a / 1e6, 1))


def refine_case(h, pk):
    from weclip_b200 import _lib
    import ctypes
    L = _lib.lib()
    P, nl, K = h * h, 6, 2
    N = max(1, min(32, (64 << 20) // (P * P * 4 * (nl + 2))))
    g = torch.Generator(device="cuda").manual_seed(h)
    attn = torch.rand(nl, N, P + 1, P + 1, device="cuda", generator=g) + 0.01
    ap = torch.rand(N, P, P, device="cuda", generator=g)
    cams = torch.rand(N * K, P, device="cuda", generator=g)
    out = torch.empty_like(cams)
    k_off = [i * K for i in range(N + 1)]
    ws = torch.empty(L.wc_affinity_workspace_bytes(N, h, h, N * K, nl), dtype=torch.uint8, device="cuda")

    def run():
        rc = L.wc_affinity_refine(_lib.ptr(attn), nl, None, _lib.ptr(ap), N, h, h, 1, _lib.ptr(cams), _lib.int_array(k_off),
                                  ctypes.c_double(0.4), _lib.ptr(out), None, None, _lib.ptr(ws), ctypes.c_size_t(ws.numel()),
                                  _lib.stream_ptr())
        _lib.check(rc, "wc_affinity_refine")
    ms = timed(run)
    alg = N * (4 * P * P * (nl + 2) + 7 * 4 * P * P)
    return dict(P=P, N=N, ms=round(ms, 3), gbs=round(alg / ms / 1e6, 1), frac=round(alg / ms / 1e6 / pk, 3))


def main():
    pk = peak()
    print(f"# PAR / affinity-refinement sweep vs HBM roofline; peak {pk} GB/s (MEASURED_PEAKS.json copy bandwidth)")
    print("# PAR, full dilation set: whole PAR.forward (affinity build + 20 steps) and one propagation step alone")
    for dt in (torch.float16, torch.float32):
        for Kc in (2, 3, 6):
            for H in (128, 192, 256, 320, 384, 512, 640, 768, 1024):
                print(json.dumps(par_case(H, Kc, DIL, dt, pk)), flush=True)
    print("# PAR, each single dilation (generic kernels), Kc = 3, 512 x 512")
    for dt in (torch.float16, torch.float32):
        for d in DIL:
            print(json.dumps(par_case(512, 3, [d], dt, pk)), flush=True)
    print("# attention-affinity refinement (wc_affinity_refine), RFM window 6, K = 2")
    for h in (8, 12, 16, 20, 24, 32, 40, 48, 64):
        print(json.dumps(refine_case(h, pk)), flush=True)


if __name__ == "__main__":
    main()

"""BASELINE.json configs[4] (GPU box): PAR sweep against the HBM roofline.
H = W in {128 .. 1024}, Kc in {2, 3, 6}, the module's dilation set 1/2/4/8/12/24, 20 iterations, fp16 and fp32 affinity storage,
batch 1 (the reference's call shape) and a batch that fills the GPU.  Algorithmic bytes per SURVEY.md §8(d):
    B * H * W * [12 + 48 s_a + 20 (48 s_a + 8 Kc)]      (s_a = bytes per stored affinity)
usage: python tools/par_sweep.py > profiles/<name>.txt"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from weclip_b200 import synth  # noqa: E402
from weclip_b200.par import PAR  # noqa: E402

DIL = [1, 2, 4, 8, 12, 24]


def peak():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    return json.load(open(p))["hbm_gbs"] if os.path.exists(p) else 6650.0


def run(B, H, Kc, dt, iters=5):
    W = H
    s_a = 2 if dt == torch.float16 else 4
    imgs = synth.make_images(H + Kc, B, H, W).cuda()
    masks = torch.rand(B, Kc, H, W, generator=torch.Generator().manual_seed(H)).cuda()
    par = PAR(DIL, 20, affinity_dtype=dt).cuda()
    for _ in range(2):
        par(imgs, masks)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ms = []
    for _ in range(iters):
        flush.zero_()                                   # L2 flush between timed calls
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); par(imgs, masks); e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    t = sorted(ms)[len(ms) // 2]
    alg = B * H * W * (12 + 48 * s_a + 20 * (48 * s_a + 8 * Kc))
    gbs = alg / (t * 1e-3) / 1e9
    return t, gbs


def main():
    pk = peak()
    print(f"# PAR sweep, 20 iterations, dilations {DIL}; HBM peak (measured copy) {pk:.0f} GB/s; median of 5, L2 flushed")
    print(f"# {'B':>3} {'H=W':>5} {'Kc':>3} {'aff':>5} {'ms':>9} {'GB/s':>9} {'frac':>6}")
    for H in (128, 192, 256, 320, 384, 512, 640, 768, 1024):
        for Kc in (2, 3, 6):
            for dt in (torch.float16, torch.float32):
                for B in (1, max(1, (2048 * 2048) // (H * H))):
                    if B * H * H * 48 * (2 if dt == torch.float16 else 4) > 6e9:
                        continue
                    t, g = run(B, H, Kc, dt)
                    print(f"  {B:3d} {H:5d} {Kc:3d} {'f16' if dt == torch.float16 else 'f32':>5} {t:9.3f} {g:9.1f} {g / pk:6.3f}", flush=True)


if __name__ == "__main__":
    main()

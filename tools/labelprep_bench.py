"""GPU box: time the label-prep / loss kernels (labelprep.cu) at the cfg-2 step size against the HBM roofline.
    python tools/labelprep_bench.py > gpurun_out/labelprep_bench.txt
CUDA events on the launching stream, median of 20 after 3 warm-ups, L2 flushed between repeats."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from weclip_b200 import losses as L  # noqa: E402


def timed(fn, flush, reps=20, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        flush.fill_(1.0)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    peak = 6545.0
    try:
        peak = float(json.load(open("MEASURED_PEAKS.json")).get("hbm_gbs", peak))
    except Exception:
        pass
    dev = "cuda"
    flush = torch.empty(256 << 20, dtype=torch.float32, device=dev)
    print(f"# labelprep kernels vs HBM copy peak {peak:.0f} GB/s; median of 20, L2 flushed")
    print("# case                                   ms      algorithmic MB    GB/s   frac")
    for B, H, W, C in [(32, 320, 320, 21), (32, 512, 512, 21), (8, 512, 512, 81)]:
        P = (H // 16) * (W // 16)
        cam = (torch.randint(0, 3, (B, H // 32, W // 32), device=dev).repeat_interleave(32, 1).repeat_interleave(32, 2) * 7)
        cam[:, -40:] = 255
        attn = torch.rand(B, P, P, device=dev)
        pred = torch.randn(B, C, H, W, device=dev)
        mask = L.radius_mask(H // 16, W // 16, 8)
        rows = []
        rows.append(("affinity_label", lambda: L.cams_to_affinity_label(cam, mask), 8 * B * P * P + P * P))
        t = L.cams_to_affinity_label(cam, mask)
        rows.append(("aff_loss(targets) value only", lambda: L.get_aff_loss(attn, t), 12 * B * P * P))
        rows.append(("aff_loss fused value only", lambda: L.aff_loss_from_labels(attn, cam, 8), 4 * B * P * P))
        xg = attn.clone().requires_grad_(True)
        rows.append(("aff_loss fused value+grad", lambda: L.aff_loss_from_labels(xg, cam, 8), 8 * B * P * P))
        rows.append(("seg_loss value only", lambda: L.get_seg_loss(pred, cam), 4 * B * C * H * W + 16 * B * H * W))
        pg = pred.clone().requires_grad_(True)
        rows.append(("seg_loss value+grad", lambda: L.get_seg_loss(pg, cam), 8 * B * C * H * W + 16 * B * H * W))
        for name, fn, nbytes in rows:
            ms = timed(fn, flush)
            gbs = nbytes / ms / 1e6
            print(f"B={B:<3d}{H}x{W} C={C:<3d} {name:30s} {ms:8.4f} {nbytes / 1e6:12.1f} {gbs:9.1f} {gbs / peak:6.3f}")


if __name__ == "__main__":
    main()

import os, sys, statistics, torch
sys.path.insert(0, '/root/repo')
from weclip_b200 import ops, synth
H, Kc = int(sys.argv[1]), int(sys.argv[2])
B = (4 << 20) // (H * H)
DIL = [1, 2, 4, 8, 12, 24]
imgs = synth.make_images(H, 1, H, H).cuda().expand(B, 3, H, H).contiguous()
masks = torch.rand(B * Kc, H, H, device="cuda")
ch_off = [b * Kc for b in range(B + 1)]
raw = ops.par_affinity(imgs, DIL, dtype=torch.float16, raw=True)
for _ in range(2): ops.par_propagate(raw, masks, ch_off, DIL, 20)
torch.cuda.synchronize()
ts = []
for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ops.par_propagate(raw, masks, ch_off, DIL, 20); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) / 20 * 1e3)
print(f"H={H} Kc={Kc} PIPE={os.environ.get('WC_PAR_PIPE','1')} DBG={os.environ.get('WC_PAR_DBG','0')}: {statistics.median(ts):.1f} us/step")

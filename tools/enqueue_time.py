"""Developer check (GPU box): host time to enqueue one msc+flip step vs its device time."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from weclip_b200 import bench_workloads as bw

wl = bw.WORKLOADS["mscflip"](0, torch.device("cuda"))
for _ in range(3):
    wl.step_device()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10):
    wl.step_device()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"overlap={wl.ev.overlap_scales}: host enqueue {1e3 * (t1 - t0) / 10:.2f} ms/step, total {1e3 * (t2 - t0) / 10:.2f} ms/step")
for _ in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    wl.step_device()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"  single step from an idle device: host enqueue {1e3 * (t1 - t0):.2f} ms, total {1e3 * (t2 - t0):.2f} ms")

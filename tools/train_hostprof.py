"""Developer profile (GPU box): where the host time of one training step goes (cProfile over 10 steps of bench.py's train workload)."""
import cProfile, pstats, sys, os, io
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from weclip_b200.bench_workloads import TrainStepVoc
wl = TrainStepVoc(0, torch.device("cuda:0"))
for _ in range(3):
    wl.step_device()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(10):
    wl.step_device()
torch.cuda.synchronize()
pr.disable()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45)
print(s.getvalue()[:9000])

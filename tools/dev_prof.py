"""Developer script (GPU box): per-launch device times of one msc+flip step, grouped by kernel class / GEMM shape."""
import sys, os, ctypes, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from weclip_b200 import _lib
from weclip_b200.bench_workloads import WORKLOADS, prof_summary

def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "mscflip"
    wl = WORKLOADS[name](0, torch.device("cuda", 0))
    b1 = len(sys.argv) > 2 and sys.argv[2] == "b1"       # the one-image-per-call step (8 images of 512 x 512, one call each)
    step = wl.step_b1 if b1 else wl.step_device
    for ev in (getattr(wl, "ev", None), getattr(wl, "ev_b1", None)):
        if ev is not None:       # every kernel alone on the device, launched eagerly (the profiler brackets launches)
            ev.overlap_scales = False
            ev.use_graphs = False
    for _ in range(2):
        step()
    torch.cuda.synchronize()
    L = _lib.lib()
    L.wc_prof_reset(); L.wc_prof_enable(1)
    step()
    summ = prof_summary()
    cap = 1 << 16
    kind = (ctypes.c_int * cap)(); ms = (ctypes.c_double * cap)(); work = (ctypes.c_double * cap)(); tags = (ctypes.c_int * (4 * cap))()
    L.wc_prof_records.restype = ctypes.c_long
    n = L.wc_prof_records(kind, ms, work, tags, ctypes.c_long(cap))
    L.wc_prof_enable(0)
    for k, (m, w, c) in summ.items():
        if c: print(f"{k:18s} {m:9.3f} ms  {c:5d} launches")
    groups = collections.defaultdict(lambda: [0, 0.0, 0.0])
    for i in range(n):
        if kind[i] == 0:
            key = tuple(tags[4 * i:4 * i + 4])
            g = groups[key]; g[0] += 1; g[1] += ms[i]; g[2] += work[i]
    other = collections.defaultdict(lambda: [0, 0.0])
    names = {10: "attention", 2: "softmax/headmean", 4: "par_iter", 3: "par_aff"}
    for i in range(n):
        if kind[i] in names:
            key = (names[kind[i]],) + tuple(tags[4 * i:4 * i + 4])
            other[key][0] += 1; other[key][1] += ms[i]
    for key, (c, m) in sorted(other.items(), key=lambda kv: -kv[1][1]):
        print(f"  {str(key):48s} x{c:4d} {m:9.3f} ms  {1e3 * m / c:9.1f} us each")
    print("GEMM shapes (M,N,K,batch): count, total ms, TFLOP/s")
    for key, (c, m, w) in sorted(groups.items(), key=lambda kv: -kv[1][1]):
        print(f"  {str(key):32s} x{c:4d} {m:9.3f} ms {w / (m * 1e-3) / 1e12:8.1f} TF/s")

if __name__ == "__main__":
    main()

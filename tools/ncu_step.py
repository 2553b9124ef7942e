"""Profiling target (GPU box): one step of a bench workload between cudaProfilerStart/Stop.

    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
        --log-file gpurun_out/launches.csv python tools/ncu_step.py mscflip
    ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:gemm_tc -c 3 \
        -o gpurun_out/prof python tools/ncu_step.py mscflip

Two warm-up steps run outside the profiled range.  Nothing printed by this script is a bench value.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "mscflip"
    sys.argv = sys.argv[:1]
    import bench
    wl = bench.WORKLOADS[name](0, torch.device("cuda", 0))
    for _ in range(2):
        wl.step_device()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    wl.step_device()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("ncu_step done:", name)


if __name__ == "__main__":
    main()

#!/bin/bash
# GPU box: ncu --set full captures of the hot kernels inside one msc+flip step (tools/ncu_step.py brackets the step with
# cudaProfilerStart/Stop).  Usage: tools/ncu_capture.sh <tag>; reports land in gpurun_out/ncu_<tag>_<kernel>.ncu-rep
set -u
TAG=${1:-cap}
python tools/ncu_step.py mscflip > gpurun_out/plain_step.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_step.log; exit 1; }
cap() {  # name regex skip count
  ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:$2" -s "$3" -c "$4" \
      -f -o "gpurun_out/ncu_${TAG}_$1" python tools/ncu_step.py mscflip > "gpurun_out/ncu_${TAG}_$1.log" 2>&1
  tail -1 "gpurun_out/ncu_${TAG}_$1.log"
}
cap attn_fwd attn_fwd_kernel 0 1
cap attn_headmean attn_headmean_kernel 0 1
cap par_iter par_iter_tile_kernel 5 1
cap par_aff par_aff_tile_kernel 0 1
cap gemm gemm_tc_kernel 1 4
ls -la gpurun_out/*.ncu-rep

#!/bin/bash
# GPU box: ncu captures of one msc+flip step (tools/ncu_step.py brackets the step with cudaProfilerStart/Stop).
#   1. per-launch DRAM traffic + duration of EVERY kernel of the step  -> gpurun_out/traffic_<tag>.csv
#   2. --set full captures of the hot kernels                          -> gpurun_out/ncu_<tag>_<kernel>.ncu-rep
# Usage: tools/ncu_capture.sh <tag>.  Nothing timed under ncu is a bench value.
set -u
TAG=${1:-cap}
python tools/ncu_step.py mscflip > gpurun_out/plain_step.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_step.log; exit 1; }
ncu --profile-from-start off --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv \
    --log-file "gpurun_out/traffic_${TAG}.csv" python tools/ncu_step.py mscflip > "gpurun_out/traffic_${TAG}.log" 2>&1
cap() {  # name regex skip count
  ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:$2" -s "$3" -c "$4" \
      -f -o "gpurun_out/ncu_${TAG}_$1" python tools/ncu_step.py mscflip > "gpurun_out/ncu_${TAG}_$1.log" 2>&1
  tail -1 "gpurun_out/ncu_${TAG}_$1.log"
}
cap gemm_pair gemm_tc2_kernel 2 8
cap attn_fwd attn_fwd_kernel 2 1
cap attn_headmean attn_headmean_kernel 0 1
cap attn_bwd attn_bwd_colsum_kernel 0 1
cap par_iter par_iter_tile_kernel 5 1
cap par_aff par_aff_tile_kernel 0 1
ls -la gpurun_out/*.ncu-rep

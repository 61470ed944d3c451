#!/bin/bash
# ncu --set full of the per-image kernels on the bench workload.  usage: tools/gpu_ncu.sh TAG [profile] [kernels...]
TAG=${1:-run}; PROFILE=${2:-uniform}; shift; shift
KS=${@:-map_update decode_nms}
mkdir -p gpurun_out
for K in $KS; do
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:$K -s 2 -c 1 -f -o gpurun_out/${TAG}_$K \
      python tools/run_kernels.py --reps 3 --profile $PROFILE > gpurun_out/${TAG}_ncu_$K.log 2>&1
  echo "ncu $K rc=$?"
done

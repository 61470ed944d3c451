#!/bin/bash
# ncu --set full of one per-image kernel + source dump.  usage: tools/gpu_prof.sh TAG KERNEL [profile]
TAG=$1; K=$2; PROFILE=${3:-uniform}
mkdir -p gpurun_out
timeout 600 ncu --set full --import-source on --clock-control none -k regex:$K -s 2 -c 1 -f -o gpurun_out/${TAG}_$K \
    python tools/run_kernels.py --reps 3 --profile $PROFILE > gpurun_out/${TAG}_ncu_$K.log 2>&1
echo "ncu $K rc=$?"

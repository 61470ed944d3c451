#!/bin/bash
# One GPU-box visit: parity tests, kernel timings, fuzz, and ncu captures of the per-image kernels.
# usage: tools/gpu_round.sh TAG [fuzz_rounds]
TAG=${1:-run}
FUZZ=${2:-150}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${TAG}_pytest.log
tail -5 gpurun_out/${TAG}_pytest.log
python tools/time_kernels.py > gpurun_out/${TAG}_times.log 2>&1; cat gpurun_out/${TAG}_times.log
timeout 600 python tools/fuzz_parity.py --rounds $FUZZ > gpurun_out/${TAG}_fuzz.log 2>&1; tail -3 gpurun_out/${TAG}_fuzz.log
for K in map_update decode_nms; do
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:$K -s 2 -c 1 -f -o gpurun_out/${TAG}_$K \
      python tools/run_kernels.py --reps 3 > gpurun_out/${TAG}_ncu_$K.log 2>&1
  echo "ncu $K rc=$?"
done

python -m pytest tests/test_gpu_parity.py tests/test_training_step.py tests/test_gpu_edges.py -x -q -m gpu 2>&1 | tail -2
python tools/time_loss.py 2>&1 | tail -1
for v in "-DYH_LOSS_MINBLOCKS_BULK=6" "-DYH_LOSS_NO_BULK" "-DYH_LOSS_NO_BULK -DYH_LOSS_MINBLOCKS=6" "-DYH_LOSS_NO_BULK -DYH_LOSS_MINBLOCKS=7"; do
  YH_NVCC_EXTRA="$v" python -m yolo_v1_b200.build >/dev/null 2>&1
  YH_NVCC_EXTRA="$v" python tools/time_loss.py 2>&1 | tail -1
done
python -m yolo_v1_b200.build >/dev/null 2>&1

python -m pytest tests/test_gpu_parity.py tests/test_training_step.py -x -q -m gpu -k "loss or training or autocast or ddp or parameter" 2>&1 | tail -2
python tools/kernel_table.py 2>&1 | grep "loss_kernel\|^sum"
python tools/time_loss.py 2>&1 | tail -3

python -m pytest tests/test_gpu_edges.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
python tools/kernel_table.py 2>&1 | grep "radix\|^sum"
PROFILE=realistic python tools/kernel_table.py 2>&1 | grep "radix\|^sum"
python tools/time_ap.py 2>&1 | tail -1

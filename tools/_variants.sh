python -m pytest tests/test_gpu_edges.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
python tools/kernel_table.py 2>&1 | grep "emit_prec\|^sum"
PROFILE=correlated CONF=0.01 python tools/kernel_table.py 2>&1 | grep "emit_prec\|^sum"

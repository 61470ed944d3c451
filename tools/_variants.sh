python -m pytest tests/test_gpu_edges.py tests/test_gpu_parity.py tests/test_reference_suite.py -x -q -m gpu 2>&1 | tail -2
python tools/kernel_table.py 2>&1 | grep "scan_tile\|tile_bit\|emit\|^sum"
python tools/fuzz_parity.py --rounds 300 --seed0 7000 | tail -2

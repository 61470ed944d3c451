python tools/fuzz_parity.py --rounds 150 2>&1 | tail -8
python tools/sanitize.py 2>&1 | tail -3
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize.py > gpurun_out/memcheck.log 2>&1; echo "memcheck rc=$?"; grep -c "Invalid\|Error" gpurun_out/memcheck.log; tail -3 gpurun_out/memcheck.log

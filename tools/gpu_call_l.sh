#!/bin/bash
# device-side SMF parser: parity tests (short timeout), the whole GPU suite, default bench
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_smf.py -m gpu -x -q > gpurun_out/l_smf.log 2>&1
rc=$?; echo "smf rc=$rc"; tail -15 gpurun_out/l_smf.log
if [ $rc -ne 0 ]; then exit 1; fi
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/l_pytest.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/l_pytest.log
timeout 900 python bench.py > gpurun_out/l_bench.json 2> gpurun_out/l_bench.err
echo "bench rc=$?"; tail -c 1600 gpurun_out/l_bench.json; tail -5 gpurun_out/l_bench.err

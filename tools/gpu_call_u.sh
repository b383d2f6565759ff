#!/bin/bash
# end-of-round confirmation at HEAD: whole GPU suite + default bench
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/u_pytest.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/u_pytest.log
timeout 400 python bench.py > gpurun_out/u_bench.json 2> gpurun_out/u_bench.err
echo "bench rc=$?"; tail -c 900 gpurun_out/u_bench.json; tail -3 gpurun_out/u_bench.err

#!/bin/bash
# round-2 checkpoint call: the allocation-free loss entry alone in a fresh process, the whole GPU suite, both bench arms
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/i_noalloc_alone.log 2>&1
echo "noalloc alone rc=$?" | tee -a gpurun_out/i_noalloc_alone.log
tail -4 gpurun_out/i_noalloc_alone.log
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/i_pytest.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/i_pytest.log
tail -6 gpurun_out/i_pytest.log
timeout 900 python bench.py > gpurun_out/i_bench.json 2> gpurun_out/i_bench.err
echo "bench rc=$?"; tail -c 1500 gpurun_out/i_bench.json
timeout 900 python bench.py --impl reference > gpurun_out/i_bench_ref.json 2> gpurun_out/i_bench_ref.err
echo "bench ref rc=$?"; tail -c 600 gpurun_out/i_bench_ref.json

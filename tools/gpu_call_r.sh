#!/bin/bash
# final evidence of the round: whole GPU suite, default bench, reference arm, ncu launch list of a short bench run
mkdir -p gpurun_out
timeout 420 python -m pytest tests -m gpu -x -q > gpurun_out/r_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/r_pytest.log
timeout 420 python bench.py > gpurun_out/r_bench.json 2> gpurun_out/r_bench.err
echo "bench rc=$?"; tail -c 1300 gpurun_out/r_bench.json
timeout 240 python bench.py --impl reference > gpurun_out/r_bench_ref.json 2> gpurun_out/r_bench_ref.err
echo "ref rc=$?"; tail -c 300 gpurun_out/r_bench_ref.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r_launches_bench.csv \
  python bench.py --steps 2 --warmup 3 --train-steps 1 --no-long --file-songs 0 > gpurun_out/r_ncu.log 2>&1
echo "ncu rc=$?"

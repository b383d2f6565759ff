#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/d_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/d_pytest.log
tail -12 gpurun_out/d_pytest.log
MST_PRECISION=bf16 timeout 600 python tools/train_probe.py 16,64 4,200,4 5 > gpurun_out/d_probe_bf16.log 2>&1; cat gpurun_out/d_probe_bf16.log
MST_PRECISION=bf16 MST_BF16_NO_GEMM=1 timeout 600 python tools/train_probe.py 64 4,200,4 5 > gpurun_out/d_probe_bf16_tails_only.log 2>&1; cat gpurun_out/d_probe_bf16_tails_only.log
timeout 1500 python bench.py --steps 10 --warmup 3 > gpurun_out/d_bench.json 2> gpurun_out/d_bench.err; echo "bench rc=$?"; tail -c 1800 gpurun_out/d_bench.json; tail -5 gpurun_out/d_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/d_bench_ref.json 2> gpurun_out/d_bench_ref.err; echo "ref rc=$?"; tail -c 900 gpurun_out/d_bench_ref.json

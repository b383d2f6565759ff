#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/e_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/e_pytest.log
tail -12 gpurun_out/e_pytest.log
cat gpurun_out/bf16_train_parity.txt
MST_PRECISION=bf16 timeout 600 python tools/train_probe.py 64 4,200,4 5 > gpurun_out/e_probe_bf16.log 2>&1; cat gpurun_out/e_probe_bf16.log
MST_PRECISION=bf16 MST_BF16_NO_TMA=1 timeout 600 python tools/train_probe.py 64 4,200,4 5 > gpurun_out/e_probe_bf16_notma.log 2>&1; cat gpurun_out/e_probe_bf16_notma.log
MST_PRECISION=bf16 MST_PSA_FWD_TC=1 timeout 600 python tools/train_probe.py 64 4,200,4 5 > gpurun_out/e_probe_bf16_fwdtc.log 2>&1; cat gpurun_out/e_probe_bf16_fwdtc.log
timeout 1500 python bench.py --steps 10 --warmup 3 --no-long > gpurun_out/e_bench.json 2> gpurun_out/e_bench.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/e_bench.json; tail -5 gpurun_out/e_bench.err
MST_PRECISION=bf16 MST_PSA_FWD_TC=1 timeout 300 python tools/train_probe.py 16 4,200,4 2 > gpurun_out/e_plain.log 2>&1 && \
MST_PRECISION=bf16 MST_PSA_FWD_TC=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 900 -c 900 --csv --log-file gpurun_out/launches_bf16_N16_v2.csv python tools/train_probe.py 16 4,200,4 2 > gpurun_out/e_ncu.log 2>&1
python tools/launch_summary.py gpurun_out/launches_bf16_N16_v2.csv > gpurun_out/launches_bf16_N16_v2_summary.txt 2>&1; head -24 gpurun_out/launches_bf16_N16_v2_summary.txt

#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/b_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/b_pytest.log
tail -15 gpurun_out/b_pytest.log
timeout 600 python tools/bf16_diag.py 4 12 4 0.003 > gpurun_out/b_diag_small.log 2>&1
timeout 600 python tools/bf16_diag.py 4 100 4 0.003 > gpurun_out/b_diag_big.log 2>&1
cat gpurun_out/b_diag_small.log gpurun_out/b_diag_big.log
for prec in bf16; do
  MST_PRECISION=$prec timeout 600 python tools/train_probe.py 1,16,64 4,200,4 5 > gpurun_out/b_probe_$prec.log 2>&1
  cat gpurun_out/b_probe_$prec.log
done
MST_PRECISION=bf16 MST_BF16_NO_TAILS=1 timeout 600 python tools/train_probe.py 64 4,200,4 5 > gpurun_out/b_probe_bf16_gemm_only.log 2>&1; cat gpurun_out/b_probe_bf16_gemm_only.log
MST_PRECISION=bf16 MST_BF16_NO_GEMM=1 timeout 600 python tools/train_probe.py 64 4,200,4 5 > gpurun_out/b_probe_bf16_tails_only.log 2>&1; cat gpurun_out/b_probe_bf16_tails_only.log
MST_PRECISION=bf16 timeout 300 python tools/train_probe.py 16 4,200,4 2 > gpurun_out/b_plain.log 2>&1 && \
MST_PRECISION=bf16 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 900 -c 900 --csv --log-file gpurun_out/launches_bf16_N16.csv python tools/train_probe.py 16 4,200,4 2 > gpurun_out/b_ncu.log 2>&1
python tools/launch_summary.py gpurun_out/launches_bf16_N16.csv > gpurun_out/launches_bf16_N16_summary.txt 2>&1; head -40 gpurun_out/launches_bf16_N16_summary.txt

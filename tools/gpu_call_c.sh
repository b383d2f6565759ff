#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_transfer.py tests/test_gpu_dropin.py tests/test_gpu_configs.py tests/test_gpu_decode.py -q > gpurun_out/c_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c_pytest.log
tail -15 gpurun_out/c_pytest.log
MST_DIAG_MASKS=1,2,4,8,16,7,24 timeout 900 python tools/bf16_diag.py 4 100 4 0.003 2>&1 | grep -v Warning | grep -v "tot = " | grep -v "Consider" > gpurun_out/c_diag_masks.log
cat gpurun_out/c_diag_masks.log
MST_PRECISION=bf16 timeout 300 python tools/train_probe.py 4 4,200,4 1 > gpurun_out/c_plain.log 2>&1 && \
MST_PRECISION=bf16 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_psa_.*_tc -s 4 -c 2 -o gpurun_out/prof_psa_tc python tools/train_probe.py 4 4,200,4 1 > gpurun_out/c_ncu.log 2>&1
tail -5 gpurun_out/c_ncu.log
ls -la gpurun_out/*.ncu-rep

#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_bf16.py tests/test_dp_gloo.py tests/test_gpu_encode.py -m gpu -q > gpurun_out/f_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/f_pytest.log
tail -8 gpurun_out/f_pytest.log; cat gpurun_out/bf16_train_parity.txt | tail -6
bash tools/gpu_call_f.sh memcheck

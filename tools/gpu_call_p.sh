#!/bin/bash
mkdir -p gpurun_out
MST_MODEL_STREAMS=1 timeout 300 python tools/n1_probe.py 1,16,64 20 fp32 2>&1 | grep -v Warning | tee gpurun_out/p_probe.txt
MST_MODEL_STREAMS=1 timeout 300 python tools/n1_probe.py 1,64 20 bf16 2>&1 | grep -v Warning | tee -a gpurun_out/p_probe.txt
for i in 1 2; do
  timeout 600 python -m pytest tests/test_gpu_train.py tests/test_gpu_model.py tests/test_gpu_bf16.py tests/test_gpu_dropin.py tests/test_gpu_configs.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/p_pytest_$i.log 2>&1
  echo "pytest $i rc=$?"; tail -2 gpurun_out/p_pytest_$i.log
done

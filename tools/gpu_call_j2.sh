#!/bin/bash
# branch concurrency: timing with / without side streams, then the model-side GPU tests three times and the whole suite once
mkdir -p gpurun_out
rm -f gpurun_out/j_probe.txt
for s in 0 1; do
  MST_MODEL_STREAMS=$s timeout 600 python tools/n1_probe.py 1,16,64 20 fp32 2>&1 | grep -v Warning | tee -a gpurun_out/j_probe.txt
done
MST_MODEL_STREAMS=1 timeout 600 python tools/n1_probe.py 1,64 20 bf16 2>&1 | grep -v Warning | tee -a gpurun_out/j_probe.txt
for i in 1 2 3; do
  timeout 1200 python -m pytest tests/test_gpu_train.py tests/test_gpu_model.py tests/test_gpu_bf16.py tests/test_gpu_dropin.py tests/test_gpu_configs.py -m gpu -x -q > gpurun_out/j_pytest_$i.log 2>&1
  echo "pytest $i rc=$?"; tail -3 gpurun_out/j_pytest_$i.log
done
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/j_pytest_full.log 2>&1
echo "pytest full rc=$?"; tail -3 gpurun_out/j_pytest_full.log

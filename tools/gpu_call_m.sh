#!/bin/bash
mkdir -p gpurun_out
timeout 420 python tools/smf_debug.py > gpurun_out/m_smf_debug.txt 2>&1
echo "debug rc=$?"; cat gpurun_out/m_smf_debug.txt | tail -30
MST_MODEL_STREAMS=1 timeout 300 python tools/n1_probe.py 1,64 20 fp32 2>&1 | grep -v Warning | tee gpurun_out/m_probe.txt
timeout 600 python -m pytest tests/test_gpu_train.py tests/test_gpu_model.py -m gpu -x -q > gpurun_out/m_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/m_pytest.log

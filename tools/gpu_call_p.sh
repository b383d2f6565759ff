#!/bin/bash
# short, strictly bounded validation of the model-side schedule
mkdir -p gpurun_out
MST_MODEL_STREAMS=1 timeout 90 python tools/n1_probe.py 1,64 20 fp32 2>&1 | grep -v Warning | tee gpurun_out/p_probe.txt
if [ ! -s gpurun_out/p_probe.txt ]; then echo "probe produced nothing"; exit 1; fi
timeout 240 python -m pytest tests/test_gpu_train.py tests/test_gpu_model.py tests/test_gpu_bf16.py tests/test_gpu_dropin.py tests/test_gpu_configs.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/p_pytest_1.log 2>&1
echo "pytest rc=$?"; tail -2 gpurun_out/p_pytest_1.log
timeout 60 python tools/n1_timeline.py 1 > gpurun_out/q_timeline_graph.txt 2>&1
head -1 gpurun_out/q_timeline_graph.txt

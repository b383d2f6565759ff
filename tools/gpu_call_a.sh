#!/bin/bash
# round-2 GPU call A: full GPU suite (bf16 mode + configs[0] parity included), then the training-step probe in both modes
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/a_gpu.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/a_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -40 gpurun_out/a_pytest.log
for prec in fp32 bf16; do
  MST_PRECISION=$prec timeout 600 python tools/train_probe.py 1,16,64 4,200,4 5 > gpurun_out/a_probe_$prec.log 2>&1
  echo "probe $prec rc=$?" >> gpurun_out/a_probe_$prec.log
  cat gpurun_out/a_probe_$prec.log
done

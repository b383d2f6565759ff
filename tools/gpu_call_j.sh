#!/bin/bash
# branch-concurrency check: N=1 / N=16 / N=64 timing with and without side streams, then the training parity tests three times
mkdir -p gpurun_out
(lscpu | grep -i 'numa\|socket\|^CPU(s)\|model name'; nvidia-smi topo -m; for d in /sys/bus/pci/devices/*; do if [ "$(cat $d/vendor)" = "0x10de" ]; then echo $d $(cat $d/numa_node) $(cat $d/local_cpulist); fi; done; numactl -H 2>/dev/null; cat /sys/devices/system/node/node*/cpulist) > gpurun_out/j_topology.txt 2>&1
for s in 0 1; do
  MST_MODEL_STREAMS=$s timeout 600 python tools/n1_probe.py 1,16,64 20 fp32 2>&1 | grep -v Warning | tee -a gpurun_out/j_probe.txt
done
MST_MODEL_STREAMS=1 timeout 600 python tools/n1_probe.py 1,64 20 bf16 2>&1 | grep -v Warning | tee -a gpurun_out/j_probe.txt
for i in 1 2 3; do
  timeout 1200 python -m pytest tests/test_gpu_train.py tests/test_gpu_model.py tests/test_gpu_bf16.py tests/test_gpu_dropin.py tests/test_gpu_configs.py -m gpu -x -q > gpurun_out/j_pytest_$i.log 2>&1
  echo "pytest $i rc=$?"; tail -3 gpurun_out/j_pytest_$i.log
done
# launch list of the N = 1 iteration (eager, one stream): which kernels make up the serial chain
MST_MODEL_STREAMS=0 MST_PROBE_GRAPH=0 MST_PROBE_WARMUP=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv \
  --log-file gpurun_out/j_launches_n1.csv python tools/n1_probe.py 1 2 fp32 > gpurun_out/j_ncu.log 2>&1
echo "ncu rc=$?"

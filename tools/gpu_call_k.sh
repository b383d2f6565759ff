#!/bin/bash
# 8-GPU host-path diagnosis: topology, concurrent H2D bandwidth per rank (as launched / NUMA-bound), encode-only bench
N=${1:-8}
mkdir -p gpurun_out
(lscpu | grep -i 'numa\|socket\|^CPU(s)\|model name'; nvidia-smi topo -m; cat /sys/devices/system/node/node*/cpulist; free -g) > gpurun_out/k_topology_${N}gpu.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 tools/h2d_probe.py > gpurun_out/k_h2d_${N}gpu.txt 2>&1
echo "probe rc=$?"; cat gpurun_out/k_h2d_${N}gpu.txt | grep -v "^\*\|OMP_NUM\|^$" | tail -60
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 20 --warmup 3 --no-train --file-songs 0 > gpurun_out/k_bench_${N}gpu.json 2> gpurun_out/k_bench_${N}gpu.err
echo "bench rc=$?"; python - <<PY
import json
d=json.loads(open('gpurun_out/k_bench_${N}gpu.json').read().strip().splitlines()[-1])
print(d['value'], json.dumps(d['e2e'])[:900])
PY

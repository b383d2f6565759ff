#!/bin/bash
# multi-GPU bench: bash tools/gpu_call_h.sh <n_gpus>  (torchrun as the driver launches it)
N=${1:-2}
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 10 --warmup 3 --no-long --file-songs 0 > gpurun_out/h_bench_${N}gpu.json 2> gpurun_out/h_bench_${N}gpu.err
echo "bench $N rc=$?"
tail -c 1200 gpurun_out/h_bench_${N}gpu.json
tail -5 gpurun_out/h_bench_${N}gpu.err

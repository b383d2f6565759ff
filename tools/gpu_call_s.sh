#!/bin/bash
mkdir -p gpurun_out
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s_smoke.log 2>&1
echo "smoke rc=$?"; tail -4 gpurun_out/s_smoke.log
timeout 200 python -m pytest tests/test_gpu_smf.py -m gpu -x -q 2>&1 | tail -2

#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/smf_probe.py 512 5 2>&1 | grep -v Warn | tee gpurun_out/o_probe.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/o_launches.csv python tools/smf_probe.py 512 1 > gpurun_out/o_ncu.log 2>&1
echo "ncu rc=$?"
python tools/launch_summary.py gpurun_out/o_launches.csv | head -40

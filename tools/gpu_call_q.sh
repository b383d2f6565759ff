#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/n1_timeline.py 1 > gpurun_out/q_timeline_graph.txt 2>&1
echo "rc=$?"; head -3 gpurun_out/q_timeline_graph.txt; tail -3 gpurun_out/q_timeline_graph.txt
timeout 300 python tools/n1_timeline.py 0 > gpurun_out/q_timeline_eager.txt 2>&1
echo "rc=$?"; head -3 gpurun_out/q_timeline_eager.txt

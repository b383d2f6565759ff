#!/bin/bash
# one ncu --set full capture of the device-side SMF parser kernels (after smf_probe has run clean without ncu in call o)
mkdir -p gpurun_out
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_smf -c 5 -o gpurun_out/t_smf --force-overwrite \
  python tools/smf_probe.py 512 1 > gpurun_out/t_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/t_ncu.log; ls -la gpurun_out/t_smf.ncu-rep

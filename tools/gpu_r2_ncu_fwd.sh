#!/bin/bash
# ncu --set full capture of the prefill kernel on C3 (one launch), after the same command has run clean without ncu.
mkdir -p gpurun_out
python tools/prof_cmd.py fwd > gpurun_out/r2_ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fwd_kernel -s 1 -c 1 -f -o gpurun_out/r2_fwd_prof python tools/prof_cmd.py fwd > gpurun_out/r2_ncu_fwd.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/r2_ncu_fwd.log; ls -la gpurun_out/*.ncu-rep

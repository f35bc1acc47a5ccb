#!/bin/bash
# Sweep of the scheduler's head-group size (probe library: SFA_FWD_GROUP) on C5, C3 and one rank's share of each.
mkdir -p gpurun_out
{
for g in 1 2 4 8 32; do SFA_PROBES_LIB=1 SFA_FWD_GROUP=$g timeout 300 python tools/fwd_time.py c5 6; done
for g in 2 4 8 16 64; do SFA_PROBES_LIB=1 SFA_FWD_GROUP=$g timeout 300 python tools/fwd_time.py c3 10; done
for g in 2 4 8 32; do SFA_PROBES_LIB=1 SFA_FWD_GROUP=$g timeout 300 python tools/fwd_time.py c3r8 20; done
for g in 4 16 64; do SFA_PROBES_LIB=1 SFA_FWD_GROUP=$g timeout 300 python tools/fwd_time.py c2 10; done
} > gpurun_out/r2_group_sweep.log 2>&1
grep fwd_time gpurun_out/r2_group_sweep.log | sed 's/| sustained.*//'

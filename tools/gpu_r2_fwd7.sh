#!/bin/bash
# Round 2, seventh GPU pass: spin (test_wait) instead of suspend (try_wait) on the critical waits.
mkdir -p gpurun_out
{
for cfg in c3 c2 c5; do
  SFA_PROBES_LIB=1 timeout 300 python tools/fwd_time.py $cfg 10
  SFA_PROBES_LIB=1 SFA_FWD_POLL=1 timeout 300 python tools/fwd_time.py $cfg 10
  SFA_PROBES_LIB=1 SFA_FWD_POLL=3 timeout 300 python tools/fwd_time.py $cfg 10
done
} > gpurun_out/r2_fwd7_time.log 2>&1
grep fwd_time gpurun_out/r2_fwd7_time.log

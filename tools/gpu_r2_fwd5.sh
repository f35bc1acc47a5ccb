#!/bin/bash
# Round 2, fifth GPU pass: A/B of the exp2 consumer delay and of the deferred epilogue (probe library), product timing.
mkdir -p gpurun_out
{
for cfg in c3 c2; do
  timeout 300 python tools/fwd_time.py $cfg 10
  SFA_PROBES_LIB=1 timeout 300 python tools/fwd_time.py $cfg 10
  SFA_PROBES_LIB=1 SFA_FWD_DELAY=0 timeout 300 python tools/fwd_time.py $cfg 10
  SFA_PROBES_LIB=1 SFA_FWD_DELAY=3 timeout 300 python tools/fwd_time.py $cfg 10
  SFA_PROBES_LIB=1 SFA_FWD_DEFER=0 timeout 300 python tools/fwd_time.py $cfg 10
done
timeout 300 python tools/fwd_time.py c5 10
SFA_PROBES_LIB=1 SFA_FWD_DEFER=0 timeout 300 python tools/fwd_time.py c5 10
} > gpurun_out/r2_fwd5_time.log 2>&1
grep fwd_time gpurun_out/r2_fwd5_time.log
( timeout 600 python -m pytest tests/test_fwd_gpu.py -q -x 2>&1 | tail -5 ) > gpurun_out/r2_fwd5_pytest.log 2>&1
tail -3 gpurun_out/r2_fwd5_pytest.log

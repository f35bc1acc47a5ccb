#!/bin/bash
# Round 2, eighth GPU pass: smaller head groups (L2), scalar polynomial variants.
mkdir -p gpurun_out
{
for cfg in c3 c2 c5; do timeout 300 python tools/fwd_time.py $cfg 10; done
SFA_PROBES_LIB=1 timeout 300 python tools/fwd_time.py c3 10
SFA_PROBES_LIB=1 SFA_FWD_POLY=12 timeout 300 python tools/fwd_time.py c3 10
SFA_PROBES_LIB=1 SFA_FWD_POLY=11 timeout 300 python tools/fwd_time.py c3 10
SFA_PROBES_LIB=1 timeout 300 python tools/fwd_time.py c2 10
SFA_PROBES_LIB=1 SFA_FWD_POLY=12 timeout 300 python tools/fwd_time.py c2 10
SFA_PROBES_LIB=1 SFA_FWD_POLY=13 timeout 300 python tools/fwd_time.py c2 10
SFA_PROBES_LIB=1 SFA_FWD_POLY=0 timeout 300 python tools/fwd_time.py c2 10
} > gpurun_out/r2_fwd8_time.log 2>&1
grep fwd_time gpurun_out/r2_fwd8_time.log
( timeout 600 python -m pytest tests/test_fwd_gpu.py -q -x 2>&1 | tail -3 ) > gpurun_out/r2_fwd8_pytest.log 2>&1
tail -2 gpurun_out/r2_fwd8_pytest.log

#!/bin/bash
# Round 2: product kernel with L2-sized head groups: all GPU tests, timing of every shape, v1 beside it.
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -12 ) > gpurun_out/r2_fwd11_pytest.log 2>&1
tail -3 gpurun_out/r2_fwd11_pytest.log
{
for cfg in c3 c2 c5 c3nc c3r8 c5r8; do timeout 300 python tools/fwd_time.py $cfg 10; done
for cfg in c3 c2 c5 c3r8; do SFA_FWD_V1=1 timeout 300 python tools/fwd_time.py $cfg 10; done
} > gpurun_out/r2_fwd11_time.log 2>&1
grep fwd_time gpurun_out/r2_fwd11_time.log | sed 's/SM clock median(2nd half) //'

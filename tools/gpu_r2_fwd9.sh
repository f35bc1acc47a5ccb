#!/bin/bash
# Round 2, ninth GPU pass: v4 (one issuer warp per query tile + scheduler / Q-load warp): parity, timing, trace.
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_fwd_gpu.py tests/test_capi_gpu.py -q 2>&1 | tail -12 ) > gpurun_out/r2_fwd10_pytest.log 2>&1
tail -4 gpurun_out/r2_fwd10_pytest.log
{
for cfg in c3 c2 c5 c3nc; do timeout 300 python tools/fwd_time.py $cfg 10; done
SFA_FWD_V1=1 timeout 300 python tools/fwd_time.py c3 10
SFA_FWD_V1=1 timeout 300 python tools/fwd_time.py c5 10
} > gpurun_out/r2_fwd10_time.log 2>&1
grep fwd_time gpurun_out/r2_fwd10_time.log
timeout 300 python tools/fwd_trace.py 30 2 > gpurun_out/r2_fwd10_trace.log 2>&1
grep -E "epilogue of CTA|mean cycles" gpurun_out/r2_fwd10_trace.log

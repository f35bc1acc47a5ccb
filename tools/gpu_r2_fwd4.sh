#!/bin/bash
# Round 2, fourth GPU pass: kernel v3 (role split, deferred epilogue): parity, A/B timing, trace.
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_fwd_gpu.py tests/test_capi_gpu.py -q -x 2>&1 | tail -30 ) > gpurun_out/r2_fwd4_pytest.log 2>&1
tail -6 gpurun_out/r2_fwd4_pytest.log
{
for cfg in c3 c2 c5 c3nc; do
  timeout 300 python tools/fwd_time.py $cfg 10
  SFA_FWD_V1=1 timeout 300 python tools/fwd_time.py $cfg 10
done
SFA_PROBES_LIB=1 timeout 300 python tools/fwd_time.py c3 10
SFA_PROBES_LIB=1 SFA_FWD_POLY=0 timeout 300 python tools/fwd_time.py c2 10
SFA_PROBES_LIB=1 SFA_FWD_POLY=3 timeout 300 python tools/fwd_time.py c2 10
SFA_PROBES_LIB=1 SFA_FWD_POLY=2 timeout 300 python tools/fwd_time.py c3 10
SFA_PROBES_LIB=1 SFA_FWD_STAGGER=0 timeout 300 python tools/fwd_time.py c2 10
} > gpurun_out/r2_fwd4_time.log 2>&1
grep fwd_time gpurun_out/r2_fwd4_time.log
timeout 300 python tools/fwd_trace.py 30 2 > gpurun_out/r2_fwd4_trace.log 2>&1
tail -5 gpurun_out/r2_fwd4_trace.log
( timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -30 ) > gpurun_out/r2_fwd4_pytest_all.log 2>&1
tail -4 gpurun_out/r2_fwd4_pytest_all.log

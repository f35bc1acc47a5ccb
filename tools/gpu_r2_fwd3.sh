#!/bin/bash
# Round 2, third GPU pass: all GPU tests (no -x), prefill A/B (persistent kernel vs round-1 kernel), event trace.
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -40 ) > gpurun_out/r2_fwd3_pytest.log 2>&1
tail -6 gpurun_out/r2_fwd3_pytest.log
{
for cfg in c3 c2 c5; do
  timeout 300 python tools/fwd_time.py $cfg 10
  SFA_FWD_V1=1 timeout 300 python tools/fwd_time.py $cfg 10
done
SFA_PROBES_LIB=1 timeout 300 python tools/fwd_time.py c3 10
} > gpurun_out/r2_fwd3_time.log 2>&1
grep fwd_time gpurun_out/r2_fwd3_time.log
timeout 300 python tools/fwd_trace.py 30 2 > gpurun_out/r2_fwd3_trace.log 2>&1
tail -5 gpurun_out/r2_fwd3_trace.log

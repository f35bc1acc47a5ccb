#!/bin/bash
# Round 2, first GPU pass of the persistent prefill kernel: parity, then A/B timing against the round-1 kernel.
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_fwd_gpu.py -x -q -s 2>&1 | tail -25 ) > gpurun_out/r2_fwd1_pytest.log 2>&1
tail -5 gpurun_out/r2_fwd1_pytest.log
{
for cfg in c3 c2 c5 c3nc; do
  timeout 300 python tools/fwd_time.py $cfg 10
  SFA_FWD_V1=1 timeout 300 python tools/fwd_time.py $cfg 10
  SFA_PROBES_LIB=1 SFA_FWD_SOFTREL=1 timeout 300 python tools/fwd_time.py $cfg 10
done
} > gpurun_out/r2_fwd1_time.log 2>&1
grep fwd_time gpurun_out/r2_fwd1_time.log
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/r2_fwd1_pytest_all.log 2>&1
tail -5 gpurun_out/r2_fwd1_pytest_all.log

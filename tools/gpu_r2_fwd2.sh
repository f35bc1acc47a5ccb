#!/bin/bash
# Round 2, second GPU pass: all GPU tests, prefill A/B (persistent kernel vs round-1 kernel), first bench line.
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 ) > gpurun_out/r2_fwd2_pytest.log 2>&1
tail -4 gpurun_out/r2_fwd2_pytest.log
{
for cfg in c3 c2 c5; do
  timeout 300 python tools/fwd_time.py $cfg 10
  SFA_FWD_V1=1 timeout 300 python tools/fwd_time.py $cfg 10
done
SFA_PROBES_LIB=1 SFA_FWD_GROUP=2 timeout 300 python tools/fwd_time.py c5 10
SFA_PROBES_LIB=1 SFA_FWD_GROUP=4 timeout 300 python tools/fwd_time.py c3 10
SFA_PROBES_LIB=1 SFA_FWD_GROUP=16 timeout 300 python tools/fwd_time.py c3 10
SFA_PROBES_LIB=1 SFA_FWD_STAGGER=0 timeout 300 python tools/fwd_time.py c2 10
} > gpurun_out/r2_fwd2_time.log 2>&1
grep fwd_time gpurun_out/r2_fwd2_time.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_fwd2_bench.json 2> gpurun_out/r2_fwd2_bench.err
tail -c 1500 gpurun_out/r2_fwd2_bench.err
head -c 6000 gpurun_out/r2_fwd2_bench.json

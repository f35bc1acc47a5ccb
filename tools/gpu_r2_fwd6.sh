#!/bin/bash
# Round 2, sixth GPU pass: product = role split without the deferred epilogue; timing, trace of the epilogue, tests.
mkdir -p gpurun_out
{
for cfg in c3 c2 c5 c3nc; do
  timeout 300 python tools/fwd_time.py $cfg 10
done
SFA_FWD_V1=1 timeout 300 python tools/fwd_time.py c3 10
SFA_FWD_V1=1 timeout 300 python tools/fwd_time.py c2 10
} > gpurun_out/r2_fwd6_time.log 2>&1
grep fwd_time gpurun_out/r2_fwd6_time.log
timeout 300 python tools/fwd_trace.py 30 2 > gpurun_out/r2_fwd6_trace.log 2>&1
grep -E "epilogue of CTA|mean cycles" gpurun_out/r2_fwd6_trace.log
( timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -8 ) > gpurun_out/r2_fwd6_pytest.log 2>&1
tail -3 gpurun_out/r2_fwd6_pytest.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_fwd6_bench.json 2> gpurun_out/r2_fwd6_bench.err
tail -c 600 gpurun_out/r2_fwd6_bench.err; head -c 1500 gpurun_out/r2_fwd6_bench.json

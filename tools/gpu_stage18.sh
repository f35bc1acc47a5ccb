#!/bin/bash
# Stage 18: speculative softmax (row-max scan under the exp2 stream) vs the two-phase form (SFA_FWD_DBG=7),
# test_wait polling for the MMA / softmax warps (SFA_FWD_SPIN bits).
mkdir -p gpurun_out
run() { timeout 120 "$@" 2>&1 | grep "fwd_time\|probe" | cut -c1-230 || echo "FAILED: $*"; }
timeout 600 python -m pytest tests/test_fwd_gpu.py -m gpu -q -x 2>&1 | tail -3
for cfg in c3 c2; do
  echo "== $cfg legacy";  SFA_FWD_DBG=7 run python tools/fwd_time.py $cfg 10
  for sp in 0 1 2 3; do echo "== $cfg spec spin=$sp"; SFA_FWD_SPIN=$sp run python tools/fwd_time.py $cfg 10; done
done
SFA_FWD_DBG=5 run python tools/fwd_time.py c3 5
SFA_FWD_DBG=6 timeout 120 python tools/fwd_trace.py 30 2 > gpurun_out/trace18.txt 2>&1
tail -2 gpurun_out/trace18.txt

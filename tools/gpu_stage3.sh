#!/bin/bash
# Prefill tuning sweep.  A hung kernel would hang every later process too, so bail out at the first failure.
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_fwd_gpu.py -m gpu -q -x > gpurun_out/t_fwd.log 2>&1; rc=$?
echo "fwd tests rc=$rc"; tail -4 gpurun_out/t_fwd.log
[ $rc -ne 0 ] && exit 1
run() { timeout 90 "$@" || { echo "FAILED: $*"; exit 1; }; }
for poly in 0 2 3 4; do SFA_FWD_POLY=$poly run python tools/fwd_time.py c3; done
for dbg in 1 2 3; do SFA_FWD_DBG=$dbg run python tools/fwd_time.py c3; done
run python tools/fwd_time.py c3nc
for poly in 0 3 4; do SFA_FWD_POLY=$poly run python tools/fwd_time.py c2; done
run python tools/fwd_time.py c2b
run python tools/fwd_time.py c5 5

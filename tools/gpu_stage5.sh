#!/bin/bash
timeout 200 python -m pytest tests/test_fwd_gpu.py -m gpu -q -x > gpurun_out/t_fwd.log 2>&1; rc=$?
echo "fwd tests rc=$rc"; tail -4 gpurun_out/t_fwd.log
[ $rc -ne 0 ] && exit 1
run() { timeout 90 "$@" || { echo "FAILED: $*"; exit 1; }; }
for dbg in 4 6 5 2; do SFA_FWD_DBG=$dbg run python tools/fwd_time.py c3 5; done
SFA_FWD_POLY=0 run python tools/fwd_time.py c3
SFA_FWD_POLY=2 run python tools/fwd_time.py c3
SFA_FWD_POLY=4 run python tools/fwd_time.py c3
SFA_FWD_POLY=0 run python tools/fwd_time.py c2

#!/bin/bash
timeout 200 python -m pytest tests/test_fwd_gpu.py -m gpu -q -x > gpurun_out/t_fwd.log 2>&1; rc=$?
echo "fwd tests rc=$rc"; tail -4 gpurun_out/t_fwd.log
[ $rc -ne 0 ] && exit 1
SFA_FWD_SPLIT=2 timeout 200 python -m pytest tests/test_fwd_gpu.py -m gpu -q -x > gpurun_out/t_fwd_split2.log 2>&1; rc=$?
echo "fwd tests (split 2) rc=$rc"; tail -4 gpurun_out/t_fwd_split2.log
[ $rc -ne 0 ] && exit 1
run() { timeout 90 "$@" || { echo "FAILED: $*"; exit 1; }; }
SFA_FWD_SPLIT=1 run python tools/fwd_time.py c3
SFA_FWD_SPLIT=1 SFA_FWD_POLY=2 run python tools/fwd_time.py c3
SFA_FWD_SPLIT=2 run python tools/fwd_time.py c3
SFA_FWD_SPLIT=1 SFA_FWD_DBG=5 run python tools/fwd_time.py c3 5
SFA_FWD_SPLIT=1 run python tools/fwd_time.py c2
SFA_FWD_SPLIT=1 run python tools/fwd_time.py c3nc

#!/bin/bash
# Stage 6: decoupled S/P TMEM protocol of the prefill kernel - parity first, then timing and probes.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_fwd_gpu.py -m gpu -q -x > gpurun_out/t_fwd.log 2>&1; rc=$?
echo "fwd tests rc=$rc"; tail -6 gpurun_out/t_fwd.log
[ $rc -ne 0 ] && { timeout 120 python tools/fwd_probe.py 2>&1 | tail -30; exit 1; }
run() { timeout 120 "$@" || { echo "FAILED: $*"; }; }
run python tools/fwd_time.py c3
SFA_FWD_STAGES=5 run python tools/fwd_time.py c3
SFA_FWD_DBG=5 run python tools/fwd_time.py c3 5
SFA_FWD_POLY=2 run python tools/fwd_time.py c3
SFA_FWD_POLY=3 run python tools/fwd_time.py c3
SFA_FWD_DBG=2 run python tools/fwd_time.py c3 5
SFA_FWD_DBG=4 run python tools/fwd_time.py c3 5
run python tools/fwd_time.py c2
SFA_FWD_POLY=3 run python tools/fwd_time.py c2
SFA_FWD_DBG=5 run python tools/fwd_time.py c2 5
run python tools/fwd_time.py c3nc
run python tools/fwd_time.py c5 5
timeout 400 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; echo "all gpu tests rc=$?"; tail -4 gpurun_out/t_gpu.log

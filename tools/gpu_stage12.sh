#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_fwd_gpu.py -m gpu -q -x > gpurun_out/t_fwd.log 2>&1; echo "fwd tests rc=$?"; tail -3 gpurun_out/t_fwd.log
run() { timeout 120 "$@" || echo "FAILED: $*"; }
run python tools/fwd_time.py c3
SFA_FWD_POLY=2 run python tools/fwd_time.py c3
SFA_FWD_DBG=5 run python tools/fwd_time.py c3 5
run python tools/fwd_trace.py 31 1
run python tools/fwd_time.py c2
run python tools/fwd_time.py c3nc

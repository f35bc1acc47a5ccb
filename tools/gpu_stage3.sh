#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_fwd_gpu.py -m gpu -q -x > gpurun_out/t_fwd.log 2>&1; echo "fwd tests rc=$?"; tail -4 gpurun_out/t_fwd.log
for poly in 0 2 3 4; do SFA_FWD_POLY=$poly timeout 120 python tools/fwd_time.py c3; done
for dbg in 1 2 3; do SFA_FWD_DBG=$dbg timeout 120 python tools/fwd_time.py c3; done
timeout 120 python tools/fwd_time.py c3nc
for poly in 0 3 4; do SFA_FWD_POLY=$poly timeout 120 python tools/fwd_time.py c2; done
timeout 120 python tools/fwd_time.py c2b
timeout 120 python tools/fwd_time.py c5 5

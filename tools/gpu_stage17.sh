#!/bin/bash
# Stage 17: event trace + probe of the current prefill kernel (baseline for the round's last tuning pass).
mkdir -p gpurun_out
SFA_FWD_DBG=6 timeout 120 python tools/fwd_trace.py 30 2 > gpurun_out/trace17.txt 2>&1
tail -3 gpurun_out/trace17.txt
SFA_FWD_DBG=5 timeout 120 python tools/fwd_time.py c3 5 2>&1 | grep "fwd_time\|probe" | cut -c1-200
timeout 120 python tools/fwd_time.py c3 10 2>&1 | grep "fwd_time\|probe" | cut -c1-260

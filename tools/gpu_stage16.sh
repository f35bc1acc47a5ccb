#!/bin/bash
# Stage 16: head start of softmax warpgroup 0 (SFA_FWD_STAGGER cycles) at head_dim 64 and 128.
mkdir -p gpurun_out
run() { timeout 120 "$@" 2>&1 | grep "fwd_time\|probe" | cut -c1-150 || echo "FAILED: $*"; }
for st in 0 500 900 1300 1700; do echo "== stagger $st"; SFA_FWD_STAGGER=$st run python tools/fwd_time.py c2 20; SFA_FWD_STAGGER=$st run python tools/fwd_time.py c2b 20; done
SFA_FWD_STAGGER=1300 SFA_FWD_DBG=5 run python tools/fwd_time.py c2 5
for st in 0 1200; do echo "== c3 stagger $st"; SFA_FWD_STAGGER=$st run python tools/fwd_time.py c3; done
SFA_FWD_STAGGER=1300 timeout 300 python -m pytest tests/test_fwd_gpu.py -m gpu -q -x 2>&1 | tail -2

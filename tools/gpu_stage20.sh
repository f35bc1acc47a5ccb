#!/bin/bash
# Stage 20: test_wait polling (SFA_FWD_SPIN bits: 1 MMA S-free/P-ready, 2 softmax S-full, 4 MMA K/V-full) on the product kernel.
run() { timeout 120 "$@" 2>&1 | grep "fwd_time\|probe" | cut -c1-130 || echo "FAILED: $*"; }
for sp in 0 1 2 3 5 7; do echo "== c3 spin=$sp"; SFA_FWD_SPIN=$sp run python tools/fwd_time.py c3 10; done
for sp in 0 2 7; do echo "== c2 spin=$sp"; SFA_FWD_SPIN=$sp run python tools/fwd_time.py c2 10; done

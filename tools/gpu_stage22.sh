#!/bin/bash
# Stage 22: polynomial exp2 share at head_dim 64 with the staggered warpgroups (default 1700 cycles).
run() { timeout 120 "$@" 2>&1 | grep "fwd_time" | cut -c1-110 || echo "FAILED: $*"; }
for cfg in c2 c2b; do
  for po in 0 2 4; do echo -n "poly $po: "; SFA_FWD_POLY=$po run python tools/fwd_time.py $cfg 20; done
done
timeout 600 python -m pytest tests/test_fwd_gpu.py -m gpu -q -x 2>&1 | tail -2

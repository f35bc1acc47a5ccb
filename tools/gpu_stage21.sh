#!/bin/bash
# Stage 21: head start of softmax warpgroup 0 at head_dim 64 over several shapes (is a default worth it?).
run() { timeout 120 "$@" 2>&1 | grep "fwd_time" | cut -c1-110 || echo "FAILED: $*"; }
for cfg in c2 c2b d64c d64s d64l; do
  for st in 0 1300 1700 2100 2600; do echo -n "stagger $st: "; SFA_FWD_STAGGER=$st run python tools/fwd_time.py $cfg 20; done
done

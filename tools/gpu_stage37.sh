#!/bin/bash
# Stage 37: prologue load order of the prefill kernel (Q0 K0 first against Q0 Q1 K0): parity, then C3 / C2 / causal N=2048 timing, A/B/A/B.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_fwd_gpu.py -m gpu -q -x 2>&1 | tail -2
run() { timeout 120 "$@" 2>&1 | grep "fwd_time" | cut -c1-120 || echo "FAILED: $*"; }
for rep in 1 2; do
  for lo in 1 0; do
    for cfg in c3 c2; do echo -n "order $lo: "; SFA_FWD_LOADORDER=$lo run python tools/fwd_time.py $cfg 20; done
  done
done

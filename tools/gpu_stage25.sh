#!/bin/bash
# Stage 25: decoder-step latency at the reference example's shapes (B=2, H=32 MHA, 4 layers), eager and graph.
mkdir -p gpurun_out
for cfg in "2 32 4096 4" "2 32 8192 4" "1 32 512 4"; do
  echo "== $cfg"; timeout 200 python tools/decoder_step_time.py $cfg 2>&1 | grep decoder_step
done

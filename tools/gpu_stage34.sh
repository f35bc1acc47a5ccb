#!/bin/bash
# Stage 34: programmatic dependent launch of the combine kernel (two-launch form): decode tests, sweep both ways.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_decode_gpu.py tests/test_shard_gpu.py -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; rc=$?
echo "gpu decode tests rc=$rc"; tail -3 gpurun_out/t_gpu.log
for pdl in 1 0; do
echo "== SFA_DECODE_PDL=$pdl"
SFA_DECODE_PDL=$pdl timeout 600 python tools/decode_split_sweep.py 2>&1 | grep split_sweep | cut -c1-115
echo "-- decoder step 1 8 32768 4"; SFA_DECODE_PDL=$pdl timeout 200 python tools/decoder_step_time.py 1 8 32768 4 2>&1 | grep "decoder_step" | head -3
done

#!/bin/bash
# Stage 9 (1 GPU): cross-GPU sharded decode emulated on one device + the single-rank bench line.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_shard_gpu.py -m gpu -q -x > gpurun_out/t_shard.log 2>&1; echo "shard tests rc=$?"; tail -8 gpurun_out/t_shard.log
timeout 200 python tools/shard_decode_bench.py 262144 20 2>&1 | tail -3

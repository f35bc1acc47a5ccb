#!/bin/bash
# Stage 27: in-cluster merge of the split-KV decode: decode tests, then the split / merge-form sweep.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_decode_gpu.py tests/test_shard_gpu.py -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; rc=$?
echo "gpu decode tests rc=$rc"; tail -8 gpurun_out/t_gpu.log
[ $rc -eq 124 ] && exit 1
timeout 600 python tools/decode_split_sweep.py > gpurun_out/split_sweep.log 2>&1; echo "sweep rc=$?"; grep split_sweep gpurun_out/split_sweep.log

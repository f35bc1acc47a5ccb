#!/bin/bash
# Stage 30: clusters of up to 16 splits (non-portable sizes): decode tests, sweep with 9 .. 16 cluster splits, C4 bench.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_decode_gpu.py tests/test_shard_gpu.py -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; rc=$?
echo "gpu decode tests rc=$rc"; tail -8 gpurun_out/t_gpu.log
timeout 600 python tools/decode_split_sweep.py > gpurun_out/split_sweep.log 2>&1; echo "sweep rc=$?"; grep split_sweep gpurun_out/split_sweep.log | cut -c1-700
timeout 300 python bench.py --workload c4 --steps 30 --no-secondary --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench c4 rc=$?"; cut -c1-160 gpurun_out/bench_c4.json; tail -3 gpurun_out/bench_c4.err

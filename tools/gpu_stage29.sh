#!/bin/bash
# Stage 29: one-wave threshold at 10: C4 bench (default and two-launch), decoder step, sweep (auto column), decode tests.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_decode_gpu.py tests/test_shard_gpu.py -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; rc=$?
echo "gpu decode tests rc=$rc"; tail -3 gpurun_out/t_gpu.log
for i in 1 2; do
timeout 300 python bench.py --workload c4 --steps 30 --no-secondary --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench c4 rc=$?"; cut -c1-160 gpurun_out/bench_c4.json; tail -3 gpurun_out/bench_c4.err
SFA_DECODE_CLUSTER=0 timeout 300 python bench.py --workload c4 --steps 30 --no-secondary --no-cpu-baseline > gpurun_out/bench_c4_nocluster.json 2>/dev/null; echo "bench c4 (two-launch) rc=$?"; cut -c1-160 gpurun_out/bench_c4_nocluster.json
done
for cfg in "2 32 4096 4" "8 8 32768 4"; do
  echo "== decoder step $cfg"; timeout 200 python tools/decoder_step_time.py $cfg 2>&1 | grep decoder_step
done
timeout 600 python tools/decode_split_sweep.py > gpurun_out/split_sweep.log 2>&1; echo "sweep rc=$?"; grep split_sweep gpurun_out/split_sweep.log | cut -c1-140
timeout 300 python tools/decode_paged_time.py > gpurun_out/paged_c4.log 2>&1; grep "decode_paged\]" gpurun_out/paged_c4.log

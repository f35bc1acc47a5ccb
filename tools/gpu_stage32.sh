#!/bin/bash
# Stage 32: in-kernel merge across clusters (tickets): GPU tests, sweep, decoder step timings, C4 bench.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; rc=$?
echo "gpu tests rc=$rc"; tail -12 gpurun_out/t_gpu.log
[ $rc -eq 124 ] && exit 1
timeout 600 python tools/decode_split_sweep.py > gpurun_out/split_sweep.log 2>&1; echo "sweep rc=$?"; grep split_sweep gpurun_out/split_sweep.log | cut -c1-600
for cfg in "1 8 32768 4" "2 32 4096 4"; do
  echo "== decoder step $cfg"; timeout 200 python tools/decoder_step_time.py $cfg 2>&1 | grep decoder_step
done
timeout 300 python bench.py --workload c4 --steps 30 --no-secondary --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench c4 rc=$?"; cut -c1-160 gpurun_out/bench_c4.json; tail -3 gpurun_out/bench_c4.err
timeout 200 python tools/shard_decode_bench.py 262144 10 2>&1 | tail -1 | cut -c1-600

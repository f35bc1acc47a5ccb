#!/bin/bash
# Stage 28: one-wave cluster splits as the launcher's default: all GPU tests, smoke, C4 bench both ways, decoder step, sweep.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; rc=$?
echo "gpu tests rc=$rc"; tail -8 gpurun_out/t_gpu.log
[ $rc -eq 124 ] && exit 1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 300 python bench.py --workload c4 --steps 30 --no-secondary --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench c4 rc=$?"; cut -c1-330 gpurun_out/bench_c4.json; tail -3 gpurun_out/bench_c4.err
SFA_DECODE_CLUSTER=0 timeout 300 python bench.py --workload c4 --steps 30 --no-secondary --no-cpu-baseline > gpurun_out/bench_c4_nocluster.json 2>/dev/null; echo "bench c4 (two-launch) rc=$?"; cut -c1-330 gpurun_out/bench_c4_nocluster.json
for cfg in "2 32 4096 4" "1 32 512 4" "8 8 32768 4"; do
  echo "== decoder step $cfg"; timeout 200 python tools/decoder_step_time.py $cfg 2>&1 | grep decoder_step
done
timeout 600 python tools/decode_split_sweep.py > gpurun_out/split_sweep.log 2>&1; echo "sweep rc=$?"; grep split_sweep gpurun_out/split_sweep.log | cut -c1-140
timeout 300 python tools/decode_paged_time.py > gpurun_out/paged_c4.log 2>&1; grep "decode_paged\]" gpurun_out/paged_c4.log

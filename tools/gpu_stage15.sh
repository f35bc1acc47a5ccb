#!/bin/bash
# Stage 15: transposed decode as default - whole GPU suite, C4 bench line, ncu of the decode kernels, B=1 long sequence.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; rc=$?; echo "gpu tests rc=$rc"; tail -4 gpurun_out/t_gpu.log
[ $rc -ne 0 ] && exit 1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 300 python bench.py --workload c4 --steps 30 --no-secondary > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench c4 rc=$?"; cat gpurun_out/bench_c4.json
timeout 200 python tools/shard_decode_bench.py 262144 10 2>&1 | tail -1
timeout 200 python tools/decoder_step_time.py 2>&1 | tail -5
for w in decode step; do
  python tools/prof_cmd.py $w > gpurun_out/plain_$w.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:decode_splitkv -s 3 -c 1 -f -o gpurun_out/r1f_$w python tools/prof_cmd.py $w > gpurun_out/ncu_$w.log 2>&1
  echo "ncu $w rc=$?"
done

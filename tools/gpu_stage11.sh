#!/bin/bash
# Stage 11: parity of the emulated exchange, then the ncu evidence for the round (launch list + full sets).
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_shard_gpu.py tests/test_decode_gpu.py -m gpu -q -x > gpurun_out/t_shard.log 2>&1; rc=$?; echo "shard+decode tests rc=$rc"; tail -5 gpurun_out/t_shard.log
[ $rc -ne 0 ] && exit 1
BENCH="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-secondary"
$BENCH > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1d_launches_bench_c3.csv $BENCH > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
for w in fwd decode step fwd64; do
  case $w in fwd|fwd64) pat=fwd_kernel;; *) pat=decode_splitkv;; esac
  python tools/prof_cmd.py $w > gpurun_out/plain_$w.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:$pat -s 3 -c 1 -f -o gpurun_out/r1d_$w python tools/prof_cmd.py $w > gpurun_out/ncu_$w.log 2>&1
  echo "ncu $w rc=$?"; tail -2 gpurun_out/ncu_$w.log
done
ls -la gpurun_out/*.ncu-rep

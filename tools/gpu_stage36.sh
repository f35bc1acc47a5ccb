#!/bin/bash
# Stage 36 (stage 31 again after the dependent-launch change, without the prefill capture): final pass of the round: all GPU tests, smoke, examples, every bench line, the reference arm, and the ncu
# evidence of the final kernels (launch lists of the bench commands, full capture of the cluster-merge decode kernel).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; rc=$?
echo "gpu tests rc=$rc"; tail -4 gpurun_out/t_gpu.log
[ $rc -eq 124 ] && exit 1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 120 ./build/testFlashDecoder > gpurun_out/example_cpp.log 2>&1; echo "cpp example rc=$?"; tail -6 gpurun_out/example_cpp.log
timeout 120 python examples/python/testFlashDecoder.py > gpurun_out/example_py.log 2>&1; echo "py example rc=$?"; tail -2 gpurun_out/example_py.log
timeout 400 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench c3 rc=$?"; cut -c1-200 gpurun_out/bench_c3.json; tail -3 gpurun_out/bench_c3.err
for w in c4 c2 c5; do
  timeout 300 python bench.py --workload $w --steps 20 --no-secondary > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"; cut -c1-200 gpurun_out/bench_$w.json; tail -3 gpurun_out/bench_$w.err
done
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"; cut -c1-200 gpurun_out/bench_ref.json
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1g_launches_bench_c3.csv \
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/ncu_c3.log 2>&1; echo "ncu list c3 rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1g_launches_bench_c4.csv \
  python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/ncu_c4.log 2>&1; echo "ncu list c4 rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:decode_splitkv -s 3 -c 1 -o gpurun_out/r1g_decode_cluster \
  python tools/prof_cmd.py decode > gpurun_out/ncu_decode.log 2>&1; echo "ncu full decode rc=$?"; tail -1 gpurun_out/ncu_decode.log

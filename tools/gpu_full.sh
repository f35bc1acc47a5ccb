#!/bin/bash
# Full single-GPU pass: every GPU test, smoke, both examples, the bench lines of every workload and the reference arm.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; rc=$?
echo "gpu tests rc=$rc"; tail -6 gpurun_out/t_gpu.log
[ $rc -eq 124 ] && exit 1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 120 ./build/testFlashDecoder > gpurun_out/example_cpp.log 2>&1; echo "cpp example rc=$?"; tail -4 gpurun_out/example_cpp.log
timeout 120 python examples/python/testFlashDecoder.py > gpurun_out/example_py.log 2>&1; echo "py example rc=$?"; tail -2 gpurun_out/example_py.log
timeout 400 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench c3 rc=$?"; cat gpurun_out/bench_c3.json; tail -3 gpurun_out/bench_c3.err
for w in c4 c2 c5; do
  timeout 300 python bench.py --workload $w --steps 20 --no-secondary > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"; cat gpurun_out/bench_$w.json; tail -3 gpurun_out/bench_$w.err
done
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"; cat gpurun_out/bench_ref.json
timeout 200 python tools/shard_decode_bench.py 262144 10 2>&1 | tail -1

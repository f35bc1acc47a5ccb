#!/bin/bash
# Stage 23: paged split-KV decode (scope row f1): all GPU tests, the paged-vs-dense timing at the C4 shape,
# the bench lines, and an ncu capture of the paged kernel.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/t_gpu.log 2>&1; rc=$?
echo "gpu tests rc=$rc"; tail -6 gpurun_out/t_gpu.log
[ $rc -eq 124 ] && exit 1
timeout 300 python tools/decode_paged_time.py > gpurun_out/paged_c4.log 2>&1; echo "paged time rc=$?"; cat gpurun_out/paged_c4.log
timeout 300 python tools/decode_paged_time.py 1 262144 > gpurun_out/paged_b1_256k.log 2>&1; echo "paged b1 rc=$?"; cat gpurun_out/paged_b1_256k.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 120 ./build/testFlashDecoder > gpurun_out/example_cpp.log 2>&1; echo "cpp example rc=$?"; tail -3 gpurun_out/example_cpp.log
timeout 120 python examples/python/testFlashDecoder.py > gpurun_out/example_py.log 2>&1; echo "py example rc=$?"; tail -2 gpurun_out/example_py.log
timeout 400 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench c3 rc=$?"; cat gpurun_out/bench_c3.json; tail -3 gpurun_out/bench_c3.err
timeout 300 python bench.py --workload c4 --steps 30 --no-secondary > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench c4 rc=$?"; cat gpurun_out/bench_c4.json; tail -3 gpurun_out/bench_c4.err
timeout 300 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:false, true, true>' -s 5 -c 1 -o gpurun_out/r1g_decode_paged \
  python tools/decode_paged_time.py 8 32768 > gpurun_out/ncu_paged.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_paged.log

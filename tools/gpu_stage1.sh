#!/bin/bash
# First GPU pass: parity tests, smoke, short benches.  Every stage has its own timeout so that a hang
# cannot hold the box; logs go to gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 600 python -m pytest tests/test_decode_gpu.py -m gpu -q -x > gpurun_out/t_decode.log 2>&1; echo "decode tests rc=$?"
tail -5 gpurun_out/t_decode.log
timeout 900 python -m pytest tests/test_fwd_gpu.py -m gpu -q > gpurun_out/t_fwd.log 2>&1; echo "fwd tests rc=$?"
tail -15 gpurun_out/t_fwd.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
tail -3 gpurun_out/smoke.log
timeout 300 python bench.py --workload c4 --steps 20 --warmup 5 --no-secondary > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench c4 rc=$?"
cat gpurun_out/bench_c4.json; tail -3 gpurun_out/bench_c4.err
timeout 600 python bench.py --workload c3 --steps 5 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench c3 rc=$?"
cat gpurun_out/bench_c3.json; tail -3 gpurun_out/bench_c3.err

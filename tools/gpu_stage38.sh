#!/bin/bash
# Stage 38: pitched DMA for head-sliced chunks of the host-buffer prefill entry (batch 1): parity, then the C5 e2e leg.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_fwd_gpu.py -m gpu -q -x 2>&1 | tail -2
timeout 300 python bench.py --workload c5 --steps 5 --no-secondary --no-cpu-baseline > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_c5.json')); print('c5 value', round(d['value'],1), 'e2e', d['e2e'])"; tail -2 gpurun_out/bench_c5.err

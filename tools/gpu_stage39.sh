#!/bin/bash
# Stage 39: batch x heads chunking of the host-buffer prefill entry: parity, then the e2e legs of C3, C2, C5.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_fwd_gpu.py -m gpu -q -x 2>&1 | tail -3
for w in c3 c2 c5; do
timeout 300 python bench.py --workload $w --steps 5 --no-secondary --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_$w.json')); print('$w value', round(d['value'],1), 'e2e', d['e2e'])"; tail -2 gpurun_out/bench_$w.err
done

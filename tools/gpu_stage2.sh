#!/bin/bash
# Second GPU pass: re-run the prefill parity tests, then the ncu launch list of the default bench command and one
# full-set capture of the prefill and decode kernels (ncu first runs the command plainly by itself).
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fwd_gpu.py -m gpu -q > gpurun_out/t_fwd.log 2>&1; echo "fwd tests rc=$?"
tail -5 gpurun_out/t_fwd.log
B="python bench.py --steps 5 --warmup 3 --cpu-budget 3"
timeout 600 $B > gpurun_out/plain_c3.json 2> gpurun_out/plain_c3.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_c3.csv $B > gpurun_out/ncu_c3.log 2>&1
echo "launch list rc=$?"; cat gpurun_out/plain_c3.json
P="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-secondary"
timeout 300 $P > gpurun_out/plain_p.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fwd_kernel -s 3 -c 1 -f -o gpurun_out/prof_fwd_c3 $P > gpurun_out/ncu_fwd.log 2>&1
echo "ncu fwd rc=$?"
P4="python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu-baseline --no-secondary"
timeout 300 $P4 > gpurun_out/plain_p4.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:decode_ -s 6 -c 2 -f -o gpurun_out/prof_decode_c4 $P4 > gpurun_out/ncu_dec.log 2>&1
echo "ncu decode rc=$?"
P2="python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline --no-secondary"
timeout 300 $P2 > gpurun_out/plain_p2.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fwd_kernel -s 3 -c 1 -f -o gpurun_out/prof_fwd_c2 $P2 > gpurun_out/ncu_fwd2.log 2>&1
echo "ncu fwd c2 rc=$?"
ls -la gpurun_out

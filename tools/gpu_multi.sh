#!/bin/bash
# Two-rank runs of the bench (one process per GPU, torchrun), as the driver launches them.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 400 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_c3_n$N.json 2> gpurun_out/bench_c3_n$N.err; echo "c3 n=$N rc=$?"; tail -1 gpurun_out/bench_c3_n$N.json; tail -3 gpurun_out/bench_c3_n$N.err
timeout 400 $TR bench.py --gpus $N --workload c5 --steps 5 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/bench_c5_n$N.json 2> gpurun_out/bench_c5_n$N.err; echo "c5 n=$N rc=$?"; tail -1 gpurun_out/bench_c5_n$N.json; tail -3 gpurun_out/bench_c5_n$N.err
timeout 400 $TR bench.py --gpus $N --workload c4 --steps 20 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/bench_c4_n$N.json 2> gpurun_out/bench_c4_n$N.err; echo "c4 n=$N rc=$?"; tail -1 gpurun_out/bench_c4_n$N.json; tail -3 gpurun_out/bench_c4_n$N.err
timeout 300 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/bench_c5_n1.json 2> gpurun_out/bench_c5_n1.err; echo "c5 n=1 rc=$?"; tail -1 gpurun_out/bench_c5_n1.json

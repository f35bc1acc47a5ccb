#!/bin/bash
# N-GPU pass of the final round-1 state: C4 and C3 bench lines (weak scaling), the sharded 256k decode, the
# multi-GPU GPU tests.
N=${1:-2}
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
run() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:3}" > gpurun_out/$2.json 2> gpurun_out/$2.err; echo "$2 rc=$?"; cut -c1-700 gpurun_out/$2.json; tail -2 gpurun_out/$2.err; }
run 29531 bench3_c4_n$N bench.py --gpus $N --workload c4 --steps 30 --warmup 3 --no-secondary --no-cpu-baseline
run 29532 bench3_c3_n$N bench.py --gpus $N --workload c3 --steps 10 --warmup 3 --no-secondary --no-cpu-baseline
run 29533 shard3_256k_n$N tools/shard_decode_bench.py 262144 20
if [ "$N" -ge 4 ]; then
run 29534 shard3_1m_n$N tools/shard_decode_bench.py 1048576 20
run 29535 bench3_c5_n$N bench.py --gpus $N --workload c5 --steps 10 --warmup 3 --no-secondary --no-cpu-baseline
fi
